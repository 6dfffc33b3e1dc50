/*
 * gritpack oracle -- TEST INFRASTRUCTURE ONLY (see gritpack_oracle.h).
 *
 * CPU restatement of /root/reference/pkg/gritagent/copy/copy.go.  The Go
 * toolchain is absent here and on the GPU box, so this is the closest thing to
 * "the reference's own code" that can run; the reference holds no test or
 * fixture for it (parity unpinned by the reference -- the contract checked in
 * tests/ is byte identity, mode bits and the sentinel payload).
 *
 *   orc_transfer_data        <- TransferData   copy.go:17-64
 *   copy_file                <- copyFile       copy.go:66-90
 *   orc_create_sentinel_file <- CreateSentinelFile copy.go:92-102
 */
#include "gritpack_oracle.h"

#include <dirent.h>
#include <errno.h>
#include <fcntl.h>
#include <sys/stat.h>
#include <unistd.h>

#include <algorithm>
#include <condition_variable>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

namespace {

struct Errors {
  std::mutex mu;
  std::vector<std::string> list;
  void add(const std::string &s) {
    std::lock_guard<std::mutex> g(mu);
    list.push_back(s);
  }
};

std::string sys_err(const char *op, const std::string &path) {
  return std::string(op) + " " + path + ": " + strerror(errno);
}

/* os.MkdirAll(path, os.ModePerm): 0777 filtered by the umask, parents included (copy.go:36) */
int mkdir_all(const std::string &path) {
  struct stat st;
  if (stat(path.c_str(), &st) == 0) return S_ISDIR(st.st_mode) ? 0 : (errno = ENOTDIR, -1);
  size_t slash = path.find_last_of('/');
  if (slash != std::string::npos && slash > 0 && mkdir_all(path.substr(0, slash)) != 0) return -1;
  if (mkdir(path.c_str(), 0777) != 0 && errno != EEXIST) return -1;
  return 0;
}

/* copy.go:66-90.  io.Copy between two *os.File is copy_file_range(2) in Go's
 * stdlib, with a 32 KiB read/write loop when the kernel refuses. */
bool copy_file(const std::string &src, const std::string &dst, std::string *err) {
  int in = open(src.c_str(), O_RDONLY | O_CLOEXEC); /* os.Open, follows symlinks */
  if (in < 0) return *err = sys_err("open", src), false;
  int out = open(dst.c_str(), O_WRONLY | O_CREAT | O_TRUNC | O_CLOEXEC, 0666); /* os.Create */
  if (out < 0) {
    *err = sys_err("open", dst);
    close(in);
    return false;
  }
  bool ok = true, use_cfr = true;
  std::vector<char> buf;
  for (;;) {
    if (use_cfr) {
      ssize_t n = copy_file_range(in, nullptr, out, nullptr, 1u << 30, 0);
      if (n > 0) continue;
      if (n == 0) break;
      if (errno == EINTR) continue;
      use_cfr = false; /* EXDEV, EINVAL, ENOSYS...: generic path */
    }
    if (buf.empty()) buf.resize(32 * 1024);
    ssize_t n = read(in, buf.data(), buf.size());
    if (n == 0) break;
    if (n < 0) {
      if (errno == EINTR) continue;
      *err = sys_err("read", src);
      ok = false;
      break;
    }
    for (ssize_t off = 0; off < n;) {
      ssize_t w = write(out, buf.data() + off, size_t(n - off));
      if (w < 0) {
        if (errno == EINTR) continue;
        *err = sys_err("write", dst);
        ok = false;
        break;
      }
      off += w;
    }
    if (!ok) break;
  }
  close(out);
  close(in);
  if (!ok) return false;
  struct stat st; /* os.Stat(srcFile); os.Chmod(dstFile, info.Mode()) */
  if (stat(src.c_str(), &st) != 0) return *err = sys_err("stat", src), false;
  if (chmod(dst.c_str(), st.st_mode & 07777) != 0) return *err = sys_err("chmod", dst), false;
  return true;
}

/* the 10-slot workerChan of copy.go:20,40: at most ten copies in flight */
struct Gate {
  std::mutex mu;
  std::condition_variable cv;
  int free_slots = 10;
  void acquire() {
    std::unique_lock<std::mutex> l(mu);
    cv.wait(l, [&] { return free_slots > 0; });
    --free_slots;
  }
  void release() {
    {
      std::lock_guard<std::mutex> g(mu);
      ++free_slots;
    }
    cv.notify_one();
  }
};

struct Walk {
  std::string src_root, dst_root;
  Gate gate;
  Errors errs;
  std::vector<std::thread> workers;
  std::string fatal;

  /* filepath.WalkDir: the root first, then entries of each directory in
   * lexical order; a symlink is not a directory to the walker (copy.go:35) */
  bool visit(const std::string &rel, bool is_dir) {
    std::string src = rel.empty() ? src_root : src_root + "/" + rel;
    std::string dst = rel.empty() ? dst_root : dst_root + "/" + rel;
    if (!is_dir) {
      gate.acquire();
      workers.emplace_back([this, src, dst] {
        std::string e;
        if (!copy_file(src, dst, &e)) errs.add(e);
        gate.release();
      });
      return true;
    }
    if (mkdir_all(dst) != 0) return fatal = sys_err("mkdir", dst), false;
    DIR *d = opendir(src.c_str());
    if (!d) return fatal = sys_err("open", src), false;
    std::vector<std::pair<std::string, bool>> names;
    while (struct dirent *e = readdir(d)) {
      if (!strcmp(e->d_name, ".") || !strcmp(e->d_name, "..")) continue;
      bool dir = e->d_type == DT_DIR;
      if (e->d_type == DT_UNKNOWN) {
        struct stat st;
        dir = lstat((src + "/" + e->d_name).c_str(), &st) == 0 && S_ISDIR(st.st_mode);
      }
      names.emplace_back(e->d_name, dir);
    }
    closedir(d);
    std::sort(names.begin(), names.end());
    for (auto &n : names)
      if (!visit(rel.empty() ? n.first : rel + "/" + n.first, n.second)) return false;
    return true;
  }
};

}  // namespace

extern "C" int orc_transfer_data(const char *src_dir, const char *dst_dir, char *err,
                                 size_t errlen) {
  Walk w;
  w.src_root = src_dir;
  w.dst_root = dst_dir;
  struct stat st;
  bool ok;
  if (lstat(src_dir, &st) != 0) {
    w.fatal = sys_err("lstat", src_dir);
    ok = false;
  } else {
    ok = w.visit("", S_ISDIR(st.st_mode));
  }
  /* Go returns a walk error without waiting (copy.go:56-58); joining here only
   * keeps the test process tidy and changes no observable file. */
  for (auto &t : w.workers) t.join();
  std::string msg;
  if (!ok) msg = w.fatal;
  for (auto &e : w.errs.list) msg += (msg.empty() ? "" : "; ") + e; /* multierr.Combine */
  if (err && errlen) snprintf(err, errlen, "%s", msg.c_str());
  return msg.empty() ? 0 : -1;
}

extern "C" int orc_create_sentinel_file(const char *dir, const char *file_name) {
  std::string path = std::string(dir) + "/" + file_name;
  int fd = open(path.c_str(), O_WRONLY | O_CREAT | O_TRUNC | O_CLOEXEC, 0666);
  if (fd < 0) return -1;
  static const char payload[] = "Transfer Completed\n"; /* copy.go:100 */
  ssize_t n = write(fd, payload, sizeof payload - 1);
  close(fd);
  return n == (ssize_t)(sizeof payload - 1) ? 0 : -1;
}
