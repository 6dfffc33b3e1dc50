// Directory level of the C ABI: the drop-in for TransferData.
//
//   grit_pack_dir    <- copy.TransferData in the upload direction
//                       (pkg/gritagent/checkpoint/checkpoint.go:20 -> copy.go:17-64)
//   grit_restore_dir <- copy.TransferData in the download direction
//                       (pkg/gritagent/restore/restore.go:16 -> copy.go:17-64)
//
// The reference mirrors the tree file by file (copy.go:66-90).  Here the tree
// becomes one page stream -- every regular file, in filepath.WalkDir order
// (copy.go:23), padded with zeros to a 4 KiB multiple -- that goes through the
// GPU pack pipeline; the PV receives
//     pages-NNN.bin      the pages that had to be stored: slot order, striped 1-8 MiB-wise over 8 files
//                        (a file system takes one writer per file; eight files take eight)
//     pagemap.bin        u32 ref[n_pages] then u32 crc32c[n_pages]
//     hashes.bin         u64 hash64[n_pages]   (input of the next incremental pack)
//     gritpack.manifest  tree layout + sizes + checksums, written last (rename)
// and restore rebuilds paths, bytes and mode bits exactly (copy.go:36,84-89).
#include <dirent.h>
#include <errno.h>
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <sys/vfs.h>
#include <unistd.h>

#include <algorithm>
#include <atomic>
#include <chrono>
#include <condition_variable>
#include <cstring>
#include <deque>
#include <functional>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "hostpipe.h"

using namespace grit;

namespace {

constexpr char kMagic[8] = {'G', 'R', 'I', 'T', 'P', 'A', 'K', '1'};
constexpr uint32_t kFormatVersion = 3;
const char* const kManifest = "gritpack.manifest";

struct Entry {
  uint8_t type;  // 0 directory, 1 regular file
  uint32_t mode;
  uint64_t size, first_page;
  std::string path;  // relative to the root
  uint64_t pages() const { return (size + kPage - 1) / kPage; }
};

struct Manifest {
  std::vector<Entry> entries;
  uint64_t n_pages = 0, n_slots = 0, bytes_logical = 0;
  uint32_t pagemap_crc = 0, hashes_crc = 0;
  uint32_t n_shards = 8, stripe_slots = 2048;  // layout of the stored pages over pages-NNN.bin
  std::string parent;
  std::vector<std::string> peers;  // packed directories of ranks 0..n-1 that PEER references point into
};

std::string errno_msg(const char* op, const std::string& path) {
  return std::string(op) + " " + path + ": " + strerror(errno);
}

double now_s() {
  return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

// GRITPACK_DEBUG=1: wall-clock phases of the directory calls on stderr
struct Phases {
  const char* what;
  double last = now_s();
  std::string line;
  explicit Phases(const char* w) : what(w) {}
  void mark(const char* name) {
    const double t = now_s();
    char b[64];
    snprintf(b, sizeof b, " %s %.1f ms,", name, (t - last) * 1e3);
    line += b;
    last = t;
  }
  ~Phases() {
    if (getenv("GRITPACK_DEBUG")) fprintf(stderr, "[gritpack] %s:%s\n", what, line.c_str());
  }
};

// os.MkdirAll(path, os.ModePerm), copy.go:36
int mkdir_all(const std::string& path) {
  struct stat st;
  if (stat(path.c_str(), &st) == 0) return S_ISDIR(st.st_mode) ? 0 : (errno = ENOTDIR, -1);
  const size_t slash = path.find_last_of('/');
  if (slash != std::string::npos && slash > 0 && mkdir_all(path.substr(0, slash)) != 0) return -1;
  if (mkdir(path.c_str(), 0777) != 0 && errno != EEXIST) return -1;
  return 0;
}

// filepath.WalkDir order (copy.go:23): a directory, then its entries sorted by name.
// A symlink is followed like os.Open does (copy.go:67); anything that is then
// neither a regular file nor a directory is refused.
int walk(const std::string& root, const std::string& rel, std::vector<Entry>* out, std::string* msg) {
  const std::string dir = rel.empty() ? root : root + "/" + rel;
  DIR* d = opendir(dir.c_str());
  if (!d) return *msg = errno_msg("open", dir), GRIT_EIO;
  std::vector<std::string> names;
  while (struct dirent* e = readdir(d))
    if (strcmp(e->d_name, ".") && strcmp(e->d_name, "..")) names.emplace_back(e->d_name);
  closedir(d);
  std::sort(names.begin(), names.end());
  for (const auto& n : names) {
    const std::string r = rel.empty() ? n : rel + "/" + n, full = root + "/" + r;
    struct stat lst, st;
    if (lstat(full.c_str(), &lst) != 0) return *msg = errno_msg("lstat", full), GRIT_EIO;
    if (S_ISDIR(lst.st_mode)) {
      out->push_back({0, (uint32_t)(lst.st_mode & 07777), 0, 0, r});
      if (int rc = walk(root, r, out, msg)) return rc;
      continue;
    }
    if (stat(full.c_str(), &st) != 0) return *msg = errno_msg("stat", full), GRIT_EIO;
    if (!S_ISREG(st.st_mode)) return *msg = full + ": not a regular file or directory", GRIT_EINVAL;
    out->push_back({1, (uint32_t)(st.st_mode & 07777), (uint64_t)st.st_size, 0, r});
  }
  return GRIT_OK;
}

// ---- little-endian (de)serialisation of the manifest
void put(std::string* b, const void* p, size_t n) { b->append(static_cast<const char*>(p), n); }
template <class T>
void put(std::string* b, T v) {
  put(b, &v, sizeof v);
}
struct Reader {
  const std::string& b;
  size_t off = 0;
  bool ok = true;
  template <class T>
  T get() {
    T v{};
    if (off + sizeof v > b.size()) return ok = false, v;
    memcpy(&v, b.data() + off, sizeof v);
    off += sizeof v;
    return v;
  }
  std::string str(size_t n) {
    if (off + n > b.size()) return ok = false, std::string();
    std::string s = b.substr(off, n);
    off += n;
    return s;
  }
};

std::string encode_manifest(const Manifest& m) {
  std::string b;
  put(&b, kMagic, 8);
  put<uint32_t>(&b, kFormatVersion);
  put<uint32_t>(&b, (uint32_t)m.entries.size());
  put<uint64_t>(&b, m.n_pages);
  put<uint64_t>(&b, m.n_slots);
  put<uint64_t>(&b, m.bytes_logical);
  put<uint32_t>(&b, m.pagemap_crc);
  put<uint32_t>(&b, m.hashes_crc);
  put<uint32_t>(&b, m.n_shards);
  put<uint32_t>(&b, m.stripe_slots);
  put<uint32_t>(&b, (uint32_t)m.parent.size());
  put(&b, m.parent.data(), m.parent.size());
  put<uint32_t>(&b, (uint32_t)m.peers.size());
  for (const auto& p : m.peers) {
    put<uint32_t>(&b, (uint32_t)p.size());
    put(&b, p.data(), p.size());
  }
  for (const auto& e : m.entries) {
    put<uint8_t>(&b, e.type);
    put<uint32_t>(&b, e.mode);
    put<uint64_t>(&b, e.size);
    put<uint64_t>(&b, e.first_page);
    put<uint32_t>(&b, (uint32_t)e.path.size());
    put(&b, e.path.data(), e.path.size());
  }
  put<uint32_t>(&b, grit_crc32c(0, b.data(), b.size()));
  return b;
}

int read_file(const std::string& path, std::string* out, std::string* msg) {
  const int fd = open(path.c_str(), O_RDONLY | O_CLOEXEC);
  if (fd < 0) return *msg = errno_msg("open", path), GRIT_EIO;
  struct stat st;
  if (fstat(fd, &st) != 0) {
    *msg = errno_msg("stat", path);
    close(fd);
    return GRIT_EIO;
  }
  out->resize((size_t)st.st_size);
  size_t off = 0;
  while (off < out->size()) {
    const ssize_t n = pread(fd, &(*out)[off], out->size() - off, (off_t)off);
    if (n <= 0) {
      if (n < 0 && errno == EINTR) continue;
      *msg = errno_msg("read", path);
      close(fd);
      return GRIT_EIO;
    }
    off += (size_t)n;
  }
  close(fd);
  return GRIT_OK;
}

int write_file(const std::string& path, const void* data, size_t len, std::string* msg) {
  const int fd = open(path.c_str(), O_WRONLY | O_CREAT | O_TRUNC | O_CLOEXEC, 0666);
  if (fd < 0) return *msg = errno_msg("open", path), GRIT_EIO;
  const char* p = static_cast<const char*>(data);
  size_t off = 0;
  while (off < len) {
    const ssize_t n = write(fd, p + off, len - off);
    if (n < 0) {
      if (errno == EINTR) continue;
      *msg = errno_msg("write", path);
      close(fd);
      return GRIT_EIO;
    }
    off += (size_t)n;
  }
  if (fsync(fd) != 0 || close(fd) != 0) return *msg = errno_msg("fsync", path), GRIT_EIO;
  return GRIT_OK;
}

// a manifest path must stay inside the destination: relative, no empty / "." / ".." component
bool safe_relative(const std::string& p) {
  if (p.empty() || p[0] == '/') return false;
  for (size_t a = 0; a <= p.size();) {
    const size_t b = std::min(p.find('/', a), p.size());
    const std::string c = p.substr(a, b - a);
    if (c.empty() || c == "." || c == "..") return false;
    a = b + 1;
  }
  return true;
}

bool has_manifest(const std::string& dir) {
  struct stat st;
  return stat((dir + "/" + kManifest).c_str(), &st) == 0 && S_ISREG(st.st_mode);
}

int load_manifest(const std::string& dir, Manifest* m, std::string* msg) {
  std::string b;
  if (int rc = read_file(dir + "/" + kManifest, &b, msg)) return rc;
  if (b.size() < 8 + 4 * 2 + 8 * 3 + 4 * 6 + 4 || memcmp(b.data(), kMagic, 8))
    return *msg = dir + ": not a gritpack manifest", GRIT_EFORMAT;
  uint32_t want;
  memcpy(&want, b.data() + b.size() - 4, 4);
  if (grit_crc32c(0, b.data(), b.size() - 4) != want) return *msg = dir + ": manifest checksum mismatch", GRIT_EFORMAT;
  Reader r{b, 8};
  if (r.get<uint32_t>() != kFormatVersion) return *msg = dir + ": unknown manifest version", GRIT_EFORMAT;
  const uint32_t n_entries = r.get<uint32_t>();
  m->n_pages = r.get<uint64_t>();
  m->n_slots = r.get<uint64_t>();
  m->bytes_logical = r.get<uint64_t>();
  m->pagemap_crc = r.get<uint32_t>();
  m->hashes_crc = r.get<uint32_t>();
  m->n_shards = r.get<uint32_t>();
  m->stripe_slots = r.get<uint32_t>();
  if (m->n_shards == 0 || m->n_shards > 256 || m->stripe_slots == 0) return *msg = dir + ": bad page store layout", GRIT_EFORMAT;
  m->parent = r.str(r.get<uint32_t>());
  const uint32_t n_peer_dirs = r.get<uint32_t>();
  if (n_peer_dirs > GRIT_MAX_PEERS) return *msg = dir + ": bad peer list", GRIT_EFORMAT;
  for (uint32_t i = 0; i < n_peer_dirs && r.ok; i++) m->peers.push_back(r.str(r.get<uint32_t>()));
  for (uint32_t i = 0; i < n_entries && r.ok; i++) {
    Entry e;
    e.type = r.get<uint8_t>();
    e.mode = r.get<uint32_t>();
    e.size = r.get<uint64_t>();
    e.first_page = r.get<uint64_t>();
    e.path = r.str(r.get<uint32_t>());
    if (!safe_relative(e.path)) return *msg = dir + ": manifest holds an unsafe path", GRIT_EFORMAT;
    m->entries.push_back(std::move(e));
  }
  if (!r.ok || r.off + 4 != b.size() || m->n_pages > GRIT_MAX_PAGES) return *msg = dir + ": truncated manifest", GRIT_EFORMAT;
  // the files tile the page stream in manifest order, without gaps or overlap (how grit_pack_dir numbers them)
  uint64_t next_page = 0, bytes = 0;
  for (const auto& e : m->entries) {
    if (e.type > 1) return *msg = dir + ": manifest holds an unknown entry type", GRIT_EFORMAT;
    if (e.type == 0) continue;
    if (e.first_page != next_page || e.size > (uint64_t)GRIT_MAX_PAGES * kPage) return *msg = dir + ": manifest entries do not tile the image", GRIT_EFORMAT;
    next_page += e.pages();
    bytes += e.size;
  }
  if (next_page != m->n_pages || bytes != m->bytes_logical || m->n_slots > m->n_pages)
    return *msg = dir + ": manifest totals do not match its entries", GRIT_EFORMAT;
  return GRIT_OK;
}

// per-page arrays of a packed directory
int load_pagemap(const std::string& dir, const Manifest& m, std::vector<uint32_t>* ref, std::vector<uint32_t>* crc,
                 std::string* msg) {
  std::string b;
  if (int rc = read_file(dir + "/pagemap.bin", &b, msg)) return rc;
  if (b.size() != m.n_pages * 8 || grit_crc32c(0, b.data(), b.size()) != m.pagemap_crc)
    return *msg = dir + "/pagemap.bin: size or checksum mismatch", GRIT_EFORMAT;
  ref->resize(m.n_pages);
  crc->resize(m.n_pages);
  memcpy(ref->data(), b.data(), m.n_pages * 4);
  memcpy(crc->data(), b.data() + m.n_pages * 4, m.n_pages * 4);
  return GRIT_OK;
}

int load_hashes(const std::string& dir, const Manifest& m, std::vector<uint64_t>* hash, std::string* msg) {
  std::string b;
  if (int rc = read_file(dir + "/hashes.bin", &b, msg)) return rc;
  if (b.size() != m.n_pages * 8 || grit_crc32c(0, b.data(), b.size()) != m.hashes_crc)
    return *msg = dir + "/hashes.bin: size or checksum mismatch", GRIT_EFORMAT;
  hash->resize(m.n_pages);
  memcpy(hash->data(), b.data(), b.size());
  return GRIT_OK;
}

// One contiguous piece of file I/O between a file and a staging buffer.
struct Piece {
  const Entry* e;
  uint64_t file_off, len;
  char* buf;
};

// Piece size that spreads one chunk of `pages` pages over `threads` threads: 1 to 8 MiB.
uint64_t piece_bytes_for(uint64_t pages, unsigned threads) {
  const uint64_t mib = 1ull << 20, per = (pages * kPage / std::max(1u, threads) + mib - 1) / mib * mib;
  return std::min<uint64_t>(8 * mib, std::max<uint64_t>(mib, per));
}

// the pieces of page range [p0, p1) of the stream, each at most max_piece bytes
void pieces_for(const std::vector<const Entry*>& files, uint64_t p0, uint64_t p1, char* buf, uint64_t max_piece,
                std::vector<Piece>* out) {
  auto it = std::upper_bound(files.begin(), files.end(), p0,
                             [](uint64_t p, const Entry* e) { return p < e->first_page + e->pages(); });
  for (; it != files.end() && (*it)->first_page < p1; ++it) {
    const Entry* e = *it;
    const uint64_t a = std::max(p0, e->first_page), b = std::min(p1, e->first_page + e->pages());
    uint64_t off = (a - e->first_page) * kPage;
    const uint64_t end = std::min<uint64_t>(e->size, (b - e->first_page) * kPage);
    char* dst = buf + (a - p0) * kPage;
    while (off < end) {
      const uint64_t n = std::min<uint64_t>(end - off, max_piece);
      out->push_back({e, off, n, dst});
      off += n;
      dst += n;
    }
  }
}

// zero the bytes between a file's end and the end of its last page, for the files ending in [p0, p1)
void zero_tails(const std::vector<const Entry*>& files, uint64_t p0, uint64_t p1, char* buf) {
  auto it = std::upper_bound(files.begin(), files.end(), p0,
                             [](uint64_t p, const Entry* e) { return p < e->first_page + e->pages(); });
  for (; it != files.end() && (*it)->first_page < p1; ++it) {
    const Entry* e = *it;
    const uint64_t last = e->first_page + e->pages() - 1, tail = e->size % kPage;
    if (tail && last >= p0 && last < p1) memset(buf + (last - p0) * kPage + tail, 0, kPage - tail);
  }
}

// The I/O threads of one directory call.  for_each(n, fn) runs fn(0) .. fn(n-1) on them (the caller works too) and
// returns when all are done; it may be called from several threads at once (pack: reader and writer side).
class IoPool {
 public:
  explicit IoPool(unsigned threads) {
    for (unsigned i = 1; i < std::max(1u, threads); i++) th_.emplace_back([this] { work(); });
  }
  ~IoPool() {
    {
      std::lock_guard<std::mutex> l(mu_);
      stop_ = true;
    }
    cv_.notify_all();
    for (auto& t : th_) t.join();
  }
  void for_each(size_t n, const std::function<void(size_t)>& fn) {
    if (!n) return;
    Batch b{&fn, 0, n, 0};
    std::unique_lock<std::mutex> l(mu_);
    q_.push_back(&b);
    cv_.notify_all();
    for (;;) {
      if (b.next < b.n) {
        const size_t i = take(&b);
        l.unlock();
        fn(i);
        l.lock();
        b.done++;
      } else if (b.done < b.n) {
        done_cv_.wait(l);
      } else {
        return;
      }
    }
  }

 private:
  struct Batch {
    const std::function<void(size_t)>* fn;
    size_t next, n, done;
  };
  size_t take(Batch* b) {  // mu_ held
    const size_t i = b->next++;
    if (b->next == b->n) q_.erase(std::find(q_.begin(), q_.end(), b));
    return i;
  }
  void work() {
    std::unique_lock<std::mutex> l(mu_);
    for (;;) {
      cv_.wait(l, [&] { return stop_ || !q_.empty(); });
      if (q_.empty()) return;
      Batch* b = q_.front();
      const size_t i = take(b);
      l.unlock();
      (*b->fn)(i);
      l.lock();
      if (++b->done == b->n) done_cv_.notify_all();  // b may be gone once the lock is released
    }
  }
  std::mutex mu_;
  std::condition_variable cv_, done_cv_;
  std::deque<Batch*> q_;
  std::vector<std::thread> th_;
  bool stop_ = false;
};

// run fn over the pieces on the pool; first error wins
template <class Fn>
int for_pieces(IoPool* pool, const std::vector<Piece>& ps, std::string* msg, Fn fn) {
  std::mutex emu;
  std::string first;
  std::atomic<bool> failed{false};
  pool->for_each(ps.size(), [&](size_t i) {
    if (failed.load(std::memory_order_relaxed)) return;
    std::string e;
    fn(ps[i], &e);
    if (!e.empty()) {
      std::lock_guard<std::mutex> l(emu);
      if (first.empty()) first = e;
      failed = true;
    }
  });
  if (failed) return *msg = first, GRIT_EIO;
  return GRIT_OK;
}

void read_piece(const std::string& root, const Piece& p, std::string* err) {
  const std::string path = root + "/" + p.e->path;
  const int fd = open(path.c_str(), O_RDONLY | O_CLOEXEC);
  if (fd < 0) return void(*err = errno_msg("open", path));
  uint64_t done = 0;
  while (done < p.len) {
    const ssize_t n = pread(fd, p.buf + done, p.len - done, (off_t)(p.file_off + done));
    if (n < 0 && errno == EINTR) continue;
    if (n <= 0) {  // the file shrank under us: the image would not be the one the manifest describes
      *err = n < 0 ? errno_msg("read", path) : path + ": file changed while packing";
      break;
    }
    done += (uint64_t)n;
  }
  close(fd);
}

void write_piece(const std::string& root, const Piece& p, std::string* err) {
  const std::string path = root + "/" + p.e->path;
  const int fd = open(path.c_str(), O_WRONLY | O_CLOEXEC);
  if (fd < 0) return void(*err = errno_msg("open", path));
  uint64_t done = 0;
  while (done < p.len) {
    const ssize_t n = pwrite(fd, p.buf + done, p.len - done, (off_t)(p.file_off + done));
    if (n < 0 && errno == EINTR) continue;
    if (n <= 0) {
      *err = errno_msg("write", path);
      break;
    }
    done += (uint64_t)n;
  }
  close(fd);
}

// The stored pages on the PV.  Slot s lives in stripe t = s / stripe, which is stripe number t / k of
// file pages-(t % k).bin: consecutive stripes go to different files, so a chunk of output is written
// (and read) by several threads without two of them sharing a file.
struct PageStore {
  std::string dir;
  uint32_t k = 8, stripe = 2048;
  std::vector<int> fds;
  ~PageStore() {
    for (int fd : fds)
      if (fd >= 0) close(fd);
  }
  std::string path(uint32_t i) const {
    char n[32];
    snprintf(n, sizeof n, "/pages-%03u.bin", i);
    return dir + n;
  }
  int open_all(bool writing, std::string* msg) {
    for (uint32_t i = 0; i < k; i++) {
      const int fd = writing ? open(path(i).c_str(), O_WRONLY | O_CREAT | O_TRUNC | O_CLOEXEC, 0666)
                             : open(path(i).c_str(), O_RDONLY | O_CLOEXEC);
      if (fd < 0) return *msg = errno_msg("open", path(i)), GRIT_EIO;
      fds.push_back(fd);
    }
    return GRIT_OK;
  }
  // bytes file i holds when the store has n_slots pages
  uint64_t shard_bytes(uint32_t i, uint64_t n_slots) const {
    const uint64_t full = n_slots / stripe, rem = n_slots % stripe;
    uint64_t slots = full / k * stripe + (i < full % k ? stripe : 0);
    if (i == full % k) slots += rem;
    return slots * kPage;
  }
  int check_sizes(uint64_t n_slots, std::string* msg) const {
    for (uint32_t i = 0; i < k; i++) {
      struct stat st;
      if (fstat(fds[i], &st) != 0 || (uint64_t)st.st_size != shard_bytes(i, n_slots))
        return *msg = path(i) + ": size does not match the manifest", GRIT_EFORMAT;
    }
    return GRIT_OK;
  }
  // read or write slots [s0, s0 + n) from / to buf, one piece per stripe, on the pool
  int io(uint64_t s0, uint64_t n, char* buf, IoPool* pool, bool write) const {
    struct P {
      int fd;
      uint64_t off, len;
      char* buf;
    };
    std::vector<P> ps;
    for (uint64_t s = s0; s < s0 + n;) {
      const uint64_t t = s / stripe, in = s % stripe, m = std::min<uint64_t>(stripe - in, s0 + n - s);
      ps.push_back({fds[t % k], (t / k * stripe + in) * kPage, m * kPage, buf + (s - s0) * kPage});
      s += m;
    }
    std::atomic<int> rc{0};
    pool->for_each(ps.size(), [&](size_t i) {
      uint64_t done = 0;
      while (done < ps[i].len && !rc.load(std::memory_order_relaxed)) {
        const ssize_t r = write ? pwrite(ps[i].fd, ps[i].buf + done, ps[i].len - done, (off_t)(ps[i].off + done))
                                : pread(ps[i].fd, ps[i].buf + done, ps[i].len - done, (off_t)(ps[i].off + done));
        if (r < 0 && errno == EINTR) continue;
        if (r <= 0) {
          rc = GRIT_EIO;
          return;
        }
        done += (uint64_t)r;
      }
    });
    return rc;
  }
  int sync_all() const {
    for (int fd : fds)
      if (fsync(fd) != 0) return GRIT_EIO;
    return 0;
  }
};

struct Opts {
  int device = 0;
  bool verify = true;
  unsigned io_threads = std::min(16u, std::max(8u, std::thread::hardware_concurrency()));
  int lanes = 4;             // download of a big tree: interleaved runs of the page range (GRITPACK_LANES)
  uint64_t chunk_pages = 0;  // 0: chosen from the image size
  std::string parent;
  uint32_t n_peers = 0;
  const grit_table* peer_tables = nullptr;
  std::vector<std::string> peer_dirs;
  grit_comm* comm = nullptr;  // ranked pack: the tables come over NCCL instead of through peer_tables
  // Chunks of 1/128 of the image, 4 to 16 MiB.  The pinned staging is six chunks, and pinning memory costs about
  // 0.5 ms per MiB (cudaHostAlloc, and again to free): with 64 MiB chunks that was 0.2-0.4 s per call for nothing,
  // since the I/O threads split a chunk among themselves whatever its size and a PCIe copy of 4 MiB is at full rate.
  uint64_t chunk_for(uint64_t n_pages) const {
    if (chunk_pages) return chunk_pages;
    const uint64_t lo = (4ull << 20) / kPage, hi = (16ull << 20) / kPage;
    return std::min(hi, std::max(lo, (n_pages / 128 + 255) / 256 * 256));
  }
};

Opts read_opts(const grit_opts* o) {
  Opts r;
  if (!o) return r;
  r.device = o->device;
  r.verify = o->verify != 0;
  if (o->io_threads) r.io_threads = o->io_threads;
  else if (const char* e = getenv("GRITPACK_IO_THREADS")) r.io_threads = std::max(1, atoi(e));
  if (const char* e = getenv("GRITPACK_LANES")) r.lanes = std::max(1, atoi(e));
  if (o->chunk_bytes >= (1u << 20)) r.chunk_pages = o->chunk_bytes / kPage;
  if (o->parent_dir) r.parent = o->parent_dir;
  const bool has_comm = o->struct_size >= offsetof(grit_opts, comm) + sizeof(o->comm) && o->comm;
  if (o->struct_size >= offsetof(grit_opts, peer_dirs) + sizeof(o->peer_dirs) && o->n_peers && o->peer_dirs &&
      (o->peer_tables || has_comm)) {
    r.n_peers = has_comm ? 0 : o->n_peers;
    r.peer_tables = has_comm ? nullptr : o->peer_tables;
    for (uint32_t i = 0; i < o->n_peers; i++) r.peer_dirs.emplace_back(o->peer_dirs[i] ? o->peer_dirs[i] : "");
  }
  if (has_comm) r.comm = o->comm;
  return r;
}

// criu-extents.txt: which stream pages belong to which CRIU pagemap entry (defined with the CRIU code at the end)
int write_criu_extents(const std::string& src, const std::string& dst, const Manifest& m, std::string* msg);

struct CtxGuard {
  grit_ctx* ctx = nullptr;
  void release() {
    grit_ctx_destroy(ctx);
    ctx = nullptr;
  }
  ~CtxGuard() { release(); }
};

// Fully restored images of other packed directories (parents of incremental checkpoints, lower ranks of a
// multi-GPU job), kept on the device while a restore refers to them.  An image may itself refer to a parent
// and to lower ranks; each directory is loaded once.
struct ImageCache {
  struct Img {
    void* d = nullptr;
    uint64_t pages = 0;
  };
  // a deque: entries keep their address while later ones are added (load_dependencies holds a pointer to the
  // parent's entry across the loads of the peers)
  std::deque<std::pair<std::string, Img>> items;
  void release() {
    for (auto& it : items) cudaFree(it.second.d);
    items.clear();
  }
  ~ImageCache() { release(); }
  const Img* find(const std::string& dir) const {
    for (auto& it : items)
      if (it.first == dir) return &it.second;
    return nullptr;
  }
};

std::string resolve_parent(const std::string& child_dir, const std::string& parent) {
  if (has_manifest(parent)) return parent;
  // the PV may be mounted elsewhere at restore time: try the sibling of the same name
  const size_t s1 = parent.find_last_not_of('/');
  const size_t s0 = parent.find_last_of('/', s1);
  const std::string base = parent.substr(s0 == std::string::npos ? 0 : s0 + 1, s1 - (s0 == std::string::npos ? 0 : s0 + 1) + 1);
  const size_t c = child_dir.find_last_of('/', child_dir.find_last_not_of('/'));
  const std::string sib = (c == std::string::npos ? std::string(".") : child_dir.substr(0, c)) + "/" + base;
  return has_manifest(sib) ? sib : parent;
}

// which lower ranks the page map refers to; an error if it names a rank the manifest does not list
int peers_used(const Manifest& m, const std::vector<uint32_t>& ref, bool (&used)[GRIT_MAX_PEERS], std::string* msg) {
  for (uint64_t p = 0; p < m.n_pages; p++)
    if (GRIT_REF_KIND(ref[p]) == GRIT_KIND_PEER) {
      const uint32_t rk = GRIT_REF_VALUE(ref[p]) >> 26;
      if (rk >= m.peers.size())
        return *msg = "page " + std::to_string(p) + " refers to rank " + std::to_string(rk) + ", which the manifest does not name",
               GRIT_EFORMAT;
      used[rk] = true;
    }
  return GRIT_OK;
}

int image_to_device(grit_ctx* ctx, const std::string& dir, int verify, ImageCache* cache, const ImageCache::Img** out,
                    int depth, grit_stats* stats, std::string* msg);

// parent and peer images a page map needs, from the cache
int load_dependencies(grit_ctx* ctx, const std::string& dir, const Manifest& m, const std::vector<uint32_t>& ref, int verify,
                      ImageCache* cache, int depth, grit_stats* stats, const ImageCache::Img** parent, PeerImages* peers,
                      std::string* msg) {
  *parent = nullptr;
  if (!m.parent.empty())
    if (int rc = image_to_device(ctx, resolve_parent(dir, m.parent), verify, cache, parent, depth + 1, stats, msg)) return rc;
  bool used[GRIT_MAX_PEERS] = {};
  if (int rc = peers_used(m, ref, used, msg)) return *msg = dir + ": " + *msg, rc;
  *peers = PeerImages{};
  peers->n_peers = (uint32_t)m.peers.size();
  for (uint32_t r = 0; r < m.peers.size(); r++) {
    if (!used[r]) continue;
    const ImageCache::Img* img = nullptr;
    if (int rc = image_to_device(ctx, resolve_parent(dir, m.peers[r]), verify, cache, &img, depth + 1, stats, msg)) return rc;
    peers->peer[r] = static_cast<const unsigned char*>(img->d);
    peers->peer_pages[r] = img->pages;
  }
  return GRIT_OK;
}

int image_to_device(grit_ctx* ctx, const std::string& dir, int verify, ImageCache* cache, const ImageCache::Img** out,
                    int depth, grit_stats* stats, std::string* msg) {
  if (const ImageCache::Img* hit = cache->find(dir)) return *out = hit, GRIT_OK;
  if (depth > 64) return *msg = "parent / peer chain too deep", GRIT_EFORMAT;
  Manifest m;
  if (int rc = load_manifest(dir, &m, msg)) return rc;
  std::vector<uint32_t> ref, crc;
  if (int rc = load_pagemap(dir, m, &ref, &crc, msg)) return rc;
  const ImageCache::Img* parent = nullptr;
  PeerImages peers{};
  if (int rc = load_dependencies(ctx, dir, m, ref, verify, cache, depth, stats, &parent, &peers, msg)) return rc;
  char err[512] = "";
  const size_t errlen = sizeof err;
  void *d_packed = nullptr, *d_out = nullptr;
  uint32_t *d_ref = nullptr, *d_crc = nullptr;
  auto cleanup = [&] {
    cudaFree(d_packed);
    cudaFree(d_ref);
    cudaFree(d_crc);
  };
  auto body = [&]() -> int {
    GRIT_CUDA(cudaMalloc(&d_packed, std::max<uint64_t>(m.n_slots, 1) * kPage));
    GRIT_CUDA(cudaMalloc(&d_out, std::max<uint64_t>(m.n_pages, 1) * kPage));
    GRIT_CUDA(cudaMalloc(&d_ref, std::max<uint64_t>(m.n_pages, 1) * 4));
    GRIT_CUDA(cudaMalloc(&d_crc, std::max<uint64_t>(m.n_pages, 1) * 4));
    // stored pages -> device through a pinned bounce buffer
    PageStore store;
    store.dir = dir;
    store.k = m.n_shards;
    store.stripe = m.stripe_slots;
    std::string smsg;
    if (int rc = store.open_all(false, &smsg)) return fail(err, errlen, rc, "%s", smsg.c_str());
    if (int rc = store.check_sizes(m.n_slots, &smsg)) return fail(err, errlen, rc, "%s", smsg.c_str());
    void* h = nullptr;
    const uint64_t step = (32ull << 20) / kPage;
    if (cudaMallocHost(&h, step * kPage) != cudaSuccess) return fail(err, errlen, GRIT_ECUDA, "cudaMallocHost failed");
    int rc = GRIT_OK;
    IoPool pool(8);
    for (uint64_t s0 = 0; s0 < m.n_slots && !rc; s0 += step) {
      const uint64_t n = std::min(step, m.n_slots - s0);
      if (store.io(s0, n, static_cast<char*>(h), &pool, false)) rc = fail(err, errlen, GRIT_EIO, "%s: short read", dir.c_str());
      if (!rc && cudaMemcpy(static_cast<char*>(d_packed) + s0 * kPage, h, n * kPage, cudaMemcpyHostToDevice) != cudaSuccess)
        rc = fail(err, errlen, GRIT_ECUDA, "H2D of the stored pages of %s failed", dir.c_str());
      if (stats) stats->h2d_bytes += (double)n * kPage;
    }
    cudaFreeHost(h);
    if (rc) return rc;
    GRIT_CUDA(cudaMemcpy(d_ref, ref.data(), m.n_pages * 4, cudaMemcpyHostToDevice));
    GRIT_CUDA(cudaMemcpy(d_crc, crc.data(), m.n_pages * 4, cudaMemcpyHostToDevice));
    const void* peer_ptrs[GRIT_MAX_PEERS];
    for (uint32_t r = 0; r < GRIT_MAX_PEERS; r++) peer_ptrs[r] = peers.peer[r];
    grit_restore_stats rs{};
    rc = grit_restore_buffer(ctx, d_packed, m.n_slots, d_ref, d_crc, m.n_pages, parent ? parent->d : nullptr,
                             parent ? parent->pages : 0, peer_ptrs, peers.peer_pages, peers.n_peers, d_out, verify, &rs,
                             nullptr, err, errlen);
    if (stats) {
      stats->gpu_ms += rs.gpu_ms;
      stats->n_bad += rs.n_bad;
    }
    return rc;
  };
  const int rc = body();
  cleanup();
  if (rc) {
    cudaFree(d_out);
    *msg = dir + ": " + err;
    return rc;
  }
  cache->items.push_back({dir, ImageCache::Img{d_out, m.n_pages}});
  *out = &cache->items.back().second;
  return GRIT_OK;
}

}  // namespace

extern "C" {

int grit_pack_dir(const char* src_dir, const char* dst_dir, const grit_opts* opts, grit_stats* stats, char* err,
                  size_t errlen) {
  if (!src_dir || !dst_dir) return fail(err, errlen, GRIT_EINVAL, "null directory");
  const double t0 = now_s();
  Phases ph("pack_dir");
  const Opts o = read_opts(opts);
  const std::string src = src_dir, dst = dst_dir;
  std::string msg;
  grit_stats st{};

  Manifest m;
  struct stat sst;
  if (stat(src.c_str(), &sst) != 0 || !S_ISDIR(sst.st_mode)) return fail(err, errlen, GRIT_EIO, "%s: not a directory", src_dir);
  if (int rc = walk(src, "", &m.entries, &msg)) return fail(err, errlen, rc, "%s", msg.c_str());
  std::vector<const Entry*> files;
  for (auto& e : m.entries) {
    if (e.type == 0) {
      st.n_dirs++;
      continue;
    }
    e.first_page = m.n_pages;
    m.n_pages += e.pages();
    m.bytes_logical += e.size;
    st.n_files++;
    if (e.size) files.push_back(&e);
  }
  if (m.n_pages > GRIT_MAX_PAGES) return fail(err, errlen, GRIT_EINVAL, "%s: image exceeds GRIT_MAX_PAGES pages", src_dir);

  // previous checkpoint's per-page hashes, for incremental packs
  std::vector<uint64_t> p_hash;
  std::vector<uint32_t> p_ref, p_crc;
  grit_table parent{};
  if (!o.parent.empty()) {
    Manifest pm;
    if (int rc = load_manifest(o.parent, &pm, &msg)) return fail(err, errlen, rc, "parent: %s", msg.c_str());
    if (int rc = load_pagemap(o.parent, pm, &p_ref, &p_crc, &msg)) return fail(err, errlen, rc, "parent: %s", msg.c_str());
    if (int rc = load_hashes(o.parent, pm, &p_hash, &msg)) return fail(err, errlen, rc, "parent: %s", msg.c_str());
    parent.hash = p_hash.data();
    parent.crc = p_crc.data();
    parent.n_pages = pm.n_pages;
    m.parent = o.parent;
  }

  if (mkdir_all(dst) != 0) return fail(err, errlen, GRIT_EIO, "%s", errno_msg("mkdir", dst).c_str());
  // a re-run (Job retry) must never leave an old manifest vouching for new, partial data
  if (unlink((dst + "/" + kManifest).c_str()) != 0 && errno != ENOENT)
    return fail(err, errlen, GRIT_EIO, "%s", errno_msg("unlink", dst + "/" + kManifest).c_str());
  // a chunk of output should land in all the shard files at once: stripes of chunk / n_shards, 1 to 8 MiB
  const uint64_t chunk_pages = o.chunk_for(m.n_pages);
  m.stripe_slots = (uint32_t)std::min<uint64_t>(2048, std::max<uint64_t>(256, chunk_pages / m.n_shards));
  PageStore store;
  store.dir = dst;
  store.k = m.n_shards;
  store.stripe = m.stripe_slots;
  if (int rc = store.open_all(true, &msg)) return fail(err, errlen, rc, "%s", msg.c_str());
  IoPool pool(o.io_threads);

  ph.mark("walk+open");
  CtxGuard cg;  // owns the context unless a communicator brings its own
  grit_ctx* ctx = o.comm ? comm_context(o.comm) : nullptr;
  if (!ctx) {
    if (int rc = grit_ctx_create(o.device, &cg.ctx, err, errlen)) return rc;
    ctx = cg.ctx;
  }
  ph.mark("ctx");
  std::vector<uint32_t> ref(std::max<uint64_t>(m.n_pages, 1)), crc(std::max<uint64_t>(m.n_pages, 1));
  std::vector<uint64_t> hash(std::max<uint64_t>(m.n_pages, 1));
  grit_pack_stats ps{};
  // A ranked pack (grit_opts.comm): the tree goes through the GPU twice.  The first pass only analyses (K1 of every
  // chunk), then the ranks exchange their tables over NCCL, then the usual pass packs against the lower ranks' tables,
  // which never leave the device.  The host link carries the tree twice; the PV, usually the slower of the two, stores
  // what no lower rank stores.
  grit_table lower[GRIT_MAX_PEERS] = {};
  const grit_table* peer_tables = o.peer_tables;
  uint32_t n_peer_tables = o.n_peers;
  bool peers_on_device = false;
  if (o.comm) {
    const int rank = comm_rank_of(o.comm);
    if ((int)o.peer_dirs.size() != rank)
      return fail(err, errlen, GRIT_EINVAL, "rank %d must name the packed directories of the %d ranks below it (peer_dirs)", rank, rank);
    double h2d_first = 0;
    {
      PackStream pa(ctx, m.n_pages, chunk_pages);
      pa.set_analyze_only();
      if (int rc = pa.begin(nullptr, err, errlen)) return rc;
      if (m.n_pages)
        if (int rc = pa.alloc_staging(true, false, err, errlen)) return rc;
      const uint64_t cp = pa.chunk_pages(), n_chunks = (m.n_pages + cp - 1) / cp;
      for (uint64_t i = 0; i < n_chunks; i++) {
        const uint64_t p0 = i * cp, p1 = std::min(m.n_pages, p0 + cp);
        if (int rc = pa.wait_in_free(i, err, errlen)) return rc;
        char* buf = static_cast<char*>(pa.in_buffer(i));
        std::vector<Piece> pieces;
        pieces_for(files, p0, p1, buf, piece_bytes_for(cp, o.io_threads), &pieces);
        zero_tails(files, p0, p1, buf);
        if (int rc = for_pieces(&pool, pieces, &msg, [&](const Piece& p, std::string* e) { read_piece(src, p, e); }))
          return fail(err, errlen, rc, "%s", msg.c_str());
        if (int rc = pa.submit(i, buf, p1 - p0, err, errlen)) return rc;
      }
      if (n_chunks)
        if (int rc = pa.wait_chunk(n_chunks - 1, err, errlen)) return rc;
      GRIT_CUDA(cudaStreamSynchronize(ctx->stream));
      if (int rc = comm_gather_tables(o.comm, pa.device_hash(), pa.device_crc(), m.n_pages, lower, err, errlen)) return rc;
      h2d_first = pa.h2d_bytes;
    }
    st.h2d_bytes += h2d_first;
    peer_tables = lower;
    n_peer_tables = (uint32_t)rank;
    peers_on_device = true;
    ph.mark("analyze pass + exchange");
  }
  {
    PackStream pk(ctx, m.n_pages, chunk_pages);
    pk.set_outputs(ref.data(), crc.data(), hash.data());
    if (n_peer_tables > GRIT_MAX_PEERS) return fail(err, errlen, GRIT_EINVAL, "n_peers %u > %u", n_peer_tables, GRIT_MAX_PEERS);
    if (int rc = pk.begin(parent.n_pages ? &parent : nullptr, err, errlen, peer_tables, n_peer_tables, peers_on_device)) return rc;
    if (m.n_pages)
      if (int rc = pk.alloc_staging(true, true, err, errlen)) return rc;
    const uint64_t cp = pk.chunk_pages(), n_chunks = (m.n_pages + cp - 1) / cp;
    ph.mark("alloc");

    // writer thread: chunk outputs -> the page store, in order
    struct Job {
      uint64_t chunk, first_slot, n_out;
    };
    std::mutex mu;
    std::condition_variable cv;
    std::deque<Job> jobs;
    uint64_t written_chunks = 0;
    bool stop = false;
    std::string werr;
    std::thread writer([&] {
      cudaSetDevice(ctx->device);
      for (;;) {
        Job j;
        {
          std::unique_lock<std::mutex> l(mu);
          cv.wait(l, [&] { return stop || !jobs.empty(); });
          if (jobs.empty()) return;
          j = jobs.front();
          jobs.pop_front();
        }
        char e2[256] = "";
        std::string local;
        if (pk.wait_out(j.chunk, e2, sizeof e2)) local = e2;
        if (local.empty() && j.n_out &&
            store.io(j.first_slot, j.n_out, static_cast<char*>(pk.out_buffer(j.chunk)), &pool, true))
          local = errno_msg("write", store.dir + "/pages-*.bin");
        {
          std::lock_guard<std::mutex> l(mu);
          if (!local.empty() && werr.empty()) werr = local;
          written_chunks++;
        }
        cv.notify_all();
      }
    });
    auto stop_writer = [&] {
      {
        std::lock_guard<std::mutex> l(mu);
        stop = true;
      }
      cv.notify_all();
      if (writer.joinable()) writer.join();
    };
    int rc = GRIT_OK;
    for (uint64_t i = 0; i <= n_chunks && !rc; i++) {
      if (i < n_chunks) {
        const uint64_t p0 = i * cp, p1 = std::min(m.n_pages, p0 + cp);
        if ((rc = pk.wait_in_free(i, err, errlen))) break;
        char* buf = static_cast<char*>(pk.in_buffer(i));
        std::vector<Piece> pieces;
        pieces_for(files, p0, p1, buf, piece_bytes_for(cp, o.io_threads), &pieces);
        zero_tails(files, p0, p1, buf);  // the rest of a file's last page reads as zeros
        if ((rc = for_pieces(&pool, pieces, &msg, [&](const Piece& p, std::string* e) { read_piece(src, p, e); }))) {
          fail(err, errlen, rc, "%s", msg.c_str());
          break;
        }
        if ((rc = pk.submit(i, buf, p1 - p0, err, errlen))) break;
      }
      if (i >= 1) {
        const uint64_t j = i - 1;
        if (j >= (uint64_t)kPipeDepth) {  // out_buffer(j) must have been written out
          std::unique_lock<std::mutex> l(mu);
          cv.wait(l, [&] { return written_chunks + kPipeDepth > j; });
        }
        Job job{j, 0, 0};
        if ((rc = pk.collect(j, pk.out_buffer(j), &job.first_slot, &job.n_out, err, errlen))) break;
        {
          std::lock_guard<std::mutex> l(mu);
          jobs.push_back(job);
        }
        cv.notify_all();
      }
    }
    if (!rc) {
      std::unique_lock<std::mutex> l(mu);
      cv.wait(l, [&] { return written_chunks == n_chunks; });
    }
    stop_writer();
    if (!rc && !werr.empty()) rc = fail(err, errlen, GRIT_EIO, "%s", werr.c_str());
    if (rc) return rc;
    ph.mark("stream");
    if ((rc = pk.finish(&ps, err, errlen))) return rc;
    st.h2d_bytes += pk.h2d_bytes;
    st.d2h_bytes = pk.d2h_bytes;
  }
  if (ps.n_peer && !n_peer_tables) return fail(err, errlen, GRIT_EINVAL, "internal: peer references in a single-rank pack");
  m.peers = o.peer_dirs;
  ph.mark("finish+free");
  m.n_slots = ps.n_data;
  if (store.sync_all()) return fail(err, errlen, GRIT_EIO, "%s", errno_msg("fsync", store.dir + "/pages-*.bin").c_str());

  std::string pagemap(m.n_pages * 8, '\0');
  memcpy(&pagemap[0], ref.data(), m.n_pages * 4);
  memcpy(&pagemap[m.n_pages * 4], crc.data(), m.n_pages * 4);
  m.pagemap_crc = grit_crc32c(0, pagemap.data(), pagemap.size());
  m.hashes_crc = grit_crc32c(0, hash.data(), m.n_pages * 8);
  if (int rc = write_file(dst + "/pagemap.bin", pagemap.data(), pagemap.size(), &msg)) return fail(err, errlen, rc, "%s", msg.c_str());
  if (int rc = write_file(dst + "/hashes.bin", hash.data(), m.n_pages * 8, &msg)) return fail(err, errlen, rc, "%s", msg.c_str());
  if (int rc = write_criu_extents(src, dst, m, &msg)) return fail(err, errlen, rc, "%s", msg.c_str());
  // the manifest goes last, through a rename: its presence means the pack is complete
  const std::string enc = encode_manifest(m), tmp = dst + "/" + kManifest + ".tmp";
  if (int rc = write_file(tmp, enc.data(), enc.size(), &msg)) return fail(err, errlen, rc, "%s", msg.c_str());
  if (rename(tmp.c_str(), (dst + "/" + kManifest).c_str()) != 0)
    return fail(err, errlen, GRIT_EIO, "%s", errno_msg("rename", tmp).c_str());
  const int dfd = open(dst.c_str(), O_RDONLY | O_DIRECTORY | O_CLOEXEC);
  if (dfd >= 0) {
    fsync(dfd);
    close(dfd);
  }

  ph.mark("metadata+fsync");
  cg.release();
  ph.mark("ctx free");
  st.bytes_logical = m.bytes_logical;
  st.bytes_stored = m.n_slots * kPage + pagemap.size() + m.n_pages * 8 + enc.size();
  st.n_pages = ps.n_pages;
  st.n_zero = ps.n_zero;
  st.n_data = ps.n_data;
  st.n_dup = ps.n_dup;
  st.n_parent = ps.n_parent;
  st.n_peer = ps.n_peer;
  st.gpu_ms = ps.gpu_ms;
  st.seconds = now_s() - t0;
  if (stats) *stats = st;
  return GRIT_OK;
}

}  // extern "C"

namespace {

// grit_restore_dir, and with dst_dir == nullptr grit_verify_dir: the same pipeline with nothing written
int restore_or_verify(const char* src_dir, const char* dst_dir, const grit_opts* opts, grit_stats* stats, char* err,
                      size_t errlen) {
  if (!src_dir) return fail(err, errlen, GRIT_EINVAL, "null directory");
  const bool writing = dst_dir != nullptr;
  const double t0 = now_s();
  Phases ph(writing ? "restore_dir" : "verify_dir");
  Opts o = read_opts(opts);
  if (!writing) o.verify = true;
  const std::string src = src_dir, dst = writing ? dst_dir : "";
  std::string msg;
  grit_stats st{};
  if (!has_manifest(src))
    return fail(err, errlen, GRIT_EFORMAT,
                "%s holds no %s: not a complete gritpack image (a plain mirror written by the stock agent "
                "is restored by the stock agent)",
                src_dir, kManifest);
  Manifest m;
  if (int rc = load_manifest(src, &m, &msg)) return fail(err, errlen, rc, "%s", msg.c_str());
  std::vector<uint32_t> ref, crc;
  if (int rc = load_pagemap(src, m, &ref, &crc, &msg)) return fail(err, errlen, rc, "%s", msg.c_str());
  // Tree skeleton: directories 0777 & ~umask (copy.go:36), files created and sized.  It is made on a thread of its
  // own while the CUDA context, the staging buffers and the images this one refers to are set up; the first write waits.
  std::vector<const Entry*> files;
  for (const auto& e : m.entries) {
    if (e.type == 0) {
      st.n_dirs++;
    } else {
      st.n_files++;
      if (e.size && writing) files.push_back(&e);
    }
  }
  // shared mappings of the destination files (memory-backed file systems only), by entry index
  struct Maps {
    const Manifest& m;
    std::vector<char*> at;
    explicit Maps(const Manifest& mm) : m(mm), at(mm.entries.size(), nullptr) {}
    void release() {
      for (size_t i = 0; i < at.size(); i++)
        if (at[i]) munmap(at[i], m.entries[i].size), at[i] = nullptr;
    }
    ~Maps() { release(); }
  } maps(m);
  const bool map_writes = !getenv("GRITPACK_NO_MMAP");
  struct Skeleton {
    std::thread th;
    bool joined = true;
    int rc = GRIT_OK;
    std::string msg;
    std::mutex mu;  // several writer threads may arrive at once
    int wait() {
      std::lock_guard<std::mutex> l(mu);
      if (!joined) th.join();
      joined = true;
      return rc;
    }
    ~Skeleton() { wait(); }
  } skeleton;
  if (writing) {
    skeleton.joined = false;
    skeleton.th = std::thread([&] {
      auto bad = [&](const char* op, const std::string& path) { skeleton.msg = errno_msg(op, path), skeleton.rc = GRIT_EIO; };
      if (mkdir_all(dst) != 0) return bad("mkdir", dst);
      for (const auto& e : m.entries) {
        const std::string path = dst + "/" + e.path;
        if (e.type == 0) {
          if (mkdir_all(path) != 0) return bad("mkdir", path);
          continue;
        }
        const int fd = open(path.c_str(), O_RDWR | O_CREAT | O_TRUNC | O_CLOEXEC, 0666);  // os.Create, copy.go:73
        if (fd < 0) return bad("open", path);
        // reserve the blocks up front: later writes then only overwrite (faster, and no fragmentation on disk)
        const int frc = e.size ? posix_fallocate(fd, 0, (off_t)e.size) : 0;
        int trc = frc;
        if (trc != 0) trc = ftruncate(fd, (off_t)e.size);  // file systems without fallocate
        // tmpfs takes one write(2) at a time per file (3.7 GB/s whatever the thread count) but stores from any
        // number of threads through a shared mapping of pages that exist already (7.4 GB/s per file measured, then
        // bound by the per-file page-cache lock; a pwrite + mapping mix gave no more): memory-backed destinations
        // are written through a mapping
        struct statfs sfs;
        if (trc == 0 && frc == 0 && e.size && map_writes && fstatfs(fd, &sfs) == 0 && sfs.f_type == 0x01021994 /* TMPFS_MAGIC */) {
          void* a = mmap(nullptr, e.size, PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
          if (a != MAP_FAILED) maps.at[&e - m.entries.data()] = static_cast<char*>(a);
        }
        close(fd);
        if (trc != 0) return bad("truncate", path);
      }
    });
  }

  CtxGuard cg;
  if (int rc = grit_ctx_create(o.device, &cg.ctx, err, errlen)) return rc;
  grit_ctx* ctx = cg.ctx;
  ph.mark("ctx");

  // what this image refers to: the previous checkpoint, lower ranks' images on the shared PV
  ImageCache cache;
  const ImageCache::Img* parent_img = nullptr;
  PeerImages peer_imgs{};
  if (int rc = load_dependencies(ctx, src, m, ref, o.verify, &cache, 0, &st, &parent_img, &peer_imgs, &msg))
    return fail(err, errlen, rc, "%s", msg.c_str());
  void* d_parent = parent_img ? parent_img->d : nullptr;
  const uint64_t parent_pages = parent_img ? parent_img->pages : 0;

  PageStore store;
  store.dir = src;
  store.k = m.n_shards;
  store.stripe = m.stripe_slots;
  if (int rc = store.open_all(false, &msg)) return fail(err, errlen, rc, "%s", msg.c_str());
  if (int rc = store.check_sizes(m.n_slots, &msg)) return fail(err, errlen, rc, "%s", msg.c_str());

  IoPool pool(o.io_threads);
  RestoreIO io;
  io.read_packed = [&](uint64_t s0, uint64_t n, void* dstbuf) -> int {  // reader thread; split over the I/O threads
    return store.io(s0, n, static_cast<char*>(dstbuf), &pool, false);
  };
  std::mutex msg_mu;
  io.write_out = [&](uint64_t p0, uint64_t p1, const void* srcbuf) -> int {  // writer threads, several at a time
    if (!writing) return 0;  // verify only: the CRC check happened on the GPU, the bytes are dropped
    std::vector<Piece> pieces;
    std::string local;
    int rc = skeleton.wait();
    if (rc) local = skeleton.msg;
    if (!rc) {
      pieces_for(files, p0, p1, const_cast<char*>(static_cast<const char*>(srcbuf)), piece_bytes_for(p1 - p0, o.io_threads), &pieces);
      rc = for_pieces(&pool, pieces, &local, [&](const Piece& p, std::string* e) {
        if (char* base = maps.at[p.e - m.entries.data()])
          memcpy(base + p.file_off, p.buf, p.len);
        else
          write_piece(dst, p, e);
      });
    }
    if (rc) {
      std::lock_guard<std::mutex> l(msg_mu);
      if (msg.empty()) msg = local;
    }
    return rc;
  };
  grit_restore_stats rs{};
  {
    std::lock_guard<std::mutex> g(ctx->mu);
    // a big tree is restored in interleaved runs of its page range, several destination files being written at once
    const bool big = writing && m.n_pages * kPage >= (2ull << 30);
    const int rc = restore_stream(ctx, io, m.n_slots, ref.data(), crc.data(), m.n_pages, d_parent, parent_pages, true,
                                  o.verify, o.chunk_for(m.n_pages), &rs, &st.h2d_bytes, &st.d2h_bytes, err, errlen, &peer_imgs,
                                  big ? o.lanes : 1, big ? 5 : kPipeDepth);
    st.n_bad += rs.n_bad;
    if (rc) {
      if (!msg.empty() && err && errlen) snprintf(err, errlen, "%s", msg.c_str());
      if (stats) *stats = st;
      return rc;
    }
  }
  ph.mark("stream");
  if (int rc = skeleton.wait()) return fail(err, errlen, rc, "%s", skeleton.msg.c_str());  // an image without a stored byte
  maps.release();
  // os.Chmod(dstFile, info.Mode()), copy.go:84-89; then make the bytes durable before
  // the caller writes the sentinel that patched containerd polls for (restore.go:20)
  for (const auto& e : m.entries) {
    if (e.type != 1 || !writing) continue;
    const std::string path = dst + "/" + e.path;
    const int fd = open(path.c_str(), O_WRONLY | O_CLOEXEC);
    if (fd < 0) return fail(err, errlen, GRIT_EIO, "%s", errno_msg("open", path).c_str());
    const int r1 = fsync(fd), r2 = fchmod(fd, e.mode);
    close(fd);
    if (r1 != 0 || r2 != 0) return fail(err, errlen, GRIT_EIO, "%s", errno_msg("chmod", path).c_str());
  }

  ph.mark("fsync+chmod");
  cache.release();
  cg.release();
  ph.mark("ctx free");
  st.bytes_logical = m.bytes_logical;
  st.bytes_stored = m.n_slots * kPage;
  st.n_pages = m.n_pages;
  for (uint64_t p = 0; p < m.n_pages; p++) {
    switch (GRIT_REF_KIND(ref[p])) {
      case GRIT_KIND_ZERO: st.n_zero++; break;
      case GRIT_KIND_DATA: st.n_data++; break;
      case GRIT_KIND_DUP: st.n_dup++; break;
      case GRIT_KIND_PARENT: st.n_parent++; break;
      case GRIT_KIND_PEER: st.n_peer++; break;
      default: break;
    }
  }
  st.gpu_ms += rs.gpu_ms;
  st.seconds = now_s() - t0;
  if (stats) *stats = st;
  return GRIT_OK;
}

}  // namespace

extern "C" {

int grit_restore_dir(const char* src_dir, const char* dst_dir, const grit_opts* opts, grit_stats* stats, char* err,
                     size_t errlen) {
  if (!dst_dir) return fail(err, errlen, GRIT_EINVAL, "null directory");
  return restore_or_verify(src_dir, dst_dir, opts, stats, err, errlen);
}

int grit_verify_dir(const char* src_dir, const grit_opts* opts, grit_stats* stats, char* err, size_t errlen) {
  return restore_or_verify(src_dir, nullptr, opts, stats, err, errlen);
}

int grit_image_deps(const char* dir, int recursive, char* out, size_t outlen, size_t* needed, char* err, size_t errlen) {
  if (!dir) return fail(err, errlen, GRIT_EINVAL, "null directory");
  std::vector<std::string> deps, todo{dir};
  std::string msg;
  for (size_t i = 0; i < todo.size(); i++) {  // breadth first; todo[0] is the image itself
    if (i > 0 && !recursive) break;
    if (!has_manifest(todo[i]))
      return fail(err, errlen, GRIT_EFORMAT, "%s holds no %s: not a complete gritpack image", todo[i].c_str(), kManifest);
    Manifest m;
    if (int rc = load_manifest(todo[i], &m, &msg)) return fail(err, errlen, rc, "%s", msg.c_str());
    std::vector<uint32_t> ref, crc;
    if (int rc = load_pagemap(todo[i], m, &ref, &crc, &msg)) return fail(err, errlen, rc, "%s", msg.c_str());
    bool used[GRIT_MAX_PEERS] = {};
    if (int rc = peers_used(m, ref, used, &msg)) return fail(err, errlen, rc, "%s: %s", todo[i].c_str(), msg.c_str());
    auto add = [&](const std::string& d) {
      const std::string r = resolve_parent(todo[i], d);
      if (r == todo[0] || std::find(deps.begin(), deps.end(), r) != deps.end()) return;
      deps.push_back(r);
      todo.push_back(r);
    };
    if (!m.parent.empty()) add(m.parent);  // grit_restore_dir loads the parent whenever one is named
    for (uint32_t r = 0; r < m.peers.size(); r++)
      if (used[r]) add(m.peers[r]);
  }
  size_t need = 1;
  for (const auto& d : deps) need += d.size() + 1;
  if (needed) *needed = need;
  if (!out || outlen < need) return fail(err, errlen, GRIT_EINVAL, "dependency list needs %zu bytes", need);
  char* w = out;
  for (const auto& d : deps) {
    memcpy(w, d.c_str(), d.size() + 1);
    w += d.size() + 1;
  }
  *w = 0;
  return GRIT_OK;
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------------------------
// CRIU-aware ingest.  The directory this library packs is a CRIU image directory
// (pkg/gritagent/checkpoint/runtime.go:122-134 has CRIU write it); the reference treats it as opaque files.  CRIU
// itself is not under /root/reference: the image format below is the one CRIU publishes (criu.org/Images, CRIU 3.x/4.x):
//   pagemap-<id>.img   u32 0x54564319 (IMG_COMMON_MAGIC), u32 0x56084025 (PAGEMAP_MAGIC), then records
//                      { u32 size; protobuf message of `size` bytes }: first a pagemap_head { uint32 pages_id = 1 },
//                      then pagemap_entry { uint64 vaddr = 1; uint32 nr_pages = 2; bool in_parent = 3; uint32 flags = 4 }
//                      with flags PE_PARENT 1, PE_LAZY 2, PE_PRESENT 4
//   pages-<pages_id>.img   the raw 4 KiB pages of every entry that is present (not in_parent), in entry order
// CRIU's own incremental dumps mark a run "in_parent" when the parent image holds the same virtual address range
// unchanged; CRIU never fills that in here because the agent never names a parent
// (cmd/containerd-shim-grit-v1/runc/container.go:107, process/init.go:192 plumb ParentPath, nothing sets it).
// The page hashes computed on the GPU can: a page of pages-N.img whose packed reference is PARENT and whose parent
// page is the page of the same virtual address in the parent's image is exactly a page CRIU would have left out.
namespace {

constexpr uint32_t kCriuCommonMagic = 0x54564319u, kCriuPagemapMagic = 0x56084025u;
constexpr uint32_t kPeParent = 1, kPePresent = 4;
const char* const kExtents = "criu-extents.txt";

struct CriuEntry {
  uint64_t vaddr = 0;
  uint32_t nr_pages = 0, flags = 0;
  bool in_parent = false;
  bool has_flags = false;
  bool present() const { return !(in_parent || (flags & kPeParent)); }
};
struct CriuPagemap {
  uint32_t pages_id = 0;
  std::vector<CriuEntry> entries;
};

bool get_varint(const unsigned char*& p, const unsigned char* end, uint64_t* v) {
  *v = 0;
  for (int shift = 0; p < end && shift < 64; shift += 7) {
    const unsigned char b = *p++;
    *v |= (uint64_t)(b & 0x7F) << shift;
    if (!(b & 0x80)) return true;
  }
  return false;
}
void put_varint(std::string* b, uint64_t v) {
  while (v >= 0x80) {
    b->push_back((char)(v | 0x80));
    v >>= 7;
  }
  b->push_back((char)v);
}
// one protobuf message: varint fields into f[1..4], everything else skipped
bool parse_fields(const unsigned char* p, const unsigned char* end, uint64_t (&f)[5], bool (&have)[5]) {
  while (p < end) {
    uint64_t key, v;
    if (!get_varint(p, end, &key)) return false;
    const uint32_t field = (uint32_t)(key >> 3), wt = (uint32_t)(key & 7);
    if (wt == 0) {
      if (!get_varint(p, end, &v)) return false;
      if (field >= 1 && field <= 4) f[field] = v, have[field] = true;
    } else if (wt == 2) {
      if (!get_varint(p, end, &v) || v > (uint64_t)(end - p)) return false;
      p += v;
    } else if (wt == 1 || wt == 5) {
      const size_t n = wt == 1 ? 8 : 4;
      if ((size_t)(end - p) < n) return false;
      p += n;
    } else {
      return false;
    }
  }
  return true;
}

// false: not a pagemap image (the file is then just opaque bytes to the packer)
bool parse_criu_pagemap(const std::string& bytes, CriuPagemap* out) {
  if (bytes.size() < 8) return false;
  uint32_t m[2];
  memcpy(m, bytes.data(), 8);
  if (m[0] != kCriuCommonMagic || m[1] != kCriuPagemapMagic) return false;
  const unsigned char* p = reinterpret_cast<const unsigned char*>(bytes.data()) + 8;
  const unsigned char* end = p + bytes.size() - 8;
  bool head = true;
  while (p < end) {
    if (end - p < 4) return false;
    uint32_t size;
    memcpy(&size, p, 4);
    p += 4;
    if (size > (uint64_t)(end - p)) return false;
    uint64_t f[5] = {};
    bool have[5] = {};
    if (!parse_fields(p, p + size, f, have)) return false;
    p += size;
    if (head) {
      if (!have[1]) return false;
      out->pages_id = (uint32_t)f[1];
      head = false;
      continue;
    }
    if (!have[1] || !have[2]) return false;
    CriuEntry e;
    e.vaddr = f[1];
    e.nr_pages = (uint32_t)f[2];
    e.in_parent = have[3] && f[3];
    e.has_flags = have[4];
    e.flags = (uint32_t)f[4];
    out->entries.push_back(e);
  }
  return !head;
}

std::string encode_criu_pagemap(const CriuPagemap& pm) {
  std::string b;
  put(&b, kCriuCommonMagic);
  put(&b, kCriuPagemapMagic);
  auto record = [&](const std::string& msg) {
    put<uint32_t>(&b, (uint32_t)msg.size());
    b += msg;
  };
  std::string m;
  m.push_back(0x08);
  put_varint(&m, pm.pages_id);
  record(m);
  for (const auto& e : pm.entries) {
    m.clear();
    m.push_back(0x08);
    put_varint(&m, e.vaddr);
    m.push_back(0x10);
    put_varint(&m, e.nr_pages);
    if (e.in_parent) {
      m.push_back(0x18);
      put_varint(&m, 1);
    }
    if (e.has_flags || e.in_parent) {
      m.push_back(0x20);
      put_varint(&m, e.flags);
    }
    record(m);
  }
  return b;
}

}  // namespace

namespace {

struct ExtentFile {
  std::string pagemap_path, pages_path;  // relative to the tree root
  uint64_t first_page = 0;               // stream page of the pages file's first page
  CriuPagemap pm;
};

std::string dir_of(const std::string& rel) {
  const size_t s = rel.find_last_of('/');
  return s == std::string::npos ? std::string() : rel.substr(0, s + 1);
}

// Which stream pages belong to which pagemap entry: every pagemap-*.img of the tree that parses, with the pages file
// its head names.  Written beside the packed image by grit_pack_dir; an image without CRIU page maps gets none.
int collect_extents(const std::string& src, const Manifest& m, std::vector<ExtentFile>* out) {
  for (const auto& e : m.entries) {
    if (e.type != 1) continue;
    const size_t slash = e.path.find_last_of('/');
    const std::string name = slash == std::string::npos ? e.path : e.path.substr(slash + 1);
    if (name.compare(0, 8, "pagemap-") != 0 || name.size() < 5 || name.compare(name.size() - 4, 4, ".img") != 0) continue;
    if (e.size > (64u << 20)) continue;
    std::string bytes, msg;
    if (read_file(src + "/" + e.path, &bytes, &msg)) continue;
    ExtentFile x;
    if (!parse_criu_pagemap(bytes, &x.pm)) continue;
    x.pagemap_path = e.path;
    x.pages_path = dir_of(e.path) + "pages-" + std::to_string(x.pm.pages_id) + ".img";
    uint64_t present = 0;
    for (const auto& pe : x.pm.entries)
      if (pe.present()) present += pe.nr_pages;
    bool found = false;
    for (const auto& f : m.entries)
      if (f.type == 1 && f.path == x.pages_path) {
        found = f.size == present * kPage;  // the page map and its pages file must agree
        x.first_page = f.first_page;
      }
    if (found) out->push_back(std::move(x));
  }
  return GRIT_OK;
}

std::string encode_extents(const std::vector<ExtentFile>& xs) {
  std::string b = "gritpack-criu-extents 1\n";
  char line[256];
  for (const auto& x : xs) {
    // paths go last on their own lines: they may hold spaces
    snprintf(line, sizeof line, "pagemap %u %llu %zu\n", x.pm.pages_id, (unsigned long long)x.first_page, x.pm.entries.size());
    b += line;
    b += x.pagemap_path + "\n" + x.pages_path + "\n";
    for (const auto& e : x.pm.entries) {
      snprintf(line, sizeof line, "e %llx %u %u %d %d\n", (unsigned long long)e.vaddr, e.nr_pages, e.flags, e.in_parent ? 1 : 0,
               e.has_flags ? 1 : 0);
      b += line;
    }
  }
  return b;
}

int load_extents(const std::string& packed_dir, std::vector<ExtentFile>* out, std::string* msg) {
  std::string b;
  if (int rc = read_file(packed_dir + "/" + kExtents, &b, msg)) return rc;
  size_t pos = 0;
  auto next_line = [&](std::string* l) {
    if (pos >= b.size()) return false;
    const size_t e = b.find('\n', pos);
    *l = b.substr(pos, (e == std::string::npos ? b.size() : e) - pos);
    pos = e == std::string::npos ? b.size() : e + 1;
    return true;
  };
  std::string l;
  if (!next_line(&l) || l != "gritpack-criu-extents 1") return *msg = packed_dir + "/" + kExtents + ": unknown format", GRIT_EFORMAT;
  while (next_line(&l)) {
    ExtentFile x;
    unsigned long long first = 0;
    size_t n = 0;
    if (sscanf(l.c_str(), "pagemap %u %llu %zu", &x.pm.pages_id, &first, &n) != 3 || !next_line(&x.pagemap_path) ||
        !next_line(&x.pages_path))
      return *msg = packed_dir + "/" + kExtents + ": damaged", GRIT_EFORMAT;
    x.first_page = first;
    for (size_t i = 0; i < n; i++) {
      CriuEntry e;
      unsigned long long va = 0;
      int inp = 0, hf = 0;
      if (!next_line(&l) || sscanf(l.c_str(), "e %llx %u %u %d %d", &va, &e.nr_pages, &e.flags, &inp, &hf) != 5)
        return *msg = packed_dir + "/" + kExtents + ": damaged", GRIT_EFORMAT;
      e.vaddr = va;
      e.in_parent = inp != 0;
      e.has_flags = hf != 0;
      x.pm.entries.push_back(e);
    }
    out->push_back(std::move(x));
  }
  return GRIT_OK;
}

int write_criu_extents(const std::string& src, const std::string& dst, const Manifest& m, std::string* msg) {
  std::vector<ExtentFile> xs;
  collect_extents(src, m, &xs);
  const std::string path = dst + "/" + kExtents;
  if (xs.empty()) {
    unlink(path.c_str());  // a re-run over an older pack of another tree
    return GRIT_OK;
  }
  const std::string enc = encode_extents(xs);
  return write_file(path, enc.data(), enc.size(), msg);
}

}  // namespace

extern "C" {

int grit_criu_pagemap_info(const char* path, uint32_t* pages_id, uint64_t* n_entries, uint64_t* pages_present,
                           uint64_t* pages_in_parent, char* err, size_t errlen) {
  if (!path) return fail(err, errlen, GRIT_EINVAL, "null path");
  std::string bytes, msg;
  if (int rc = read_file(path, &bytes, &msg)) return fail(err, errlen, rc, "%s", msg.c_str());
  CriuPagemap pm;
  if (!parse_criu_pagemap(bytes, &pm)) return fail(err, errlen, GRIT_EFORMAT, "%s: not a CRIU pagemap image", path);
  uint64_t present = 0, inpar = 0;
  for (const auto& e : pm.entries) (e.present() ? present : inpar) += e.nr_pages;
  if (pages_id) *pages_id = pm.pages_id;
  if (n_entries) *n_entries = pm.entries.size();
  if (pages_present) *pages_present = present;
  if (pages_in_parent) *pages_in_parent = inpar;
  return GRIT_OK;
}

int grit_criu_incremental(const char* src_dir, const char* packed_dir, const char* out_dir, grit_criu_stats* stats,
                          char* err, size_t errlen) {
  if (!src_dir || !packed_dir || !out_dir) return fail(err, errlen, GRIT_EINVAL, "null directory");
  const std::string src = src_dir, packed = packed_dir, out = out_dir;
  std::string msg;
  grit_criu_stats st{};
  Manifest m;
  if (!has_manifest(packed)) return fail(err, errlen, GRIT_EFORMAT, "%s holds no %s", packed_dir, kManifest);
  if (int rc = load_manifest(packed, &m, &msg)) return fail(err, errlen, rc, "%s", msg.c_str());
  if (m.parent.empty()) return fail(err, errlen, GRIT_EINVAL, "%s was packed without a parent: nothing can be left to one", packed_dir);
  std::vector<uint32_t> ref, crc;
  if (int rc = load_pagemap(packed, m, &ref, &crc, &msg)) return fail(err, errlen, rc, "%s", msg.c_str());
  std::vector<ExtentFile> mine, theirs;
  if (int rc = load_extents(packed, &mine, &msg)) return fail(err, errlen, rc, "%s", msg.c_str());
  const std::string parent_dir = resolve_parent(packed, m.parent);
  if (int rc = load_extents(parent_dir, &theirs, &msg)) return fail(err, errlen, rc, "parent: %s", msg.c_str());
  // A PARENT reference names the lowest parent page with that content; the page at the same address may be a later
  // copy of it.  Equal (hash64, CRC-32C) in the parent's own tables is what "the same page" means here.
  Manifest pm;
  std::vector<uint32_t> p_ref, p_crc;
  std::vector<uint64_t> p_hash;
  if (int rc = load_manifest(parent_dir, &pm, &msg)) return fail(err, errlen, rc, "parent: %s", msg.c_str());
  if (int rc = load_pagemap(parent_dir, pm, &p_ref, &p_crc, &msg)) return fail(err, errlen, rc, "parent: %s", msg.c_str());
  if (int rc = load_hashes(parent_dir, pm, &p_hash, &msg)) return fail(err, errlen, rc, "parent: %s", msg.c_str());

  for (const auto& x : mine) {
    // the parent's pages of the same page map, by virtual address: (vaddr, pages, first stream page), sorted
    struct Run {
      uint64_t vaddr, nr, first;
    };
    std::vector<Run> runs;
    for (const auto& y : theirs)
      if (y.pagemap_path == x.pagemap_path) {
        uint64_t fp = y.first_page;
        for (const auto& e : y.pm.entries)
          if (e.present()) {
            runs.push_back({e.vaddr, e.nr_pages, fp});
            fp += e.nr_pages;
          }
      }
    std::sort(runs.begin(), runs.end(), [](const Run& a, const Run& b) { return a.vaddr < b.vaddr; });
    auto parent_page_at = [&](uint64_t vaddr, uint64_t* page) {
      auto it = std::upper_bound(runs.begin(), runs.end(), vaddr, [](uint64_t v, const Run& r) { return v < r.vaddr; });
      if (it == runs.begin()) return false;
      --it;
      if (vaddr >= it->vaddr + it->nr * kPage) return false;
      *page = it->first + (vaddr - it->vaddr) / kPage;
      return true;
    };
    CriuPagemap outpm;
    outpm.pages_id = x.pm.pages_id;
    std::vector<uint64_t> keep;  // pages of the source pages file that stay, in order
    uint64_t file_page = 0;
    for (const auto& e : x.pm.entries) {
      st.entries_in++;
      if (!e.present()) {  // CRIU's own in_parent run: passed through
        outpm.entries.push_back(e);
        continue;
      }
      uint32_t i = 0;
      while (i < e.nr_pages) {
        auto in_parent = [&](uint32_t k) {
          const uint64_t p = x.first_page + file_page + k;
          uint64_t q;
          if (p >= m.n_pages || GRIT_REF_KIND(ref[p]) != GRIT_KIND_PARENT || !parent_page_at(e.vaddr + (uint64_t)k * kPage, &q))
            return false;
          const uint64_t owner = GRIT_REF_VALUE(ref[p]);
          return q < pm.n_pages && owner < pm.n_pages && p_hash[q] == p_hash[owner] && p_crc[q] == p_crc[owner];
        };
        const bool par = in_parent(i);
        uint32_t j = i + 1;
        while (j < e.nr_pages && in_parent(j) == par) j++;
        CriuEntry o = e;
        o.vaddr = e.vaddr + (uint64_t)i * kPage;
        o.nr_pages = j - i;
        if (par) {
          o.in_parent = true;
          o.has_flags = true;
          o.flags = (e.flags & ~kPePresent) | kPeParent;
          st.pages_in_parent += j - i;
        } else {
          for (uint32_t k = i; k < j; k++) keep.push_back(file_page + k);
        }
        outpm.entries.push_back(o);
        i = j;
      }
      file_page += e.nr_pages;
      st.pages_total += e.nr_pages;
    }
    st.entries_out += outpm.entries.size();
    st.n_pagemaps++;
    // the rewritten pair
    const std::string pm_path = out + "/" + x.pagemap_path, pg_path = out + "/" + x.pages_path;
    if (mkdir_all(out + "/" + dir_of(x.pagemap_path)) != 0) return fail(err, errlen, GRIT_EIO, "%s", errno_msg("mkdir", out).c_str());
    const std::string enc = encode_criu_pagemap(outpm);
    if (int rc = write_file(pm_path, enc.data(), enc.size(), &msg)) return fail(err, errlen, rc, "%s", msg.c_str());
    const int in = open((src + "/" + x.pages_path).c_str(), O_RDONLY | O_CLOEXEC);
    const int of = open(pg_path.c_str(), O_WRONLY | O_CREAT | O_TRUNC | O_CLOEXEC, 0600);
    if (in < 0 || of < 0) {
      if (in >= 0) close(in);
      if (of >= 0) close(of);
      return fail(err, errlen, GRIT_EIO, "%s", errno_msg("open", in < 0 ? src + "/" + x.pages_path : pg_path).c_str());
    }
    std::vector<char> buf(256 * kPage);
    bool bad = false;
    for (size_t a = 0; a < keep.size() && !bad;) {  // runs of consecutive pages move together
      size_t b = a + 1;
      while (b < keep.size() && keep[b] == keep[b - 1] + 1 && b - a < 256) b++;
      const size_t n = (b - a) * kPage;
      bad = pread(in, buf.data(), n, (off_t)(keep[a] * kPage)) != (ssize_t)n || write(of, buf.data(), n) != (ssize_t)n;
      a = b;
    }
    close(in);
    bad = (close(of) != 0) || bad;
    if (bad) return fail(err, errlen, GRIT_EIO, "%s", errno_msg("copy", pg_path).c_str());
    st.bytes_written += keep.size() * kPage + enc.size();
  }
  if (stats) *stats = st;
  return GRIT_OK;
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------------------------
// Serving a packed checkpoint page by page.  The reference's restore lands the whole tree on local
// disk before the runtime may start (restore.go:16-20; containerd waits for the sentinel,
// contrib/containerd/grit-interceptor.diff:146-165), then CRIU reads it all again
// (cmd/containerd-shim-grit-v1/process/init_state.go:163).  A grit_image keeps the packed checkpoint open instead: the
// stored pages stream into HBM in the background, and any page range, or any byte range of any file of the tree,
// can be asked for as soon as the slots it needs have arrived -- restored and CRC-checked on the GPU, delivered to
// host memory (a lazy-pages server answering CRIU's page faults, a FUSE view of the tree) or left in device memory
// (device state that never needs to touch the host again).
struct grit_image {
  CtxGuard cg;
  std::string dir;
  Manifest m;
  std::vector<uint32_t> ref, crc;
  ImageCache cache;  // parent / lower-rank images this one refers to, restored on the device
  PeerImages imgs{};
  PageStore store;
  void* d_packed = nullptr;
  uint32_t *d_ref = nullptr, *d_crc = nullptr;
  void* d_out = nullptr;    // one chunk of restored pages
  void* h_stage = nullptr;  // pinned, one chunk
  uint64_t chunk_pages = 0;
  int verify = 1;
  // background upload of the stored pages, in slot order
  std::thread uploader;
  std::mutex mu;
  std::condition_variable cv;
  uint64_t slots_resident = 0;
  int upload_rc = 0;
  bool stop = false;
  std::mutex read_mu;  // one read at a time
  uint64_t reads = 0, pages_served = 0;
  ~grit_image() {
    {
      std::lock_guard<std::mutex> l(mu);
      stop = true;
    }
    cv.notify_all();
    if (uploader.joinable()) uploader.join();
    if (cg.ctx) {
      cudaSetDevice(cg.ctx->device);
      cudaDeviceSynchronize();
    }
    cudaFree(d_packed);
    cudaFree(d_ref);
    cudaFree(d_crc);
    cudaFree(d_out);
    cudaFreeHost(h_stage);
    cache.release();
  }
};

extern "C" {

int grit_image_open(const char* packed_dir, const grit_opts* opts, grit_image** out, char* err, size_t errlen) {
  if (!packed_dir || !out) return fail(err, errlen, GRIT_EINVAL, "null argument");
  *out = nullptr;
  const Opts o = read_opts(opts);
  std::string msg;
  if (!has_manifest(packed_dir)) return fail(err, errlen, GRIT_EFORMAT, "%s holds no %s", packed_dir, kManifest);
  std::unique_ptr<grit_image> im(new (std::nothrow) grit_image());
  if (!im) return fail(err, errlen, GRIT_ENOMEM, "out of host memory");
  im->dir = packed_dir;
  im->verify = o.verify ? 1 : 0;
  if (int rc = load_manifest(im->dir, &im->m, &msg)) return fail(err, errlen, rc, "%s", msg.c_str());
  if (int rc = load_pagemap(im->dir, im->m, &im->ref, &im->crc, &msg)) return fail(err, errlen, rc, "%s", msg.c_str());
  if (int rc = grit_ctx_create(o.device, &im->cg.ctx, err, errlen)) return rc;
  grit_ctx* ctx = im->cg.ctx;
  const Manifest& m = im->m;
  const ImageCache::Img* parent = nullptr;
  grit_stats st{};
  if (int rc = load_dependencies(ctx, im->dir, m, im->ref, im->verify, &im->cache, 0, &st, &parent, &im->imgs, &msg))
    return fail(err, errlen, rc, "%s", msg.c_str());
  im->imgs.parent = parent ? static_cast<const unsigned char*>(parent->d) : nullptr;
  im->imgs.parent_pages = parent ? parent->pages : 0;
  im->store.dir = im->dir;
  im->store.k = m.n_shards;
  im->store.stripe = m.stripe_slots;
  if (int rc = im->store.open_all(false, &msg)) return fail(err, errlen, rc, "%s", msg.c_str());
  if (int rc = im->store.check_sizes(m.n_slots, &msg)) return fail(err, errlen, rc, "%s", msg.c_str());
  im->chunk_pages = (16ull << 20) / kPage;
  GRIT_CUDA(cudaMalloc(&im->d_packed, std::max<uint64_t>(m.n_slots, 1) * kPage));
  GRIT_CUDA(cudaMalloc(&im->d_ref, std::max<uint64_t>(m.n_pages, 1) * 4));
  GRIT_CUDA(cudaMalloc(&im->d_crc, std::max<uint64_t>(m.n_pages, 1) * 4));
  GRIT_CUDA(cudaMalloc(&im->d_out, im->chunk_pages * kPage));
  GRIT_CUDA(cudaMallocHost(&im->h_stage, im->chunk_pages * kPage));
  GRIT_CUDA(cudaMemcpy(im->d_ref, im->ref.data(), m.n_pages * 4, cudaMemcpyHostToDevice));
  GRIT_CUDA(cudaMemcpy(im->d_crc, im->crc.data(), m.n_pages * 4, cudaMemcpyHostToDevice));
  grit_image* g = im.get();
  const unsigned io_threads = o.io_threads;
  g->uploader = std::thread([g, io_threads] {
    cudaSetDevice(g->cg.ctx->device);
    const uint64_t step = (8ull << 20) / kPage, n_slots = g->m.n_slots;
    void* h[2] = {nullptr, nullptr};
    cudaEvent_t ev[2] = {nullptr, nullptr};
    int rc = 0;
    for (int b = 0; b < 2 && !rc; b++)
      if (cudaMallocHost(&h[b], step * kPage) != cudaSuccess || cudaEventCreateWithFlags(&ev[b], cudaEventDisableTiming) != cudaSuccess)
        rc = GRIT_ECUDA;
    IoPool pool(io_threads);
    uint64_t enq[2] = {0, 0};  // slots covered once buffer b's copy has landed
    const uint64_t n_chunks = (n_slots + step - 1) / step;
    for (uint64_t k = 0; k < n_chunks && !rc; k++) {
      const int b = (int)(k & 1);
      if (k >= 2) {  // buffer b's previous copy must have landed before it is refilled: that also moves the watermark
        if (cudaEventSynchronize(ev[b]) != cudaSuccess) rc = GRIT_ECUDA;
        std::lock_guard<std::mutex> l(g->mu);
        g->slots_resident = std::max(g->slots_resident, enq[b]);
        g->cv.notify_all();
      }
      {
        std::lock_guard<std::mutex> l(g->mu);
        if (g->stop) break;
      }
      const uint64_t s0 = k * step, n = std::min(step, n_slots - s0);
      if (!rc && g->store.io(s0, n, static_cast<char*>(h[b]), &pool, false)) rc = GRIT_EIO;
      if (!rc && cudaMemcpyAsync(static_cast<char*>(g->d_packed) + s0 * kPage, h[b], n * kPage, cudaMemcpyHostToDevice,
                                 g->cg.ctx->h2d) != cudaSuccess)
        rc = GRIT_ECUDA;
      if (!rc && cudaEventRecord(ev[b], g->cg.ctx->h2d) != cudaSuccess) rc = GRIT_ECUDA;
      enq[b] = s0 + n;
    }
    cudaStreamSynchronize(g->cg.ctx->h2d);
    for (int b = 0; b < 2; b++) {
      if (h[b]) cudaFreeHost(h[b]);
      if (ev[b]) cudaEventDestroy(ev[b]);
    }
    std::lock_guard<std::mutex> l(g->mu);
    if (rc) g->upload_rc = rc;
    else if (!g->stop) g->slots_resident = n_slots;
    g->cv.notify_all();
  });
  *out = im.release();
  return GRIT_OK;
}

void grit_image_close(grit_image* im) { delete im; }

}  // extern "C"

namespace {

// pages [p0, p1) of the image -> dst (host memory, or device memory of the image's GPU)
int image_read_pages(grit_image* im, uint64_t p0, uint64_t p1, char* dst, bool dst_on_device, uint64_t skip_first,
                     uint64_t len_bytes, char* err, size_t errlen) {
  // skip_first / len_bytes: the byte window of the page run that the caller wants (file reads are not page aligned)
  grit_ctx* ctx = im->cg.ctx;
  const Manifest& m = im->m;
  GRIT_CUDA(cudaSetDevice(ctx->device));
  std::lock_guard<std::mutex> g(im->read_mu);
  uint64_t done_bytes = 0;
  for (uint64_t a = p0; a < p1; a += im->chunk_pages) {
    const uint64_t b = std::min(p1, a + im->chunk_pages);
    uint64_t need = 0;  // slots [0, need) must be on the device
    for (uint64_t p = a; p < b; p++) {
      uint32_t x = im->ref[p];
      if (GRIT_REF_KIND(x) == GRIT_KIND_DUP && GRIT_REF_VALUE(x) < m.n_pages) x = im->ref[GRIT_REF_VALUE(x)];
      if (GRIT_REF_KIND(x) == GRIT_KIND_DATA) need = std::max<uint64_t>(need, (uint64_t)GRIT_REF_VALUE(x) + 1);
    }
    need = std::min(need, m.n_slots);
    {
      std::unique_lock<std::mutex> l(im->mu);
      im->cv.wait(l, [&] { return im->upload_rc || im->slots_resident >= need; });
      if (im->upload_rc) return fail(err, errlen, im->upload_rc, "%s: reading the stored pages failed", im->dir.c_str());
    }
    std::lock_guard<std::mutex> cl(ctx->mu);
    cudaStream_t s = ctx->stream;
    if (int rc = reset_counters(ctx, s, err, errlen)) return rc;
    GRIT_CUDA(launch_restore(ctx->geom, ctx->d_tables, im->d_packed, m.n_slots, im->d_ref, im->d_crc, m.n_pages, a, b, im->imgs,
                             im->d_out, im->verify, ctx->d_cnt, s));
    ctx->launches++;
    // the byte window of this chunk
    const uint64_t c0 = (a - p0) * kPage, c1 = (b - p0) * kPage;
    const uint64_t w0 = std::max(c0, skip_first), w1 = std::min(c1, skip_first + len_bytes);
    if (w1 > w0) {
      const char* src = static_cast<const char*>(im->d_out) + (w0 - c0);
      if (dst_on_device) {
        GRIT_CUDA(cudaMemcpyAsync(dst + done_bytes, src, w1 - w0, cudaMemcpyDeviceToDevice, s));
      } else {
        GRIT_CUDA(cudaMemcpyAsync(im->h_stage, src, w1 - w0, cudaMemcpyDeviceToHost, s));
      }
    }
    if (int rc = read_counters(ctx, s, err, errlen)) return rc;  // syncs
    if (ctx->h_cnt->n_bad)
      return fail(err, errlen, GRIT_ECRC, "%s: %llu page(s) failed CRC-32C or had a bad reference, first page %llu", im->dir.c_str(),
                  (unsigned long long)ctx->h_cnt->n_bad, (unsigned long long)ctx->h_cnt->first_bad);
    if (w1 > w0) {
      if (!dst_on_device) memcpy(dst + done_bytes, im->h_stage, w1 - w0);
      done_bytes += w1 - w0;
    }
    im->pages_served += b - a;
  }
  im->reads++;
  return GRIT_OK;
}

}  // namespace

extern "C" {

int grit_image_read_pages(grit_image* im, uint64_t page0, uint64_t n_pages, void* dst, int dst_on_device, char* err,
                          size_t errlen) {
  if (!im || (n_pages && !dst)) return fail(err, errlen, GRIT_EINVAL, "null argument");
  if (page0 > im->m.n_pages || n_pages > im->m.n_pages - page0)
    return fail(err, errlen, GRIT_EINVAL, "pages %llu..%llu are outside the image (%llu pages)", (unsigned long long)page0,
                (unsigned long long)(page0 + n_pages), (unsigned long long)im->m.n_pages);
  return image_read_pages(im, page0, page0 + n_pages, static_cast<char*>(dst), dst_on_device != 0, 0, n_pages * kPage, err, errlen);
}

int grit_image_stat(grit_image* im, const char* relpath, uint64_t* size, uint32_t* mode, uint64_t* first_page, char* err,
                    size_t errlen) {
  if (!im || !relpath) return fail(err, errlen, GRIT_EINVAL, "null argument");
  for (const auto& e : im->m.entries)
    if (e.path == relpath) {
      if (size) *size = e.size;
      if (mode) *mode = e.mode | (e.type == 0 ? 040000u : 0100000u);
      if (first_page) *first_page = e.first_page;
      return GRIT_OK;
    }
  return fail(err, errlen, GRIT_EIO, "%s: no such file in the image", relpath);
}

int64_t grit_image_read(grit_image* im, const char* relpath, uint64_t offset, uint64_t len, void* h_dst, char* err,
                        size_t errlen) {
  if (!im || !relpath || (len && !h_dst)) return fail(err, errlen, GRIT_EINVAL, "null argument");
  for (const auto& e : im->m.entries)
    if (e.type == 1 && e.path == relpath) {
      if (offset >= e.size) return 0;  // like pread(2) at or past the end
      len = std::min(len, e.size - offset);
      const uint64_t p0 = e.first_page + offset / kPage, p1 = e.first_page + (offset + len + kPage - 1) / kPage;
      if (int rc = image_read_pages(im, p0, p1, static_cast<char*>(h_dst), false, offset % kPage, len, err, errlen)) return rc;
      return (int64_t)len;
    }
  return fail(err, errlen, GRIT_EIO, "%s: no such file in the image", relpath);
}

uint64_t grit_image_resident_slots(grit_image* im) {
  if (!im) return 0;
  std::lock_guard<std::mutex> l(im->mu);
  return im->slots_resident;
}

}  // extern "C"
