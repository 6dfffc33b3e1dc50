// grit-agent, data-mover half: the C++ mirror of cmd/grit-agent for the two actions that call the copy
// package (cmd/grit-agent/app/app.go:53-72, options.go:45-59).  Same flags for the action and the two
// directories, same ACTION environment fallback, same "run grit-agent failed: ..." line on stderr.
//   checkpoint: TransferData(SrcDir, DstDir)                          (checkpoint.go:20; the CRIU half is out of scope)
//   restore:    TransferData(SrcDir, DstDir); CreateSentinelFile(DstDir, "download-state")   (restore.go:16-20)
//
// Beside the reference's two actions, three that the packed format makes necessary (no counterpart in
// the reference, which neither verifies nor links nor deletes checkpoints):
//   verify --src-dir <packed dir>               checks a checkpoint on the PV without restoring it (grit_verify_dir)
//   deps   --src-dir <packed dir>               lists the packed directories it needs (parents, lower ranks)
//   gc     --src-dir <PV root> --keep <dir>...  deletes every packed directory under the root that no kept
//                                               checkpoint needs; --dry-run only lists them
// --stats-file <path>: the grit_stats of the call as one JSON object, for the Job to surface (Checkpoint.Status
// conditions, Prometheus) -- the reference logs a line per file and emits no throughput figure (copy.go:22,50,61).
#include <dirent.h>
#include <sys/stat.h>
#include <unistd.h>

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

#include "../../include/gritpack.h"

static const char* kDownloadSentinelFile = "download-state";  // pkg/metadata/metadata.go:9

static bool is_packed_dir(const std::string& d) {
  struct stat st;
  return stat((d + "/gritpack.manifest").c_str(), &st) == 0 && S_ISREG(st.st_mode);
}

// every packed directory below root (a packed directory holds no other)
static void find_packed(const std::string& root, int depth, std::vector<std::string>* out) {
  if (is_packed_dir(root)) return out->push_back(root);
  if (depth > 6) return;
  DIR* d = opendir(root.c_str());
  if (!d) return;
  std::vector<std::string> names;
  while (struct dirent* e = readdir(d))
    if (strcmp(e->d_name, ".") && strcmp(e->d_name, "..")) names.emplace_back(e->d_name);
  closedir(d);
  std::sort(names.begin(), names.end());
  for (const auto& n : names) {
    struct stat st;
    const std::string p = root + "/" + n;
    if (lstat(p.c_str(), &st) == 0 && S_ISDIR(st.st_mode)) find_packed(p, depth + 1, out);
  }
}

static int deps_of(const std::string& dir, std::vector<std::string>* out, char* err, size_t errlen) {
  size_t need = 0;
  grit_image_deps(dir.c_str(), 1, nullptr, 0, &need, err, errlen);
  if (!need) return GRIT_EFORMAT;
  std::string buf(need, '\0');
  if (int rc = grit_image_deps(dir.c_str(), 1, &buf[0], need, &need, err, errlen)) return rc;
  for (const char* p = buf.c_str(); *p; p += strlen(p) + 1) out->emplace_back(p);
  return GRIT_OK;
}

static std::string real(const std::string& p) {
  char b[4096];
  return realpath(p.c_str(), b) ? std::string(b) : p;
}

static int remove_tree(const std::string& dir) {  // a packed directory is flat: files only
  DIR* d = opendir(dir.c_str());
  if (!d) return -1;
  int rc = 0;
  while (struct dirent* e = readdir(d))
    if (strcmp(e->d_name, ".") && strcmp(e->d_name, "..")) rc |= unlink((dir + "/" + e->d_name).c_str());
  closedir(d);
  return rc | rmdir(dir.c_str());
}

int main(int argc, char** argv) {
  std::string action = getenv("ACTION") ? getenv("ACTION") : "", src, dst, parent, stats_file;
  std::vector<std::string> keep;
  bool dry_run = false;
  int device = 0;
  for (int i = 1; i < argc; i++) {
    std::string a = argv[i], v;
    const size_t eq = a.find('=');
    if (a == "--dry-run") {
      dry_run = true;
      continue;
    }
    if (eq != std::string::npos) {
      v = a.substr(eq + 1);
      a = a.substr(0, eq);
    } else if (i + 1 < argc) {
      v = argv[++i];
    }
    if (a == "--action") action = v;
    else if (a == "--src-dir") src = v;
    else if (a == "--dst-dir") dst = v;
    else if (a == "--parent-dir") parent = v;
    else if (a == "--device") device = atoi(v.c_str());
    else if (a == "--keep") keep.push_back(v);
    else if (a == "--stats-file") stats_file = v;
    else {
      fprintf(stderr, "unknown flag %s\n", a.c_str());
      return 2;
    }
  }
  grit_opts o;
  memset(&o, 0, sizeof o);
  o.struct_size = sizeof o;
  o.device = device;
  o.verify = 1;
  o.parent_dir = parent.empty() ? nullptr : parent.c_str();
  grit_stats st;
  memset(&st, 0, sizeof st);
  char err[1024] = "";
  int rc;
  // the first CUDA call of a process costs 0.3-0.5 s: start it now, beside the host-side set-up of the action
  std::thread warm;
  if (action == "checkpoint" || action == "restore" || action == "verify") warm = std::thread([device] { grit_warmup(device); });
  struct Join {
    std::thread& t;
    ~Join() {
      if (t.joinable()) t.join();
    }
  } join{warm};
  if (action == "deps") {
    std::vector<std::string> d;
    if ((rc = deps_of(src, &d, err, sizeof err)) == 0)
      for (const auto& x : d) printf("%s\n", x.c_str());
    if (rc == 0) return 0;
  } else if (action == "gc") {
    std::vector<std::string> all, needed;
    find_packed(src, 0, &all);
    rc = 0;
    for (const auto& k : keep) {
      needed.push_back(real(k));
      std::vector<std::string> d;
      if ((rc = deps_of(k, &d, err, sizeof err))) break;  // a kept checkpoint that cannot be read: delete nothing
      for (const auto& x : d) needed.push_back(real(x));
    }
    if (rc == 0 && keep.empty()) {
      rc = GRIT_EINVAL;
      snprintf(err, sizeof err, "gc needs at least one --keep <packed dir>");
    }
    if (rc == 0) {
      unsigned removed = 0;
      for (const auto& d : all)
        if (std::find(needed.begin(), needed.end(), real(d)) == needed.end()) {
          printf("%s %s\n", dry_run ? "would remove" : "remove", d.c_str());
          if (!dry_run && remove_tree(d) != 0) {
            rc = GRIT_EIO;
            snprintf(err, sizeof err, "cannot remove %s", d.c_str());
            break;
          }
          removed++;
        }
      if (rc == 0) {
        printf("gc: %zu packed directories under %s, %zu needed, %u %s\n", all.size(), src.c_str(), needed.size(), removed,
               dry_run ? "removable" : "removed");
        return 0;
      }
    }
  } else if (action == "verify") {
    rc = grit_verify_dir(src.c_str(), &o, &st, err, sizeof err);
  } else if (action == "checkpoint") {  // options.ActionCheckpoint
    rc = grit_pack_dir(src.c_str(), dst.c_str(), &o, &st, err, sizeof err);
  } else if (action == "restore") {  // options.ActionRestore
    rc = grit_restore_dir(src.c_str(), dst.c_str(), &o, &st, err, sizeof err);
    if (rc == 0 && grit_write_sentinel(dst.c_str(), kDownloadSentinelFile) != 0) {
      rc = GRIT_EIO;
      snprintf(err, sizeof err, "create sentinel file %s/%s failed", dst.c_str(), kDownloadSentinelFile);
    }
  } else {
    rc = GRIT_EINVAL;
    snprintf(err, sizeof err, "unknown action %s", action.c_str());  // app.go:68
  }
  if (rc != 0) {
    fprintf(stderr, "run grit-agent failed: %s\n", err);  // app.go:40
    return 1;
  }
  if (!stats_file.empty()) {
    if (FILE* f = fopen(stats_file.c_str(), "w")) {
      fprintf(f,
              "{\"action\": \"%s\", \"bytes_logical\": %llu, \"bytes_stored\": %llu, \"n_files\": %llu, \"n_pages\": %llu, "
              "\"n_zero\": %llu, \"n_data\": %llu, \"n_dup\": %llu, \"n_parent\": %llu, \"n_peer\": %llu, \"n_bad\": %llu, "
              "\"seconds\": %.6f, \"gpu_ms\": %.3f, \"gb_per_s\": %.3f}\n",
              action.c_str(), (unsigned long long)st.bytes_logical, (unsigned long long)st.bytes_stored,
              (unsigned long long)st.n_files, (unsigned long long)st.n_pages, (unsigned long long)st.n_zero,
              (unsigned long long)st.n_data, (unsigned long long)st.n_dup, (unsigned long long)st.n_parent,
              (unsigned long long)st.n_peer, (unsigned long long)st.n_bad, st.seconds, st.gpu_ms,
              st.seconds > 0 ? st.bytes_logical / st.seconds / 1e9 : 0.0);
      fclose(f);
    }
  }
  printf("data transfer completed: %llu bytes in %llu files, %llu bytes stored, %.3f s (zero %llu, dup %llu, parent %llu pages)\n",
         (unsigned long long)st.bytes_logical, (unsigned long long)st.n_files, (unsigned long long)st.bytes_stored, st.seconds,
         (unsigned long long)st.n_zero, (unsigned long long)st.n_dup, (unsigned long long)st.n_parent);
  return 0;
}
