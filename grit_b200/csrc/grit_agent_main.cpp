// grit-agent, data-mover half: the C++ mirror of cmd/grit-agent for the two actions that call the copy
// package (cmd/grit-agent/app/app.go:53-72, options.go:45-59).  Same flags for the action and the two
// directories, same ACTION environment fallback, same "run grit-agent failed: ..." line on stderr.
//   checkpoint: TransferData(SrcDir, DstDir)                          (checkpoint.go:20; the CRIU half is out of scope)
//   restore:    TransferData(SrcDir, DstDir); CreateSentinelFile(DstDir, "download-state")   (restore.go:16-20)
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>

#include "../../include/gritpack.h"

static const char* kDownloadSentinelFile = "download-state";  // pkg/metadata/metadata.go:9

int main(int argc, char** argv) {
  std::string action = getenv("ACTION") ? getenv("ACTION") : "", src, dst, parent;
  int device = 0;
  for (int i = 1; i < argc; i++) {
    std::string a = argv[i], v;
    const size_t eq = a.find('=');
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
  char err[1024] = "";
  int rc;
  if (action == "checkpoint") {  // options.ActionCheckpoint
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
  printf("data transfer completed: %llu bytes in %llu files, %llu bytes stored, %.3f s (zero %llu, dup %llu, parent %llu pages)\n",
         (unsigned long long)st.bytes_logical, (unsigned long long)st.n_files, (unsigned long long)st.bytes_stored, st.seconds,
         (unsigned long long)st.n_zero, (unsigned long long)st.n_dup, (unsigned long long)st.n_parent);
  return 0;
}
