// How fast can a tmpfs file mapping be pinned (cudaHostRegister) and copied to/from the GPU in place?
// Decides whether the directory calls can move file pages without a CPU copy.  Development aid.
//   nvcc -O2 -o /tmp/register_probe tools/register_probe.cu && /tmp/register_probe /dev/shm/probe.bin 4096 64 4
#include <cuda_runtime.h>
#include <fcntl.h>
#include <sys/mman.h>
#include <unistd.h>

#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <thread>
#include <vector>

static double now() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }

int main(int argc, char** argv) {
  const char* path = argc > 1 ? argv[1] : "/dev/shm/probe.bin";
  const size_t total = (size_t)(argc > 2 ? atoi(argv[2]) : 4096) << 20, piece = (size_t)(argc > 3 ? atoi(argv[3]) : 64) << 20;
  const int threads = argc > 4 ? atoi(argv[4]) : 4;
  cudaFree(0);
  int fd = open(path, O_RDWR | O_CREAT | O_TRUNC, 0600);
  if (fd < 0 || posix_fallocate(fd, 0, total)) return perror("file"), 1;
  char* m = (char*)mmap(nullptr, total, PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
  if (m == MAP_FAILED) return perror("mmap"), 1;
  double t0 = now();
  for (size_t o = 0; o < total; o += 4096) m[o] = 1;
  printf("touch mapping: %.2f GB/s\n", total / (now() - t0) / 1e9);
  void* d;
  cudaMalloc(&d, piece * 4);
  for (int mode = 0; mode < 2; mode++) {  // 0: one thread, 1: `threads` threads
    const int nt = mode ? threads : 1;
    t0 = now();
    std::vector<std::thread> th;
    for (int t = 0; t < nt; t++)
      th.emplace_back([&, t] {
        for (size_t o = (size_t)t * piece; o < total; o += (size_t)nt * piece)
          if (cudaHostRegister(m + o, piece, cudaHostRegisterDefault) != cudaSuccess) printf("register failed at %zu\n", o);
      });
    for (auto& x : th) x.join();
    double t1 = now();
    printf("cudaHostRegister %d thread(s), %zu MiB pieces: %.2f GB/s\n", nt, piece >> 20, total / (t1 - t0) / 1e9);
    cudaStream_t s;
    cudaStreamCreate(&s);
    t0 = now();
    for (size_t o = 0; o < total; o += piece) cudaMemcpyAsync(m + o, (char*)d, piece, cudaMemcpyDeviceToHost, s);
    cudaStreamSynchronize(s);
    printf("  D2H into the mapping: %.2f GB/s\n", total / (now() - t0) / 1e9);
    t0 = now();
    for (size_t o = 0; o < total; o += piece) cudaMemcpyAsync((char*)d, m + o, piece, cudaMemcpyHostToDevice, s);
    cudaStreamSynchronize(s);
    printf("  H2D from the mapping: %.2f GB/s\n", total / (now() - t0) / 1e9);
    t0 = now();
    th.clear();
    for (int t = 0; t < nt; t++)
      th.emplace_back([&, t] {
        for (size_t o = (size_t)t * piece; o < total; o += (size_t)nt * piece) cudaHostUnregister(m + o);
      });
    for (auto& x : th) x.join();
    printf("  cudaHostUnregister: %.2f GB/s\n", total / (now() - t0) / 1e9);
  }
  // read-only registration of a mapping opened read-only (the upload side)
  munmap(m, total);
  const char* r = (const char*)mmap(nullptr, total, PROT_READ, MAP_SHARED, fd, 0);
  t0 = now();
  int ok = 0;
  for (size_t o = 0; o < total; o += piece)
    ok += cudaHostRegister((void*)(r + o), piece, cudaHostRegisterReadOnly) == cudaSuccess;
  printf("read-only mapping, cudaHostRegisterReadOnly: %d of %zu pieces ok, %.2f GB/s\n", ok, total / piece, total / (now() - t0) / 1e9);
  printf("last error: %s\n", cudaGetErrorString(cudaGetLastError()));
  unlink(path);
  return 0;
}
