// grit_compact_pages_static: the compaction kernel for callers that must not allocate.
//
// grit_compact_pages() keeps its working memory (one status word per 256-page
// tile for the look-back scan, plus a block of counters) in buffers that a
// grit_ctx obtains from cudaMalloc / cudaMallocHost.  Some callers cannot do
// that.  The packer runs at the moment a training job is frozen for a
// checkpoint, which is exactly when its framework's caching allocator owns
// nearly all of HBM: a cudaMalloc can then fail, or succeed only by making the
// allocator give memory back.  Others run inside a latency-critical section and
// cannot afford the device-wide synchronisation cudaMalloc and cudaFree imply,
// or are handed a fixed set of buffers by a host runtime that forbids any
// further allocation.  For them the working memory lives in the array below,
// inside the library image, and costs kScratchBytes (about 137 KiB) on every
// GPU that has used it, from that GPU's first use until the process exits.
// The price is a fixed upper bound on the image size and one shared scratch
// area per device.
//
// "From first use", not "from load": the CUDA runtime loads device code lazily
// (the default since CUDA 12.2).  The array comes into being on a device when
// its address is first asked for, and the compaction kernel is loaded at its
// first launch; both are the driver's own device-memory allocations and they
// happen inside the first call on that device.  A caller that will run when
// memory is short therefore calls grit_compact_pages_static_warmup(device)
// once at start-up, while memory is plentiful: it does the set-up and runs a
// one-page compaction on buffers that live in the array itself, so that every
// later call finds array and kernel resident and allocates nothing at all.
#include <mutex>

#include "ctx.h"

using namespace grit;

constexpr uint64_t kMaxPages = GRIT_STATIC_MAX_PAGES;
constexpr uint64_t kMaxTiles = (kMaxPages + kCompactTile - 1) / kCompactTile;
constexpr uint64_t kCounterWords = 32;  // 256 bytes reserved for DevCounters, keeps the status words aligned
// behind the status words: the buffers of the warm-up's one-page compaction (image page, packed page, one reference)
constexpr uint64_t kWarmWords = (2 * kPage + 256) / sizeof(unsigned long long);
constexpr uint64_t kScratchWords = kCounterWords + kMaxTiles + kWarmWords;
constexpr uint64_t kScratchBytes = kScratchWords * sizeof(unsigned long long);
static_assert(sizeof(DevCounters) <= kCounterWords * sizeof(unsigned long long), "counter block too small");
static_assert(kScratchBytes <= 160 * 1024, "the scratch every GPU keeps resident must stay modest");

// [0, 32): DevCounters.  [32, 32 + kMaxTiles): tile status words of the look-back scan.  Then the warm-up buffers.
static __device__ __align__(256) unsigned long long g_scratch[kScratchWords];

namespace {

constexpr int kMaxDevices = 64;
struct PerDevice {
  bool ready = false;
  unsigned long long* scratch = nullptr;  // device address of g_scratch on this device
  LaunchGeom geom{};
};
PerDevice g_dev[kMaxDevices];
std::mutex g_setup_mu;

// Once per device, with that device current: ask the runtime where this
// device's copy of g_scratch lives.  In host code `g_scratch` is only a handle
// for that query; its own address is a host address and must never reach a kernel.
int setup_device(int device, PerDevice** out, char* err, size_t errlen) {
  std::lock_guard<std::mutex> g(g_setup_mu);
  PerDevice& d = g_dev[device];
  if (!d.ready) {
    int major = 0, sms = 0;
    GRIT_CUDA(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, device));
    if (major != 10)
      return fail(err, errlen, GRIT_ECUDA, "device %d is not sm_100; libgritpack is built for sm_100a only", device);
    GRIT_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device));
    void* p = nullptr;
    GRIT_CUDA(cudaGetSymbolAddress(&p, g_scratch));
    d.scratch = static_cast<unsigned long long*>(p);
    d.geom.sm_count = sms;
    d.ready = true;
  }
  *out = &d;
  return GRIT_OK;
}

// Makes `device` current for the call and puts the caller's device back on the way out.
struct DeviceScope {
  int prev = -1;
  int enter(int device, char* err, size_t errlen) {
    int n = 0;
    GRIT_CUDA(cudaGetDeviceCount(&n));
    if (device < 0 || device >= n || device >= kMaxDevices)
      return fail(err, errlen, GRIT_EINVAL, "device %d out of range (%d device(s))", device, n);
    GRIT_CUDA(cudaGetDevice(&prev));
    if (prev != device) GRIT_CUDA(cudaSetDevice(device));
    else prev = -1;
    return GRIT_OK;
  }
  ~DeviceScope() {
    if (prev >= 0) cudaSetDevice(prev);
  }
};

}  // namespace

extern "C" int grit_compact_pages_static(int device, const void* d_image, uint64_t n_pages, uint32_t* d_ref,
                                         void* d_packed, uint64_t capacity_pages, uint64_t slot_base,
                                         uint64_t* n_stored, void* stream, char* err, size_t errlen) {
  if (n_pages > kMaxPages)
    return fail(err, errlen, GRIT_EINVAL,
                "grit_compact_pages_static: %llu pages exceed GRIT_STATIC_MAX_PAGES (%llu pages, %llu GiB); "
                "use grit_compact_pages, which sizes its scratch to the image",
                (unsigned long long)n_pages, (unsigned long long)kMaxPages,
                (unsigned long long)(kMaxPages * kPage >> 30));
  if ((reinterpret_cast<uintptr_t>(d_image) | reinterpret_cast<uintptr_t>(d_packed)) & 15)
    return fail(err, errlen, GRIT_EINVAL, "buffer is not 16-byte aligned");
  if (n_pages && (!d_image || !d_ref || (!d_packed && capacity_pages))) return fail(err, errlen, GRIT_EINVAL, "null buffer");
  DeviceScope scope;
  if (int rc = scope.enter(device, err, errlen)) return rc;
  PerDevice* d = nullptr;
  if (int rc = setup_device(device, &d, err, errlen)) return rc;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  DevCounters* cnt = reinterpret_cast<DevCounters*>(d->scratch);
  unsigned long long* status = d->scratch + kCounterWords;

  // The array still holds whatever the previous call left: the counters (tile ticket, totals,
  // overflow) and the status words of the tiles this image has must start from zero.
  const uint64_t tiles = (n_pages + kCompactTile - 1) / kCompactTile;
  GRIT_CUDA(cudaMemsetAsync(cnt, 0, kCounterWords * sizeof(unsigned long long), s));
  if (tiles) GRIT_CUDA(cudaMemsetAsync(status, 0, tiles * sizeof(unsigned long long), s));
  GRIT_CUDA(launch_compact(d->geom, d_image, n_pages, d_ref, d_packed, capacity_pages, slot_base, false, status, cnt, s));
  DevCounters host{};  // plain stack memory: the copy below is staged by the driver, nothing is allocated here
  GRIT_CUDA(cudaMemcpyAsync(&host, cnt, sizeof host, cudaMemcpyDeviceToHost, s));
  GRIT_CUDA(cudaStreamSynchronize(s));
  if (n_stored) *n_stored = host.n_stored;
  if (host.overflow)
    return fail(err, errlen, GRIT_EINVAL, "packed buffer too small: %llu pages did not fit capacity %llu",
                (unsigned long long)host.overflow, (unsigned long long)capacity_pages);
  return GRIT_OK;
}

// Start-up call for grit_compact_pages_static: brings the scratch array and the compaction kernel onto `device`
// (the driver allocates for both the first time they are used) by compacting one non-zero page that lives in the
// array itself.  Nothing is allocated by this library here either; what is allocated is the driver's, and only once.
extern "C" int grit_compact_pages_static_warmup(int device, char* err, size_t errlen) {
  DeviceScope scope;
  if (int rc = scope.enter(device, err, errlen)) return rc;
  PerDevice* d = nullptr;
  if (int rc = setup_device(device, &d, err, errlen)) return rc;
  unsigned char* warm = reinterpret_cast<unsigned char*>(d->scratch + kCounterWords + kMaxTiles);
  uint32_t* ref = reinterpret_cast<uint32_t*>(warm + 2 * kPage);
  const uint32_t keep = kRefKeep;
  GRIT_CUDA(cudaMemsetAsync(warm, 0x5A, kPage, nullptr));
  GRIT_CUDA(cudaMemcpyAsync(ref, &keep, sizeof keep, cudaMemcpyHostToDevice, nullptr));
  uint64_t stored = 0;
  if (int rc = grit_compact_pages_static(device, warm, 1, ref, warm + kPage, 1, 0, &stored, nullptr, err, errlen)) return rc;
  if (stored != 1) return fail(err, errlen, GRIT_ECUDA, "warm-up compaction stored %llu pages, expected 1", (unsigned long long)stored);
  return GRIT_OK;
}
