// Cross-rank page dedup behind the C ABI: one process per GPU, each packs its own rank's image,
// and the only thing exchanged is the per-page (hash64, CRC-32C) table -- 12 bytes per 4 KiB page -- with NCCL over
// NVLink, straight into the device buffers K2 probes.  No host round trip, no torch.
//
// The reference has no counterpart: one agent Job per pod and no cross-pod step
// (docs/proposals/20250221-gpu-workload-checkpointing-and-restoration.md:63-65).
//
// How a ranked pack runs (grit_pack_buffer_ranked):
//   kernel stream   K1(slice 0) K1(slice 1) ... K1(last)  insert own hashes | probe with lower ranks' lists  resolve  K3
//   comm stream          all-gather(slice 0) all-gather(slice 1) ...   all-gather(last) ^
// K1 runs in slices of 2^shift pages; as soon as a slice's hashes exist its all-gather starts on the comm stream, so
// all but the last slice's exchange hides behind K1.  K1 leaves 16 SMs free for NCCL's 16 CTAs while a communicator
// is in use (its grid is one persistent CTA per SM and would otherwise keep NCCL's kernels waiting for a free SM).
// Gathered slices stay where NCCL put them ([slice][rank][2^shift]); K2 reads them through ListView.
//
// NCCL is loaded with dlopen at the first grit_comm_* call, not linked: a single-GPU agent never maps the 400 MB
// library, and inside a process that already holds an NCCL (PyTorch's) the same copy is used.
#include <dlfcn.h>
#include <nccl.h>

#include <algorithm>
#include <cstdlib>
#include <mutex>
#include <new>
#include <vector>

#include "ctx.h"

using namespace grit;

namespace {

struct NcclApi {
  void* handle = nullptr;
  decltype(&ncclGetUniqueId) GetUniqueId = nullptr;
  decltype(&ncclCommInitRankConfig) CommInitRankConfig = nullptr;
  decltype(&ncclCommDestroy) CommDestroy = nullptr;
  decltype(&ncclGetErrorString) GetErrorString = nullptr;
  decltype(&ncclAllGather) AllGather = nullptr;
  decltype(&ncclBroadcast) Broadcast = nullptr;
  decltype(&ncclGroupStart) GroupStart = nullptr;
  decltype(&ncclGroupEnd) GroupEnd = nullptr;
  decltype(&ncclGetVersion) GetVersion = nullptr;
};
NcclApi g_nccl;
std::mutex g_nccl_mu;

int load_nccl(char* err, size_t errlen) {
  std::lock_guard<std::mutex> g(g_nccl_mu);
  if (g_nccl.handle) return GRIT_OK;
  const char* names[] = {getenv("GRITPACK_NCCL_LIB"), "libnccl.so.2", "libnccl.so"};
  void* h = nullptr;
  for (const char* n : names)
    if (n && *n && (h = dlopen(n, RTLD_NOW | RTLD_GLOBAL))) break;
  if (!h) return fail(err, errlen, GRIT_ECUDA, "cannot load NCCL (libnccl.so.2): %s", dlerror());
  NcclApi a;
  a.handle = h;
#define SYM(field, name)                                                                   \
  a.field = reinterpret_cast<decltype(a.field)>(dlsym(h, name));                            \
  if (!a.field) return fail(err, errlen, GRIT_ECUDA, "NCCL library has no symbol %s", name)
  SYM(GetUniqueId, "ncclGetUniqueId");
  SYM(CommInitRankConfig, "ncclCommInitRankConfig");
  SYM(CommDestroy, "ncclCommDestroy");
  SYM(GetErrorString, "ncclGetErrorString");
  SYM(AllGather, "ncclAllGather");
  SYM(Broadcast, "ncclBroadcast");
  SYM(GroupStart, "ncclGroupStart");
  SYM(GroupEnd, "ncclGroupEnd");
  SYM(GetVersion, "ncclGetVersion");
#undef SYM
  g_nccl = a;
  return GRIT_OK;
}

#define GRIT_NCCL(call)                                                                                          \
  do {                                                                                                           \
    ncclResult_t r_ = (call);                                                                                    \
    if (r_ != ncclSuccess)                                                                                       \
      return fail(err, errlen, GRIT_ECUDA, "%s: %s (%s:%d)", #call, g_nccl.GetErrorString(r_), __FILE__, __LINE__); \
  } while (0)

// How much of the GPU the exchange gets.  Measured, slowest rank's pack of the configs[2] image: on 8 x B200 8 NCCL CTAs
// on 8 SMs K1 leaves free with 8 slices 4.98 ms, 16 / 16 / 8 4.69, 32 CTAs on 8 SMs 5.15 (NCCL's CTAs wait for SMs),
// 8 / 8 / 16 5.14, 16 / 16 / 16 4.55; on 2 x B200 8 / 8 / 8 4.42 ms, 16 / 16 / 16 4.68 (the tables are a quarter of the
// size and the SMs are missed by rank 0's write-through K1).  So: 8 up to two ranks, 16 beyond.
int env_int(const char* name, int lo, int hi, int dflt) {
  const char* e = getenv(name);
  const int n = e ? atoi(e) : lo - 1;
  return n >= lo && n <= hi ? n : dflt;
}
int comm_ctas(int world) { return env_int("GRITPACK_COMM_CTAS", 1, 32, world <= 2 ? 8 : 16); }  // NCCL's CTAs
int comm_reserve(int world) { return env_int("GRITPACK_COMM_RESERVE", 0, 64, comm_ctas(world)); }  // SMs K1 leaves free
int comm_slices(int world) { return env_int("GRITPACK_COMM_SLICES", 1, 64, world <= 2 ? 8 : 16); }  // slices per image

}  // namespace

struct grit_comm {
  grit_ctx* ctx = nullptr;
  int rank = 0, world = 1;
  ncclComm_t nccl = nullptr;
  cudaStream_t cs = nullptr;  // the exchange runs here, beside the kernels
  cudaEvent_t ev_done = nullptr;
  std::vector<cudaEvent_t> ev_slice;
  unsigned long long* d_counts = nullptr;  // [world] page counts, device (all-gather target)
  unsigned long long* h_counts = nullptr;  // pinned mirror
  uint64_t exchanges = 0;
};

namespace grit {

grit_ctx* comm_context(grit_comm* c) { return c ? c->ctx : nullptr; }
int comm_rank_of(const grit_comm* c) { return c ? c->rank : -1; }

int comm_gather_tables(grit_comm* c, const uint64_t* d_hash, const uint32_t* d_crc, uint64_t n_pages, grit_table* lower,
                       char* err, size_t errlen) {
  grit_ctx* ctx = c->ctx;
  const int W = c->world;
  cudaStream_t cs = c->cs;
  GRIT_CUDA(cudaSetDevice(ctx->device));
  GRIT_CUDA(cudaMemcpyAsync(c->d_counts + c->rank, &n_pages, sizeof(uint64_t), cudaMemcpyHostToDevice, cs));
  GRIT_NCCL(g_nccl.AllGather(c->d_counts + c->rank, c->d_counts, 1, ncclUint64, c->nccl, cs));
  GRIT_CUDA(cudaMemcpyAsync(c->h_counts, c->d_counts, W * sizeof(uint64_t), cudaMemcpyDeviceToHost, cs));
  GRIT_CUDA(cudaStreamSynchronize(cs));
  uint64_t L = 1;
  for (int r = 0; r < W; r++) L = std::max<uint64_t>(L, c->h_counts[r]);
  if (L > GRIT_MAX_PAGES) return fail(err, errlen, GRIT_EINVAL, "a rank's image exceeds GRIT_MAX_PAGES");
  void *ph = nullptr, *pc = nullptr, *pp = nullptr;
  if (int rc = ctx_dev(ctx, kDevCommHash, (uint64_t)W * L * 8, &ph, err, errlen)) return rc;
  if (int rc = ctx_dev(ctx, kDevCommCrc, (uint64_t)W * L * 4, &pc, err, errlen)) return rc;
  if (int rc = ctx_dev(ctx, kDevCommPad, L * 12, &pp, err, errlen)) return rc;
  uint64_t* pad_hash = static_cast<uint64_t*>(pp);
  uint32_t* pad_crc = reinterpret_cast<uint32_t*>(pad_hash + L);
  GRIT_CUDA(cudaMemsetAsync(pp, 0, L * 12, cs));  // every rank sends L entries: zero padding behind its own (hash64 is never 0)
  if (n_pages) {
    GRIT_CUDA(cudaMemcpyAsync(pad_hash, d_hash, n_pages * 8, cudaMemcpyDeviceToDevice, cs));
    GRIT_CUDA(cudaMemcpyAsync(pad_crc, d_crc, n_pages * 4, cudaMemcpyDeviceToDevice, cs));
  }
  GRIT_NCCL(g_nccl.GroupStart());
  GRIT_NCCL(g_nccl.AllGather(pad_hash, ph, L, ncclUint64, c->nccl, cs));
  GRIT_NCCL(g_nccl.AllGather(pad_crc, pc, L, ncclUint32, c->nccl, cs));
  GRIT_NCCL(g_nccl.GroupEnd());
  GRIT_CUDA(cudaStreamSynchronize(cs));
  for (int r = 0; r < c->rank; r++) {
    lower[r].hash = static_cast<const uint64_t*>(ph) + (uint64_t)r * L;
    lower[r].crc = static_cast<const uint32_t*>(pc) + (uint64_t)r * L;
    lower[r].n_pages = c->h_counts[r];
  }
  c->exchanges++;
  return GRIT_OK;
}

}  // namespace grit

extern "C" {

int grit_comm_new_id(void* id, char* err, size_t errlen) {
  if (!id) return fail(err, errlen, GRIT_EINVAL, "id is null");
  if (int rc = load_nccl(err, errlen)) return rc;
  static_assert(sizeof(ncclUniqueId) == GRIT_COMM_ID_BYTES, "GRIT_COMM_ID_BYTES must be sizeof(ncclUniqueId)");
  ncclUniqueId u;
  GRIT_NCCL(g_nccl.GetUniqueId(&u));
  memcpy(id, &u, sizeof u);
  return GRIT_OK;
}

void grit_comm_destroy(grit_comm* c) {
  if (!c) return;
  if (c->ctx) cudaSetDevice(c->ctx->device);
  if (c->cs) cudaStreamSynchronize(c->cs);
  if (c->nccl) g_nccl.CommDestroy(c->nccl);
  for (auto e : c->ev_slice) cudaEventDestroy(e);
  if (c->ev_done) cudaEventDestroy(c->ev_done);
  if (c->d_counts) cudaFree(c->d_counts);
  if (c->h_counts) cudaFreeHost(c->h_counts);
  if (c->cs) cudaStreamDestroy(c->cs);
  delete c;
}

int grit_comm_create(grit_ctx* ctx, int rank, int world, const void* id, grit_comm** out, char* err, size_t errlen) {
  if (!out) return fail(err, errlen, GRIT_EINVAL, "out is null");
  *out = nullptr;
  if (!ctx || !id) return fail(err, errlen, GRIT_EINVAL, "null context or id");
  if (world < 1 || world > (int)GRIT_MAX_PEERS || rank < 0 || rank >= world)
    return fail(err, errlen, GRIT_EINVAL, "rank %d of %d: a job has 1..%u ranks", rank, world, GRIT_MAX_PEERS);
  if (int rc = load_nccl(err, errlen)) return rc;
  GRIT_CUDA(cudaSetDevice(ctx->device));
  grit_comm* c = new (std::nothrow) grit_comm();
  if (!c) return fail(err, errlen, GRIT_ENOMEM, "out of host memory");
  c->ctx = ctx;
  c->rank = rank;
  c->world = world;
  auto bail = [&](int rc) {
    grit_comm_destroy(c);
    return rc;
  };
  cudaError_t e = cudaStreamCreateWithFlags(&c->cs, cudaStreamNonBlocking);
  if (e == cudaSuccess) e = cudaEventCreateWithFlags(&c->ev_done, cudaEventDisableTiming);
  if (e == cudaSuccess) e = cudaMalloc(&c->d_counts, GRIT_MAX_PEERS * sizeof(unsigned long long));
  if (e == cudaSuccess) e = cudaMallocHost(&c->h_counts, GRIT_MAX_PEERS * sizeof(unsigned long long));
  if (e != cudaSuccess) return bail(fail(err, errlen, GRIT_ECUDA, "communicator set-up: %s", cudaGetErrorString(e)));
  ncclUniqueId u;
  memcpy(&u, id, sizeof u);
  ncclConfig_t cfg = NCCL_CONFIG_INITIALIZER;
  cfg.maxCTAs = comm_ctas(world);  // the table exchange is small; K1 leaves exactly this many SMs free for it
  const ncclResult_t r = g_nccl.CommInitRankConfig(&c->nccl, world, u, rank, &cfg);
  if (r != ncclSuccess) return bail(fail(err, errlen, GRIT_ECUDA, "ncclCommInitRankConfig: %s", g_nccl.GetErrorString(r)));
  *out = c;
  return GRIT_OK;
}

int grit_comm_rank(const grit_comm* c) { return c ? c->rank : -1; }
int grit_comm_world(const grit_comm* c) { return c ? c->world : 0; }

int grit_comm_broadcast(grit_comm* c, void* d_buf, uint64_t bytes, int root, void* stream, char* err, size_t errlen) {
  if (!c || root < 0 || root >= c->world) return fail(err, errlen, GRIT_EINVAL, "bad communicator or root");
  GRIT_CUDA(cudaSetDevice(c->ctx->device));
  if (bytes) GRIT_NCCL(g_nccl.Broadcast(d_buf, d_buf, bytes, ncclUint8, root, c->nccl, static_cast<cudaStream_t>(stream)));
  return GRIT_OK;
}

int grit_pack_buffer_ranked(grit_comm* c, const void* d_image, uint64_t n_pages, uint64_t max_pages,
                            const grit_table* parent, void* d_packed, uint64_t capacity_pages, uint32_t* d_ref,
                            uint32_t* d_crc, uint64_t* d_hash, grit_pack_stats* stats, void* stream, char* err,
                            size_t errlen) {
  if (!c) return fail(err, errlen, GRIT_EINVAL, "null communicator");
  grit_ctx* ctx = c->ctx;
  GRIT_CUDA(cudaSetDevice(ctx->device));
  if (n_pages > GRIT_MAX_PAGES || max_pages > GRIT_MAX_PAGES) return fail(err, errlen, GRIT_EINVAL, "image exceeds GRIT_MAX_PAGES");
  if (max_pages && n_pages > max_pages) return fail(err, errlen, GRIT_EINVAL, "n_pages %llu above max_pages %llu", (unsigned long long)n_pages, (unsigned long long)max_pages);
  if ((reinterpret_cast<uintptr_t>(d_image) | reinterpret_cast<uintptr_t>(d_packed)) & 15)
    return fail(err, errlen, GRIT_EINVAL, "buffer is not 16-byte aligned");
  if (n_pages && (!d_image || !d_ref || !d_crc || !d_hash || (!d_packed && capacity_pages)))
    return fail(err, errlen, GRIT_EINVAL, "null buffer");
  if (parent && parent->n_pages > GRIT_MAX_PAGES) return fail(err, errlen, GRIT_EINVAL, "parent too large");
  std::lock_guard<std::mutex> g(ctx->mu);
  cudaStream_t s = static_cast<cudaStream_t>(stream), cs = c->cs;
  const int W = c->world;
  if (int rc = reset_counters(ctx, s, err, errlen)) return rc;
  ctx->through_image = nullptr;
  GRIT_CUDA(cudaEventRecord(ctx->ev0, s));

  // every rank must issue the same collectives: agree on the longest image first, unless the caller vouches for it
  if (!max_pages) {
    GRIT_CUDA(cudaMemcpyAsync(c->d_counts + c->rank, &n_pages, sizeof(uint64_t), cudaMemcpyHostToDevice, cs));
    GRIT_NCCL(g_nccl.AllGather(c->d_counts + c->rank, c->d_counts, 1, ncclUint64, c->nccl, cs));
    GRIT_CUDA(cudaMemcpyAsync(c->h_counts, c->d_counts, W * sizeof(uint64_t), cudaMemcpyDeviceToHost, cs));
    GRIT_CUDA(cudaStreamSynchronize(cs));
    for (int r = 0; r < W; r++) max_pages = std::max<uint64_t>(max_pages, c->h_counts[r]);
    if (max_pages > GRIT_MAX_PAGES) return fail(err, errlen, GRIT_EINVAL, "a rank's image exceeds GRIT_MAX_PAGES");
  }
  // slices of 2^shift pages: about eight per image, at least 4096 pages each
  uint32_t shift = 12;
  while (shift < 26 && (max_pages >> shift) > (uint64_t)comm_slices(W)) shift++;
  const uint64_t L = 1ull << shift, n_slices = (max_pages + L - 1) >> shift;
  uint64_t *hash_all = nullptr, *pad_hash = nullptr;
  uint32_t *crc_all = nullptr, *pad_crc = nullptr;
  if (W > 1 && n_slices) {
    void* p = nullptr;
    if (int rc = ctx_dev(ctx, kDevCommHash, n_slices * W * L * 8, &p, err, errlen)) return rc;
    hash_all = static_cast<uint64_t*>(p);
    if (int rc = ctx_dev(ctx, kDevCommCrc, n_slices * W * L * 4, &p, err, errlen)) return rc;
    crc_all = static_cast<uint32_t*>(p);
    if (int rc = ctx_dev(ctx, kDevCommPad, L * 12, &p, err, errlen)) return rc;
    pad_hash = static_cast<uint64_t*>(p);
    pad_crc = reinterpret_cast<uint32_t*>(pad_hash + L);
    while (c->ev_slice.size() < n_slices) {
      cudaEvent_t e;
      GRIT_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
      c->ev_slice.push_back(e);
    }
  }
  LaunchGeom k1_geom = ctx->geom;  // leave NCCL its SMs while slices are still in flight
  if (W > 1) k1_geom.sm_count = std::max(1, ctx->geom.sm_count - comm_reserve(W));
  // Without a parent the image may be dense (weights, optimizer state) and K1 can store the pages to their own slots
  // as in grit_pack_buffer.  Rank 0 refers to no other rank: its own sample decides.  A rank above 0 may hold what a
  // lower rank holds (replicated weights: nearly all PEER references, nothing to write through) or not (sharded
  // state): it analyses its first slice plainly, and once that slice's tables have been exchanged a look at the
  // lower ranks' hashes at the same page indices settles it for the slices that follow.
  const bool try_through = !(parent && parent->n_pages) && n_pages && d_packed && write_through_pack() &&
                           (c->rank == 0 || n_slices >= 4);
  const uint64_t through_from = try_through && c->rank > 0 ? L : 0;
  if (try_through) {
    GRIT_CUDA(cudaMemsetAsync(&ctx->d_cnt->first_drop, 0xFF, sizeof(unsigned long long), s));
    GRIT_CUDA(cudaMemcpyAsync(&ctx->d_cnt->through_from, &through_from, sizeof(uint64_t), cudaMemcpyHostToDevice, s));
    GRIT_CUDA(cudaMemsetAsync(&ctx->d_cnt->n_nonkeep, 0, sizeof(unsigned int), s));
    GRIT_CUDA(launch_probe_dense(d_image, n_pages, ctx->d_cnt, s));
    ctx->launches++;
  }

  for (uint64_t i = 0; i < std::max<uint64_t>(n_slices, 1); i++) {
    const uint64_t p0 = i << shift, have = p0 < n_pages ? std::min(L, n_pages - p0) : 0;
    if (have && try_through && p0 >= through_from) {
      if (i == 1 && c->rank > 0) {  // the first slice of every lower rank has arrived: do they hold what this rank holds?
        cudaEvent_t first_done = c->ev_slice[n_slices - 1];  // a spare event: the last slice records it again later
        GRIT_CUDA(cudaEventRecord(first_done, cs));
        GRIT_CUDA(cudaStreamWaitEvent(s, first_done, 0));
        ProbeLists first{};
        for (int r = 0; r < c->rank; r++) {
          first.hash[1 + r].base = reinterpret_cast<const unsigned long long*>(hash_all) + (uint64_t)r * L;
          first.n[1 + r] = L;
        }
        GRIT_CUDA(launch_peer_overlap_hint(reinterpret_cast<const unsigned long long*>(d_hash), first, std::min(L, n_pages),
                                           ctx->d_cnt, s));
        ctx->launches++;
      }
      GRIT_CUDA(launch_analyze_auto(k1_geom, ctx->d_tables, static_cast<const char*>(d_image) + p0 * kPage, have, d_crc + p0,
                                    d_hash + p0, d_ref + p0, static_cast<char*>(d_packed) + p0 * kPage, capacity_pages, p0,
                                    ctx->d_cnt, s));
      ctx->launches += 2;
    } else if (have) {
      GRIT_CUDA(launch_analyze(k1_geom, ctx->d_tables, static_cast<const char*>(d_image) + p0 * kPage, have, d_crc + p0,
                               d_hash + p0, d_ref + p0, s));
      ctx->launches++;
    }
    if (W == 1 || !n_slices) continue;
    if (have) {
      GRIT_CUDA(cudaEventRecord(c->ev_slice[i], s));
      GRIT_CUDA(cudaStreamWaitEvent(cs, c->ev_slice[i], 0));
    }
    const uint64_t* sh = d_hash + p0;
    const uint32_t* sc = d_crc + p0;
    if (have < L) {  // the last slice of a short image, or nothing at all: zero padding (hash64 is never 0)
      GRIT_CUDA(cudaMemsetAsync(pad_hash, 0, L * 12, cs));
      if (have) {
        GRIT_CUDA(cudaMemcpyAsync(pad_hash, d_hash + p0, have * 8, cudaMemcpyDeviceToDevice, cs));
        GRIT_CUDA(cudaMemcpyAsync(pad_crc, d_crc + p0, have * 4, cudaMemcpyDeviceToDevice, cs));
      }
      sh = pad_hash;
      sc = pad_crc;
    }
    GRIT_NCCL(g_nccl.GroupStart());
    GRIT_NCCL(g_nccl.AllGather(sh, hash_all + i * W * L, L, ncclUint64, c->nccl, cs));
    GRIT_NCCL(g_nccl.AllGather(sc, crc_all + i * W * L, L, ncclUint32, c->nccl, cs));
    GRIT_NCCL(g_nccl.GroupEnd());
  }
  if (W > 1 && n_slices) GRIT_CUDA(cudaEventRecord(c->ev_done, cs));
  c->exchanges++;
  if (try_through) {  // K3 on these buffers leaves in place what K1 stored
    ctx->through_image = d_image;
    ctx->through_packed = d_packed;
    ctx->through_pages = n_pages;
  }
  GRIT_CUDA(cudaEventRecord(ctx->evA, s));

  // K2: the parent and the ranks below this one may own a page (lowest origin wins)
  PeerCrcs crcs{};
  ProbeLists lists{};
  if (parent && parent->n_pages) {
    crcs.parent.base = parent->crc;
    lists.hash[0].base = reinterpret_cast<const unsigned long long*>(parent->hash);
    lists.n[0] = parent->n_pages;
  }
  for (int r = 0; r < c->rank && n_slices; r++) {
    crcs.peer[r].base = crc_all + (uint64_t)r * L;
    lists.hash[1 + r].base = reinterpret_cast<const unsigned long long*>(hash_all) + (uint64_t)r * L;
    crcs.peer[r].shift = lists.hash[1 + r].shift = shift;
    crcs.peer[r].stride = lists.hash[1 + r].stride = (uint64_t)W * L;
    lists.n[1 + r] = max_pages;  // padding reads as hash 0 and is skipped
  }
  if (int rc = dedup_views_async(ctx, n_pages, d_crc, d_hash, d_ref, crcs, lists, s, err, errlen,
                                 (W > 1 && n_slices) ? c->ev_done : nullptr))
    return rc;
  GRIT_CUDA(cudaEventRecord(ctx->evB, s));
  if (int rc = compact_async(ctx, d_image, n_pages, d_ref, d_packed, capacity_pages, 0, s, err, errlen)) return rc;
  // rank 0 probes nothing, but its hashes are still being sent: the caller's arrays stay in use until the exchange ends
  if (W > 1 && n_slices) GRIT_CUDA(cudaStreamWaitEvent(s, c->ev_done, 0));
  GRIT_CUDA(cudaEventRecord(ctx->ev1, s));
  if (int rc = read_counters(ctx, s, err, errlen)) return rc;
  const DevCounters& k = *ctx->h_cnt;
  if (stats) {
    stats->n_pages = n_pages;
    stats->n_zero = k.kind[0];
    stats->n_data = k.kind[1];
    stats->n_dup = k.kind[2];
    stats->n_parent = k.kind[3];
    stats->n_peer = k.kind[4];
    float ms = 0;
    cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1);
    stats->gpu_ms = ms;
    cudaEventElapsedTime(&ms, ctx->ev0, ctx->evA);
    stats->stage_ms[0] = ms;
    cudaEventElapsedTime(&ms, ctx->evA, ctx->evB);
    stats->stage_ms[1] = ms;
    cudaEventElapsedTime(&ms, ctx->evB, ctx->ev1);
    stats->stage_ms[2] = ms;
    stats->wrote_through = try_through && k.dense_hint;
  }
  if (k.overflow)
    return fail(err, errlen, GRIT_EINVAL, "packed buffer too small: %llu pages did not fit capacity %llu",
                (unsigned long long)k.overflow, (unsigned long long)capacity_pages);
  return GRIT_OK;
}

}  // extern "C"
