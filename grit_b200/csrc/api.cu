// C ABI over the page kernels: context, stage entry points, pack/restore of a
// device-resident image.  Declarations and reference citations: include/gritpack.h.
#include <algorithm>
#include <cstddef>
#include <cstdlib>
#include <new>

#include "ctx.h"

using namespace grit;

namespace grit {

int ensure_table(grit_ctx* ctx, uint64_t n_items, DedupTable* t, char* err, size_t errlen) {
  uint64_t cap = 1024;
  while (cap < 2 * n_items) cap <<= 1;
  if (cap > ctx->table_cap) {
    if (ctx->d_slots) cudaFree(ctx->d_slots);
    ctx->d_slots = nullptr;
    ctx->table_cap = 0;
    GRIT_CUDA(cudaMalloc(&ctx->d_slots, cap * sizeof(TableSlot)));
    ctx->table_cap = cap;
  }
  t->slots = ctx->d_slots;
  t->mask = cap - 1;
  return GRIT_OK;
}

int ensure_status(grit_ctx* ctx, uint64_t n_pages, char* err, size_t errlen) {
  const uint64_t tile = std::min<uint64_t>(kCompactTile, onepass_tile_pages());
  const uint64_t tiles = (n_pages + tile - 1) / tile + 1;
  if (tiles > ctx->status_cap) {
    if (ctx->d_status) cudaFree(ctx->d_status);
    ctx->d_status = nullptr;
    ctx->status_cap = 0;
    GRIT_CUDA(cudaMalloc(&ctx->d_status, tiles * sizeof(unsigned long long)));
    ctx->status_cap = tiles;
  }
  return GRIT_OK;
}

int reset_counters(grit_ctx* ctx, cudaStream_t s, char* err, size_t errlen) {
  GRIT_CUDA(cudaMemsetAsync(ctx->d_cnt, 0, offsetof(DevCounters, first_drop), s));  // not the write-through state
  GRIT_CUDA(cudaMemsetAsync(&ctx->d_cnt->first_bad, 0xFF, sizeof(unsigned long long), s));
  return GRIT_OK;
}

int read_counters(grit_ctx* ctx, cudaStream_t s, char* err, size_t errlen) {
  GRIT_CUDA(cudaMemcpyAsync(ctx->h_cnt, ctx->d_cnt, sizeof(DevCounters), cudaMemcpyDeviceToHost, s));
  GRIT_CUDA(cudaStreamSynchronize(s));
  return GRIT_OK;
}

// GRITPACK_WRITE_THROUGH=0 turns the write-through K1 off (A/B measurements).
bool write_through_pack() {
  static const bool v = [] {
    const char* e = getenv("GRITPACK_WRITE_THROUGH");
    return !(e && *e == '0');
  }();
  return v;
}

int ctx_dev(grit_ctx* ctx, int id, size_t bytes, void** out, char* err, size_t errlen) {
  grit_ctx::Buf& b = ctx->dev[id];
  if (bytes > b.cap) {
    if (b.p) cudaFree(b.p);
    b.p = nullptr;
    b.cap = 0;
    GRIT_CUDA(cudaMalloc(&b.p, bytes));
    b.cap = bytes;
  }
  *out = b.p;
  return GRIT_OK;
}

int ctx_pin(grit_ctx* ctx, int id, size_t bytes, void** out, char* err, size_t errlen) {
  grit_ctx::Buf& b = ctx->pin[id];
  if (bytes > b.cap) {
    if (b.p) cudaFreeHost(b.p);
    b.p = nullptr;
    b.cap = 0;
    GRIT_CUDA(cudaMallocHost(&b.p, bytes));
    b.cap = bytes;
  }
  *out = b.p;
  return GRIT_OK;
}

// GRITPACK_ONEPASS=1 makes grit_pack_buffer pack every image the sample does not find dense with the single-pass
// kernel (k_pack_onepass: pages parked in TMEM, N + U of DRAM traffic).  It is bit-exact but on B200 still slower
// than K1, K2, K3 as separate kernels (N + 2U): the verdict of a tile needs every earlier tile's hashes, and with two
// tiles of TMEM per SM that chain sits on the critical path (profiles/README.md).  Opt-in until it wins.
bool onepass_pack() {
  static const bool v = [] {
    const char* e = getenv("GRITPACK_ONEPASS");
    return e && *e && *e != '0';
  }();
  return v;
}

// The staged dedup: this image's hashes enter a table sized by this image alone, the other images' lists (parent,
// lower ranks) only probe it -- a hit lowers the owner, a miss touches nothing -- then every page reads its owner.
int dedup_views_async(grit_ctx* ctx, uint64_t n_pages, const uint32_t* d_crc, const uint64_t* d_hash, uint32_t* d_ref,
                      const PeerCrcs& crcs, const ProbeLists& lists, cudaStream_t s, char* err, size_t errlen,
                      cudaEvent_t lists_ready) {
  DedupTable t{};
  // K2 counts the references it makes (K3 picks its tiles by it): from zero for every image, also for callers of the
  // stage entry points, whose K1 does not reset it
  GRIT_CUDA(cudaMemsetAsync(&ctx->d_cnt->n_nonkeep, 0, sizeof(unsigned int), s));
  bool peers_on = false;
  for (int o = 1; o < 9; o++) peers_on = peers_on || (lists.n[o] && lists.hash[o]);
  if (!peers_on && lists.n[0] && lists.hash[0] && lists.hash[0].stride == 0 && crcs.parent) {
    // An incremental pack: most pages are the parent's.  One table is built from the parent, and a page whose hash
    // is found there is settled after one read; the rest (the dirtied pages) meet in a second table.  Both are sized
    // for the worst case, so an unrelated parent costs time and never a crowded table.
    uint64_t cap_a = 1024, cap_b = 1024;
    while (cap_a < 2 * lists.n[0]) cap_a <<= 1;
    while (cap_b < 2 * n_pages) cap_b <<= 1;
    if (int rc = ensure_table(ctx, (cap_a + cap_b) / 2, &t, err, errlen)) return rc;
    const DedupTable ta{t.slots, cap_a - 1}, tb{t.slots + cap_a, cap_b - 1};
    GRIT_CUDA(launch_table_clear(ctx->geom, DedupTable{t.slots, cap_a + cap_b - 1}, s));
    if (lists_ready) GRIT_CUDA(cudaStreamWaitEvent(s, lists_ready, 0));
    GRIT_CUDA(launch_table_insert(ctx->geom, ta, reinterpret_cast<const uint64_t*>(lists.hash[0].base), nullptr, lists.n[0], 0,
                                  0, s));
    GRIT_CUDA(launch_resolve_parent(ctx->geom, ta, n_pages, d_crc, d_hash, d_ref, crcs.parent, ctx->d_cnt, s));
    GRIT_CUDA(launch_table_insert(ctx->geom, tb, d_hash, d_ref, n_pages, kOriginSelf, 0, s, true));
    GRIT_CUDA(launch_resolve(ctx->geom, tb, 0, n_pages, d_crc, d_hash, d_ref, crcs, ctx->d_cnt, s, true));
    ctx->launches += 5;
    return GRIT_OK;
  }
  if (int rc = ensure_table(ctx, n_pages, &t, err, errlen)) return rc;
  GRIT_CUDA(launch_table_clear(ctx->geom, t, s));
  ctx->launches++;
  GRIT_CUDA(launch_table_insert(ctx->geom, t, d_hash, d_ref, n_pages, kOriginSelf, 0, s));
  ctx->launches++;
  bool any = false;
  for (int o = 0; o < 9; o++) any = any || (lists.n[o] && lists.hash[o]);
  if (any) {
    if (lists_ready) GRIT_CUDA(cudaStreamWaitEvent(s, lists_ready, 0));  // the lists may still be arriving
    GRIT_CUDA(launch_table_probe_lists(ctx->geom, t, lists, s));
    ctx->launches++;
  }
  GRIT_CUDA(launch_resolve(ctx->geom, t, 0, n_pages, d_crc, d_hash, d_ref, crcs, ctx->d_cnt, s));
  ctx->launches++;
  return GRIT_OK;
}

}  // namespace grit

namespace {

// `stream` is the caller's cudaStream_t; NULL is the CUDA default stream, as everywhere in the runtime API,
// so work is ordered after whatever the caller enqueued there (the inputs' producers).
cudaStream_t pick(grit_ctx*, void* stream) { return static_cast<cudaStream_t>(stream); }

int enter(grit_ctx* ctx, char* err, size_t errlen) {
  if (!ctx) return fail(err, errlen, GRIT_EINVAL, "null context");
  GRIT_CUDA(cudaSetDevice(ctx->device));
  return GRIT_OK;
}

int check_pages(uint64_t n_pages, const void* p, char* err, size_t errlen) {
  if (n_pages > GRIT_MAX_PAGES) return fail(err, errlen, GRIT_EINVAL, "n_pages %llu exceeds GRIT_MAX_PAGES", (unsigned long long)n_pages);
  if (reinterpret_cast<uintptr_t>(p) & 15) return fail(err, errlen, GRIT_EINVAL, "buffer is not 16-byte aligned");
  return GRIT_OK;
}

// table build + resolve, asynchronous on s
int dedup_async(grit_ctx* ctx, uint64_t n_pages, const uint32_t* d_crc, const uint64_t* d_hash, uint32_t* d_ref,
                const grit_table* parent, const grit_table* peers, uint32_t n_peers, cudaStream_t s, char* err,
                size_t errlen, DedupTable* table_out = nullptr, PeerCrcs* crcs_out = nullptr) {
  if (n_peers > GRIT_MAX_PEERS) return fail(err, errlen, GRIT_EINVAL, "n_peers %u > %u", n_peers, GRIT_MAX_PEERS);
  PeerCrcs crcs{};
  ProbeLists lists{};
  uint64_t items = n_pages;
  const bool parent_on = parent && parent->n_pages;
  if (parent_on) {
    if (parent->n_pages > GRIT_MAX_PAGES) return fail(err, errlen, GRIT_EINVAL, "parent too large");
    crcs.parent.base = parent->crc;
    lists.hash[0].base = reinterpret_cast<const unsigned long long*>(parent->hash);
    lists.n[0] = parent->n_pages;
  }
  for (uint32_t r = 0; r < n_peers; r++) {
    if (peers[r].n_pages > GRIT_MAX_PAGES) return fail(err, errlen, GRIT_EINVAL, "peer too large");
    crcs.peer[r].base = peers[r].crc;
    lists.hash[1 + r].base = reinterpret_cast<const unsigned long long*>(peers[r].hash);
    lists.n[1 + r] = peers[r].n_pages;
  }
  if (!table_out) return dedup_views_async(ctx, n_pages, d_crc, d_hash, d_ref, crcs, lists, s, err, errlen);
  // The fused kernel inserts this image's pages itself, after the others: its table must hold them all.
  if (parent_on) items += parent->n_pages;
  for (uint32_t r = 0; r < n_peers; r++) items += peers[r].n_pages;
  DedupTable t{};
  if (int rc = ensure_table(ctx, items, &t, err, errlen)) return rc;
  GRIT_CUDA(launch_table_clear(ctx->geom, t, s));
  ctx->launches++;
  if (parent_on) {
    GRIT_CUDA(launch_table_insert(ctx->geom, t, parent->hash, nullptr, parent->n_pages, 0, 0, s));
    ctx->launches++;
  }
  for (uint32_t r = 0; r < n_peers; r++)
    if (peers[r].n_pages) {
      GRIT_CUDA(launch_table_insert(ctx->geom, t, peers[r].hash, nullptr, peers[r].n_pages, 1 + r, 0, s));
      ctx->launches++;
    }
  *table_out = t;
  *crcs_out = crcs;
  return GRIT_OK;
}

// K1 with the dense sample and write-through into d_packed; remembers the buffers for the K3 that follows
int analyze_through_async(grit_ctx* ctx, const void* d_image, uint64_t n_pages, uint32_t* d_crc, uint64_t* d_hash,
                          uint32_t* d_ref, void* d_packed, uint64_t capacity_pages, cudaStream_t s, char* err,
                          size_t errlen, bool probed = false) {
  GRIT_CUDA(cudaMemsetAsync(&ctx->d_cnt->first_drop, 0xFF, sizeof(unsigned long long), s));
  GRIT_CUDA(cudaMemsetAsync(&ctx->d_cnt->through_from, 0, sizeof(unsigned long long), s));
  GRIT_CUDA(cudaMemsetAsync(&ctx->d_cnt->n_nonkeep, 0, sizeof(unsigned int), s));
  if (!probed) GRIT_CUDA(launch_probe_dense(d_image, n_pages, ctx->d_cnt, s));  // sets dense_hint
  GRIT_CUDA(launch_analyze_auto(ctx->geom, ctx->d_tables, d_image, n_pages, d_crc, d_hash, d_ref, d_packed, capacity_pages, 0,
                                ctx->d_cnt, s));
  ctx->launches += 3;  // probe + the two K1 variants, one of which returns at once
  ctx->through_image = d_image;
  ctx->through_packed = d_packed;
  ctx->through_pages = n_pages;
  return GRIT_OK;
}

}  // namespace

namespace grit {
int compact_async(grit_ctx* ctx, const void* d_image, uint64_t n_pages, uint32_t* d_ref, void* d_packed,
                  uint64_t capacity_pages, uint64_t slot_base, cudaStream_t s, char* err, size_t errlen) {
  // pages that the write-through K1 of these very buffers put at their own index stay where they are
  const bool in_place = n_pages && slot_base == 0 && ctx->through_image == d_image && ctx->through_packed == d_packed &&
                        ctx->through_pages == n_pages;
  ctx->through_image = ctx->through_packed = nullptr;
  ctx->through_pages = 0;
  if (int rc = ensure_status(ctx, n_pages, err, errlen)) return rc;
  const uint64_t tiles = (n_pages + kCompactTile - 1) / kCompactTile;
  GRIT_CUDA(cudaMemsetAsync(ctx->d_status, 0, tiles * sizeof(unsigned long long), s));
  GRIT_CUDA(launch_compact(ctx->geom, d_image, n_pages, d_ref, d_packed, capacity_pages, slot_base, false, ctx->d_status,
                           ctx->d_cnt, s, in_place));
  ctx->launches += 2;  // k_compact and k_compact_wide, one of which returns at once
  return GRIT_OK;
}
}  // namespace grit

extern "C" {

int grit_version(void) { return GRIT_ABI_VERSION; }

int grit_ctx_create(int device, grit_ctx** out, char* err, size_t errlen) {
  if (!out) return fail(err, errlen, GRIT_EINVAL, "out is null");
  *out = nullptr;
  int n_dev = 0;
  cudaError_t e = cudaGetDeviceCount(&n_dev);
  if (e != cudaSuccess || n_dev == 0)
    return fail(err, errlen, GRIT_ECUDA, "no CUDA device (%s); gritpack has no CPU path",
                e == cudaSuccess ? "device count 0" : cudaGetErrorString(e));
  if (device < 0 || device >= n_dev) return fail(err, errlen, GRIT_EINVAL, "device %d out of range (0..%d)", device, n_dev - 1);
  GRIT_CUDA(cudaSetDevice(device));
  cudaDeviceProp prop{};
  GRIT_CUDA(cudaGetDeviceProperties(&prop, device));
  if (prop.major != 10)
    return fail(err, errlen, GRIT_ECUDA, "device %d is sm_%d%d; libgritpack is built for sm_100a only", device, prop.major,
                prop.minor);
  grit_ctx* ctx = new (std::nothrow) grit_ctx();
  if (!ctx) return fail(err, errlen, GRIT_ENOMEM, "out of host memory");
  ctx->device = device;
  ctx->geom.sm_count = prop.multiProcessorCount;
  auto bail = [&](int rc) {
    grit_ctx_destroy(ctx);
    return rc;
  };
#define CTX_CUDA(call)                                                                                  \
  do {                                                                                                  \
    cudaError_t e_ = (call);                                                                            \
    if (e_ != cudaSuccess) return bail(fail(err, errlen, GRIT_ECUDA, "%s: %s", #call, cudaGetErrorString(e_))); \
  } while (0)
  CTX_CUDA(cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking));
  CTX_CUDA(cudaStreamCreateWithFlags(&ctx->h2d, cudaStreamNonBlocking));
  CTX_CUDA(cudaStreamCreateWithFlags(&ctx->d2h, cudaStreamNonBlocking));
  CTX_CUDA(cudaEventCreate(&ctx->ev0));
  CTX_CUDA(cudaEventCreate(&ctx->ev1));
  CTX_CUDA(cudaEventCreate(&ctx->evA));
  CTX_CUDA(cudaEventCreate(&ctx->evB));
  CTX_CUDA(cudaMalloc(&ctx->d_tables, sizeof(Gf2Tables)));
  CTX_CUDA(cudaMalloc(&ctx->d_cnt, sizeof(DevCounters)));
  CTX_CUDA(cudaMemset(ctx->d_cnt, 0, sizeof(DevCounters)));
  CTX_CUDA(cudaMallocHost(&ctx->h_cnt, sizeof(DevCounters)));
  {
    Gf2Tables* h = new (std::nothrow) Gf2Tables();
    if (!h) return bail(fail(err, errlen, GRIT_ENOMEM, "out of host memory"));
    build_gf2_tables(h);
    cudaError_t ce = cudaMemcpy(ctx->d_tables, h, sizeof(Gf2Tables), cudaMemcpyHostToDevice);
    delete h;
    CTX_CUDA(ce);
  }
  CTX_CUDA(configure_kernels());
#undef CTX_CUDA
  *out = ctx;
  return GRIT_OK;
}

void grit_ctx_destroy(grit_ctx* ctx) {
  if (!ctx) return;
  cudaSetDevice(ctx->device);
  cudaDeviceSynchronize();
  for (auto& b : ctx->dev)
    if (b.p) cudaFree(b.p);
  for (auto& b : ctx->pin)
    if (b.p) cudaFreeHost(b.p);
  if (ctx->d_slots) cudaFree(ctx->d_slots);
  if (ctx->d_status) cudaFree(ctx->d_status);
  if (ctx->d_tables) cudaFree(ctx->d_tables);
  if (ctx->d_cnt) cudaFree(ctx->d_cnt);
  if (ctx->h_cnt) cudaFreeHost(ctx->h_cnt);
  if (ctx->ev0) cudaEventDestroy(ctx->ev0);
  if (ctx->ev1) cudaEventDestroy(ctx->ev1);
  if (ctx->evA) cudaEventDestroy(ctx->evA);
  if (ctx->evB) cudaEventDestroy(ctx->evB);
  if (ctx->stream) cudaStreamDestroy(ctx->stream);
  if (ctx->h2d) cudaStreamDestroy(ctx->h2d);
  if (ctx->d2h) cudaStreamDestroy(ctx->d2h);
  delete ctx;
}

int grit_warmup(int device) {
  int n_dev = 0;
  if (cudaGetDeviceCount(&n_dev) != cudaSuccess || device < 0 || device >= n_dev) return GRIT_ECUDA;
  if (cudaSetDevice(device) != cudaSuccess || cudaFree(nullptr) != cudaSuccess) return GRIT_ECUDA;  // primary context
  return configure_kernels() == cudaSuccess ? GRIT_OK : GRIT_ECUDA;                                   // loads the module
}

uint64_t grit_ctx_launches(const grit_ctx* ctx) { return ctx ? ctx->launches : 0; }

int grit_ctx_set_option(grit_ctx* ctx, const char* name, uint64_t value) {
  if (!ctx || !name) return GRIT_EINVAL;
  if (!strcmp(name, "verify_dup_bytes")) {
    ctx->verify_dup_bytes = value != 0;
    return GRIT_OK;
  }
  return GRIT_EINVAL;
}

uint64_t grit_ctx_dup_mismatches(grit_ctx* ctx) {
  if (!ctx) return 0;
  unsigned long long v = 0;
  cudaSetDevice(ctx->device);
  cudaMemcpy(&v, &ctx->d_cnt->dup_mismatch, sizeof v, cudaMemcpyDeviceToHost);
  return v;
}

int grit_analyze_pages(grit_ctx* ctx, const void* d_image, uint64_t n_pages, uint32_t* d_crc, uint64_t* d_hash,
                       uint32_t* d_ref, void* stream, char* err, size_t errlen) {
  if (int rc = enter(ctx, err, errlen)) return rc;
  if (int rc = check_pages(n_pages, d_image, err, errlen)) return rc;
  if (n_pages && (!d_image || !d_crc || !d_hash || !d_ref)) return fail(err, errlen, GRIT_EINVAL, "null buffer");
  std::lock_guard<std::mutex> g(ctx->mu);
  ctx->through_image = nullptr;
  GRIT_CUDA(launch_analyze(ctx->geom, ctx->d_tables, d_image, n_pages, d_crc, d_hash, d_ref, pick(ctx, stream)));
  ctx->launches++;
  return GRIT_OK;
}

int grit_analyze_pages_through(grit_ctx* ctx, const void* d_image, uint64_t n_pages, uint32_t* d_crc, uint64_t* d_hash,
                               uint32_t* d_ref, void* d_packed, uint64_t capacity_pages, void* stream, char* err,
                               size_t errlen) {
  if (int rc = enter(ctx, err, errlen)) return rc;
  if (int rc = check_pages(n_pages, d_image, err, errlen)) return rc;
  if (int rc = check_pages(0, d_packed, err, errlen)) return rc;
  if (n_pages && (!d_image || !d_crc || !d_hash || !d_ref || !d_packed)) return fail(err, errlen, GRIT_EINVAL, "null buffer");
  std::lock_guard<std::mutex> g(ctx->mu);
  if (!n_pages) return GRIT_OK;
  return analyze_through_async(ctx, d_image, n_pages, d_crc, d_hash, d_ref, d_packed, capacity_pages, pick(ctx, stream), err,
                               errlen);
}

int grit_dedup_pages(grit_ctx* ctx, uint64_t n_pages, const uint32_t* d_crc, const uint64_t* d_hash, uint32_t* d_ref,
                     const grit_table* parent, const grit_table* peers, uint32_t n_peers, void* stream, char* err,
                     size_t errlen) {
  if (int rc = enter(ctx, err, errlen)) return rc;
  if (n_pages > GRIT_MAX_PAGES) return fail(err, errlen, GRIT_EINVAL, "n_pages too large");
  if (n_pages && (!d_crc || !d_hash || !d_ref)) return fail(err, errlen, GRIT_EINVAL, "null buffer");
  std::lock_guard<std::mutex> g(ctx->mu);
  return dedup_async(ctx, n_pages, d_crc, d_hash, d_ref, parent, peers, n_peers, pick(ctx, stream), err, errlen);
}

int grit_verify_dup_pages(grit_ctx* ctx, const void* d_image, uint64_t n_pages, uint32_t* d_ref, uint64_t* n_mismatch, void* stream,
                          char* err, size_t errlen) {
  if (int rc = enter(ctx, err, errlen)) return rc;
  if (int rc = check_pages(n_pages, d_image, err, errlen)) return rc;
  if (n_pages && (!d_image || !d_ref)) return fail(err, errlen, GRIT_EINVAL, "null buffer");
  std::lock_guard<std::mutex> g(ctx->mu);
  cudaStream_t s = pick(ctx, stream);
  unsigned long long before = 0, after = 0;
  GRIT_CUDA(cudaMemcpyAsync(&before, &ctx->d_cnt->dup_mismatch, sizeof before, cudaMemcpyDeviceToHost, s));
  GRIT_CUDA(launch_verify_dups(ctx->geom, d_image, n_pages, d_ref, ctx->d_cnt, s));
  ctx->launches++;
  GRIT_CUDA(cudaMemcpyAsync(&after, &ctx->d_cnt->dup_mismatch, sizeof after, cudaMemcpyDeviceToHost, s));
  GRIT_CUDA(cudaStreamSynchronize(s));
  if (n_mismatch) *n_mismatch = after - before;
  return GRIT_OK;
}

int grit_compact_pages(grit_ctx* ctx, const void* d_image, uint64_t n_pages, uint32_t* d_ref, void* d_packed,
                       uint64_t capacity_pages, uint64_t slot_base, uint64_t* n_stored, void* stream, char* err,
                       size_t errlen) {
  if (int rc = enter(ctx, err, errlen)) return rc;
  if (int rc = check_pages(n_pages, d_image, err, errlen)) return rc;
  if (int rc = check_pages(0, d_packed, err, errlen)) return rc;
  if (n_pages && (!d_image || !d_ref || (!d_packed && capacity_pages))) return fail(err, errlen, GRIT_EINVAL, "null buffer");
  std::lock_guard<std::mutex> g(ctx->mu);
  cudaStream_t s = pick(ctx, stream);
  if (int rc = reset_counters(ctx, s, err, errlen)) return rc;
  if (int rc = compact_async(ctx, d_image, n_pages, d_ref, d_packed, capacity_pages, slot_base, s, err, errlen)) return rc;
  if (int rc = read_counters(ctx, s, err, errlen)) return rc;
  if (n_stored) *n_stored = ctx->h_cnt->n_stored;
  if (ctx->h_cnt->overflow)
    return fail(err, errlen, GRIT_EINVAL, "packed buffer too small: %llu pages did not fit capacity %llu",
                (unsigned long long)ctx->h_cnt->overflow, (unsigned long long)capacity_pages);
  return GRIT_OK;
}

int grit_pack_buffer(grit_ctx* ctx, const void* d_image, uint64_t n_pages, const grit_table* parent,
                     const grit_table* peers, uint32_t n_peers, void* d_packed, uint64_t capacity_pages, uint32_t* d_ref,
                     uint32_t* d_crc, uint64_t* d_hash, grit_pack_stats* stats, void* stream, char* err, size_t errlen) {
  if (int rc = enter(ctx, err, errlen)) return rc;
  if (int rc = check_pages(n_pages, d_image, err, errlen)) return rc;
  if (int rc = check_pages(0, d_packed, err, errlen)) return rc;
  if (n_pages && (!d_image || !d_ref || !d_crc || !d_hash || (!d_packed && capacity_pages)))
    return fail(err, errlen, GRIT_EINVAL, "null buffer");
  std::lock_guard<std::mutex> g(ctx->mu);
  cudaStream_t s = pick(ctx, stream);
  if (int rc = reset_counters(ctx, s, err, errlen)) return rc;
  GRIT_CUDA(cudaEventRecord(ctx->ev0, s));
  // Three ways to pack, chosen per image:
  //   through  a full image whose sample looks dense (weights, optimizer state): K1 stores every page to its own slot
  //            while analysing it and K3 finds the pages in place -- N read + N written;
  //   staged   everything else: K1, K2, K3 as separate kernels -- N + 2U (also the streamed host pipelines);
  //   onepass  instead of staged with GRITPACK_ONEPASS=1: one kernel, pages parked in TMEM until their verdict --
  //            N read + U written, but not yet faster (see onepass_pack above).
  enum { kStaged, kThrough, kOnepass } path = kStaged;
  const bool full = !(parent && parent->n_pages) && n_peers == 0;
  bool staged = false;
  ctx->through_image = nullptr;
  if (n_pages && d_packed) {
    if (full && write_through_pack()) path = kThrough;  // the sample decides on the device which K1 variant runs
    if (onepass_pack() && !ctx->verify_dup_bytes) {  // a different kernel sequence: the host needs the sample's answer, at the price of a sync
      path = kOnepass;
      if (full && write_through_pack()) {
        GRIT_CUDA(launch_probe_dense(d_image, n_pages, ctx->d_cnt, s));
        ctx->launches++;
        GRIT_CUDA(cudaMemcpyAsync(&ctx->h_cnt->dense_hint, &ctx->d_cnt->dense_hint, sizeof(unsigned int), cudaMemcpyDeviceToHost, s));
        GRIT_CUDA(cudaStreamSynchronize(s));
        if (ctx->h_cnt->dense_hint) path = kThrough;
      }
    }
  }
  if (path == kOnepass) {
    DedupTable t{};
    PeerCrcs crcs{};
    if (int rc = dedup_async(ctx, n_pages, d_crc, d_hash, d_ref, parent, peers, n_peers, s, err, errlen, &t, &crcs)) return rc;
    if (int rc = ensure_status(ctx, n_pages, err, errlen)) return rc;
    const uint64_t tiles = (n_pages + onepass_tile_pages() - 1) / onepass_tile_pages();
    GRIT_CUDA(cudaMemsetAsync(ctx->d_status, 0, tiles * sizeof(unsigned long long), s));
    GRIT_CUDA(launch_pack_onepass(ctx->geom, ctx->d_tables, d_image, n_pages, d_crc, d_hash, d_ref, t, crcs, d_packed,
                                  capacity_pages, ctx->d_status, ctx->d_cnt, s));
    ctx->launches++;
  } else {
    if (path == kThrough) {
      if (int rc = analyze_through_async(ctx, d_image, n_pages, d_crc, d_hash, d_ref, d_packed, capacity_pages, s, err, errlen,
                                         onepass_pack()))
        return rc;
    } else {
      GRIT_CUDA(launch_analyze(ctx->geom, ctx->d_tables, d_image, n_pages, d_crc, d_hash, d_ref, s));
      ctx->launches++;
    }
    GRIT_CUDA(cudaEventRecord(ctx->evA, s));
    if (int rc = dedup_async(ctx, n_pages, d_crc, d_hash, d_ref, parent, peers, n_peers, s, err, errlen)) return rc;
    if (ctx->verify_dup_bytes) {  // references within the image are checked byte for byte; a mismatch is stored after all
      GRIT_CUDA(launch_verify_dups(ctx->geom, d_image, n_pages, d_ref, ctx->d_cnt, s));
      ctx->launches++;
    }
    GRIT_CUDA(cudaEventRecord(ctx->evB, s));
    staged = true;
    if (int rc = compact_async(ctx, d_image, n_pages, d_ref, d_packed, capacity_pages, 0, s, err, errlen)) return rc;
  }
  GRIT_CUDA(cudaEventRecord(ctx->ev1, s));
  if (int rc = read_counters(ctx, s, err, errlen)) return rc;
  const DevCounters& c = *ctx->h_cnt;
  if (stats) {
    stats->n_pages = n_pages;
    stats->n_zero = c.kind[0];
    stats->n_data = c.kind[1];
    stats->n_dup = c.kind[2];
    stats->n_parent = c.kind[3];
    stats->n_peer = c.kind[4];
    float ms = 0;
    cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1);
    stats->gpu_ms = ms;
    stats->stage_ms[0] = stats->stage_ms[1] = stats->stage_ms[2] = 0;
    if (staged && n_pages) {
      cudaEventElapsedTime(&ms, ctx->ev0, ctx->evA);
      stats->stage_ms[0] = ms;
      cudaEventElapsedTime(&ms, ctx->evA, ctx->evB);
      stats->stage_ms[1] = ms;
      cudaEventElapsedTime(&ms, ctx->evB, ctx->ev1);
      stats->stage_ms[2] = ms;
    }
    stats->wrote_through = path == kThrough ? (c.dense_hint ? 1 : 0) : (path == kOnepass ? 2 : 0);
  }
  if (c.overflow)
    return fail(err, errlen, GRIT_EINVAL, "packed buffer too small: %llu pages did not fit capacity %llu",
                (unsigned long long)c.overflow, (unsigned long long)capacity_pages);
  return GRIT_OK;
}

int grit_restore_buffer(grit_ctx* ctx, const void* d_packed, uint64_t n_slots, const uint32_t* d_ref,
                        const uint32_t* d_crc, uint64_t n_pages, const void* d_parent_image, uint64_t parent_pages,
                        const void* const* d_peer_images, const uint64_t* peer_pages, uint32_t n_peers, void* d_out,
                        int verify, grit_restore_stats* stats, void* stream, char* err, size_t errlen) {
  if (int rc = enter(ctx, err, errlen)) return rc;
  if (int rc = check_pages(n_pages, d_out, err, errlen)) return rc;
  if (int rc = check_pages(0, d_packed, err, errlen)) return rc;
  if (n_peers > GRIT_MAX_PEERS) return fail(err, errlen, GRIT_EINVAL, "n_peers %u > %u", n_peers, GRIT_MAX_PEERS);
  if (n_pages && (!d_ref || !d_out || (verify && !d_crc))) return fail(err, errlen, GRIT_EINVAL, "null buffer");
  PeerImages imgs{};
  imgs.parent = static_cast<const unsigned char*>(d_parent_image);
  imgs.parent_pages = d_parent_image ? parent_pages : 0;
  imgs.n_peers = n_peers;
  for (uint32_t r = 0; r < n_peers; r++) {
    imgs.peer[r] = static_cast<const unsigned char*>(d_peer_images ? d_peer_images[r] : nullptr);
    imgs.peer_pages[r] = (imgs.peer[r] && peer_pages) ? peer_pages[r] : 0;
  }
  std::lock_guard<std::mutex> g(ctx->mu);
  cudaStream_t s = pick(ctx, stream);
  if (int rc = reset_counters(ctx, s, err, errlen)) return rc;
  GRIT_CUDA(cudaEventRecord(ctx->ev0, s));
  GRIT_CUDA(launch_restore(ctx->geom, ctx->d_tables, d_packed, n_slots, d_ref, d_crc, n_pages, 0, n_pages, imgs, d_out, verify,
                           ctx->d_cnt, s));
  ctx->launches++;
  GRIT_CUDA(cudaEventRecord(ctx->ev1, s));
  if (int rc = read_counters(ctx, s, err, errlen)) return rc;
  const DevCounters& c = *ctx->h_cnt;
  if (stats) {
    stats->n_pages = n_pages;
    stats->n_bad = c.n_bad;
    stats->first_bad = c.first_bad;
    float ms = 0;
    cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1);
    stats->gpu_ms = ms;
  }
  if (c.n_bad)
    return fail(err, errlen, GRIT_ECRC, "restore: %llu page(s) failed CRC-32C or had a bad reference, first page %llu",
                (unsigned long long)c.n_bad, (unsigned long long)c.first_bad);
  return GRIT_OK;
}

int grit_synth_image(grit_ctx* ctx, void* d_image, uint64_t n_pages, uint64_t page_base, uint64_t seed,
                     uint32_t pct_zero, uint32_t pct_dup, uint32_t dup_pool, void* stream, char* err, size_t errlen) {
  if (int rc = enter(ctx, err, errlen)) return rc;
  if (int rc = check_pages(0, d_image, err, errlen)) return rc;
  std::lock_guard<std::mutex> g(ctx->mu);
  GRIT_CUDA(launch_synth(ctx->geom, d_image, n_pages, page_base, seed, pct_zero, pct_dup, dup_pool, pick(ctx, stream)));
  ctx->launches++;
  return GRIT_OK;
}

}  // extern "C"
