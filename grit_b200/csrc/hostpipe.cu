// Pinned, event-chained host<->device pipeline around the page kernels.
// See hostpipe.h; C ABI declarations in include/gritpack.h.
#include "hostpipe.h"

#include <algorithm>
#include <chrono>
#include <condition_variable>
#include <cstdlib>
#include <cstring>
#include <deque>
#include <thread>

using namespace grit;

namespace {
// pipeline chunk of the host-buffer entry points: 64 MiB unless GRITPACK_CHUNK_MB says otherwise
uint64_t host_chunk_pages() {
  static const uint64_t v = [] {
    const char* e = getenv("GRITPACK_CHUNK_MB");
    const uint64_t mb = e ? strtoull(e, nullptr, 10) : 0;
    return ((mb >= 1 && mb <= 4096 ? mb : 64) << 20) / kPage;
  }();
  return v;
}
double wall() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }
bool debug_on() {
  static const bool v = getenv("GRITPACK_DEBUG") != nullptr;
  return v;
}
}  // namespace

namespace grit {

bool is_pinned(const void* p) {
  cudaPointerAttributes a{};
  if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
    cudaGetLastError();
    return false;
  }
  return a.type == cudaMemoryTypeHost;
}

// ------------------------------------------------------------------ PackStream
PackStream::PackStream(grit_ctx* ctx, uint64_t total_pages, uint64_t chunk_pages)
    : ctx_(ctx), total_pages_(total_pages), chunk_pages_(std::max<uint64_t>(chunk_pages, 256)) {}

PackStream::~PackStream() {
  cudaSetDevice(ctx_->device);
  cudaStreamSynchronize(ctx_->h2d);
  cudaStreamSynchronize(ctx_->stream);
  cudaStreamSynchronize(ctx_->d2h);
  for (int b = 0; b < kPipeDepth; b++) {
    if (ev_h2d_[b]) cudaEventDestroy(ev_h2d_[b]);
    if (ev_kern_[b]) cudaEventDestroy(ev_kern_[b]);
    if (ev_d2h_[b]) cudaEventDestroy(ev_d2h_[b]);
  }
  // the buffers belong to the context's cache and stay for the next call
}

int PackStream::begin(const grit_table* parent, char* err, size_t errlen, const grit_table* peers, uint32_t n_peers,
                      bool peers_on_device) {
  GRIT_CUDA(cudaSetDevice(ctx_->device));
  if (total_pages_ > GRIT_MAX_PAGES) return fail(err, errlen, GRIT_EINVAL, "image exceeds GRIT_MAX_PAGES");
  const size_t chunk_bytes = chunk_pages_ * kPage;
  for (int b = 0; b < kPipeDepth; b++) {
    if (int rc = ctx_dev(ctx_, kDevIn0 + b, chunk_bytes, &d_in_[b], err, errlen)) return rc;
    if (int rc = ctx_dev(ctx_, kDevOut0 + b, chunk_bytes, &d_out_[b], err, errlen)) return rc;
    GRIT_CUDA(cudaEventCreateWithFlags(&ev_h2d_[b], cudaEventDisableTiming));
    GRIT_CUDA(cudaEventCreateWithFlags(&ev_kern_[b], cudaEventDisableTiming));
    GRIT_CUDA(cudaEventCreateWithFlags(&ev_d2h_[b], cudaEventDisableTiming));
  }
  if (int rc = ctx_pin(ctx_, kPinCounts, kPipeDepth * sizeof(unsigned long long), (void**)&h_counts_, err, errlen)) return rc;
  for (int b = 0; b < kPipeDepth; b++)
    if (int rc = ctx_pin(ctx_, kPinMeta0 + b, chunk_pages_ * 16, &h_meta_[b], err, errlen)) return rc;
  const uint64_t n = std::max<uint64_t>(total_pages_, 1);
  if (int rc = ctx_dev(ctx_, kDevRef, n * 4, (void**)&d_ref_, err, errlen)) return rc;
  if (int rc = ctx_dev(ctx_, kDevCrc, n * 4, (void**)&d_crc_, err, errlen)) return rc;
  if (int rc = ctx_dev(ctx_, kDevHash, n * 8, (void**)&d_hash_, err, errlen)) return rc;
  uint64_t items = total_pages_;
  if (parent && parent->n_pages) {
    if (parent->n_pages > GRIT_MAX_PAGES) return fail(err, errlen, GRIT_EINVAL, "parent exceeds GRIT_MAX_PAGES");
    items += parent->n_pages;
    if (int rc = ctx_dev(ctx_, kDevParentHash, parent->n_pages * 8, (void**)&d_parent_hash_, err, errlen)) return rc;
    if (int rc = ctx_dev(ctx_, kDevParentCrc, parent->n_pages * 4, (void**)&d_parent_crc_, err, errlen)) return rc;
    GRIT_CUDA(cudaMemcpyAsync(d_parent_hash_, parent->hash, parent->n_pages * 8, cudaMemcpyHostToDevice, ctx_->stream));
    GRIT_CUDA(cudaMemcpyAsync(d_parent_crc_, parent->crc, parent->n_pages * 4, cudaMemcpyHostToDevice, ctx_->stream));
    crcs_.parent.base = d_parent_crc_;
    h2d_bytes += parent->n_pages * 12.0;
  }
  if (n_peers > GRIT_MAX_PEERS) return fail(err, errlen, GRIT_EINVAL, "n_peers %u > %u", n_peers, GRIT_MAX_PEERS);
  for (uint32_t r = 0; r < n_peers; r++) {
    if (peers[r].n_pages > GRIT_MAX_PAGES) return fail(err, errlen, GRIT_EINVAL, "peer exceeds GRIT_MAX_PAGES");
    items += peers[r].n_pages;
  }
  if (int rc = ensure_table(ctx_, items, &table_, err, errlen)) return rc;
  if (int rc = ensure_status(ctx_, chunk_pages_, err, errlen)) return rc;
  cudaStream_t s = ctx_->stream;
  if (int rc = reset_counters(ctx_, s, err, errlen)) return rc;
  GRIT_CUDA(launch_table_clear(ctx_->geom, table_, s));
  ctx_->launches++;
  if (parent && parent->n_pages) {
    GRIT_CUDA(launch_table_insert(ctx_->geom, table_, d_parent_hash_, nullptr, parent->n_pages, 0, 0, s));
    ctx_->launches++;
  }
  for (uint32_t r = 0; r < n_peers; r++) {
    if (!peers[r].n_pages) continue;
    void *dh = const_cast<uint64_t*>(peers[r].hash), *dc = const_cast<uint32_t*>(peers[r].crc);
    if (!peers_on_device) {  // cached in the context like every other working buffer
      if (int rc = ctx_dev(ctx_, kDevPeerHash0 + (int)r, peers[r].n_pages * 8, &dh, err, errlen)) return rc;
      if (int rc = ctx_dev(ctx_, kDevPeerCrc0 + (int)r, peers[r].n_pages * 4, &dc, err, errlen)) return rc;
      GRIT_CUDA(cudaMemcpyAsync(dh, peers[r].hash, peers[r].n_pages * 8, cudaMemcpyHostToDevice, s));
      GRIT_CUDA(cudaMemcpyAsync(dc, peers[r].crc, peers[r].n_pages * 4, cudaMemcpyHostToDevice, s));
      h2d_bytes += peers[r].n_pages * 12.0;
    }
    crcs_.peer[r].base = static_cast<const uint32_t*>(dc);
    GRIT_CUDA(launch_table_insert(ctx_->geom, table_, static_cast<const uint64_t*>(dh), nullptr, peers[r].n_pages, 1 + r, 0, s));
    ctx_->launches++;
  }
  return GRIT_OK;
}

int PackStream::alloc_staging(bool in, bool out, char* err, size_t errlen) {
  for (int b = 0; b < kPipeDepth; b++) {
    if (in && !h_in_[b])
      if (int rc = ctx_pin(ctx_, kPinIn0 + b, chunk_pages_ * kPage, &h_in_[b], err, errlen)) return rc;
    if (out && !h_out_[b])
      if (int rc = ctx_pin(ctx_, kPinOut0 + b, chunk_pages_ * kPage, &h_out_[b], err, errlen)) return rc;
  }
  return GRIT_OK;
}

int PackStream::wait_in_free(uint64_t i, char* err, size_t errlen) {
  if (i >= (uint64_t)kPipeDepth) GRIT_CUDA(cudaEventSynchronize(ev_h2d_[i % kPipeDepth]));
  return GRIT_OK;
}

int PackStream::submit(uint64_t i, const void* h_src, uint64_t n_pages, char* err, size_t errlen) {
  const int b = (int)(i % kPipeDepth);
  if (n_pages > chunk_pages_ || pages_done_ + n_pages > total_pages_)
    return fail(err, errlen, GRIT_EINVAL, "chunk %llu out of range", (unsigned long long)i);
  cudaStream_t s = ctx_->stream;
  const uint64_t p0 = pages_done_;
  chunk_page0_[b] = p0;
  // the H2D may overwrite d_in[b] only after the compaction that last read it
  if (i >= (uint64_t)kPipeDepth) GRIT_CUDA(cudaStreamWaitEvent(ctx_->h2d, ev_kern_[b], 0));
  GRIT_CUDA(cudaMemcpyAsync(d_in_[b], h_src, n_pages * kPage, cudaMemcpyHostToDevice, ctx_->h2d));
  GRIT_CUDA(cudaEventRecord(ev_h2d_[b], ctx_->h2d));
  h2d_bytes += (double)n_pages * kPage;

  GRIT_CUDA(cudaStreamWaitEvent(s, ev_h2d_[b], 0));
  // d_out[b] is free once the D2H of its previous chunk is done
  if (i >= (uint64_t)kPipeDepth) GRIT_CUDA(cudaStreamWaitEvent(s, ev_d2h_[b], 0));
  if (!timing_started_) {
    GRIT_CUDA(cudaEventRecord(ctx_->ev0, s));
    timing_started_ = true;
  }
  if (analyze_only_) {  // first pass of a ranked directory pack: hashes only
    GRIT_CUDA(launch_analyze(ctx_->geom, ctx_->d_tables, d_in_[b], n_pages, d_crc_ + p0, d_hash_ + p0, d_ref_ + p0, s));
    ctx_->launches++;
    GRIT_CUDA(cudaEventRecord(ev_kern_[b], s));
    pages_done_ += n_pages;
    return GRIT_OK;
  }
  GRIT_CUDA(cudaMemsetAsync(ctx_->d_status, 0, (n_pages + kCompactTile - 1) / kCompactTile * sizeof(unsigned long long), s));
  GRIT_CUDA(cudaMemsetAsync(&ctx_->d_cnt->ticket, 0, 3 * sizeof(unsigned int), s));  // ticket and its neighbours
  GRIT_CUDA(launch_analyze(ctx_->geom, ctx_->d_tables, d_in_[b], n_pages, d_crc_ + p0, d_hash_ + p0, d_ref_ + p0, s));
  GRIT_CUDA(launch_table_insert(ctx_->geom, table_, d_hash_ + p0, d_ref_ + p0, n_pages, kOriginSelf, p0, s));
  GRIT_CUDA(launch_resolve(ctx_->geom, table_, p0, p0 + n_pages, d_crc_, d_hash_, d_ref_, crcs_, ctx_->d_cnt, s));
  GRIT_CUDA(launch_compact(ctx_->geom, d_in_[b], n_pages, d_ref_ + p0, d_out_[b], chunk_pages_, 0, true,
                           ctx_->d_status, ctx_->d_cnt, s));
  ctx_->launches += 4;
  GRIT_CUDA(cudaMemcpyAsync(&h_counts_[b], &ctx_->d_cnt->n_stored, sizeof(unsigned long long), cudaMemcpyDeviceToHost, s));
  // the chunk's per-page arrays are final now: bring them to pinned staging (u32 ref | u32 crc | u64 hash)
  char* meta = static_cast<char*>(h_meta_[b]);
  GRIT_CUDA(cudaMemcpyAsync(meta, d_ref_ + p0, n_pages * 4, cudaMemcpyDeviceToHost, s));
  GRIT_CUDA(cudaMemcpyAsync(meta + chunk_pages_ * 4, d_crc_ + p0, n_pages * 4, cudaMemcpyDeviceToHost, s));
  GRIT_CUDA(cudaMemcpyAsync(meta + chunk_pages_ * 8, d_hash_ + p0, n_pages * 8, cudaMemcpyDeviceToHost, s));
  chunk_pages_n_[b] = n_pages;
  d2h_bytes += n_pages * 16.0;
  // the next launch continues the slot numbering from here
  GRIT_CUDA(cudaMemcpyAsync(&ctx_->d_cnt->slot_base, &ctx_->d_cnt->slot_base_next, sizeof(unsigned long long),
                            cudaMemcpyDeviceToDevice, s));
  GRIT_CUDA(cudaEventRecord(ev_kern_[b], s));
  pages_done_ += n_pages;
  return GRIT_OK;
}

int PackStream::collect(uint64_t i, void* h_dst, uint64_t* first_slot, uint64_t* n_out, char* err, size_t errlen) {
  const int b = (int)(i % kPipeDepth);
  GRIT_CUDA(cudaEventSynchronize(ev_kern_[b]));
  const uint64_t n = h_counts_[b];
  {
    const char* meta = static_cast<const char*>(h_meta_[b]);
    const uint64_t p0 = chunk_page0_[b], np = chunk_pages_n_[b];
    if (out_ref_) memcpy(out_ref_ + p0, meta, np * 4);
    if (out_crc_) memcpy(out_crc_ + p0, meta + chunk_pages_ * 4, np * 4);
    if (out_hash_) memcpy(out_hash_ + p0, meta + chunk_pages_ * 8, np * 8);
  }
  *first_slot = slots_done_;
  *n_out = n;
  GRIT_CUDA(cudaStreamWaitEvent(ctx_->d2h, ev_kern_[b], 0));
  if (n) GRIT_CUDA(cudaMemcpyAsync(h_dst, d_out_[b], n * kPage, cudaMemcpyDeviceToHost, ctx_->d2h));
  GRIT_CUDA(cudaEventRecord(ev_d2h_[b], ctx_->d2h));
  d2h_bytes += (double)n * kPage;
  slots_done_ += n;
  return GRIT_OK;
}

int PackStream::collect_at(uint64_t i, void* h_dst_base, uint64_t capacity_pages, uint64_t* first_slot,
                           uint64_t* n_out, char* err, size_t errlen) {
  const int b = (int)(i % kPipeDepth);
  GRIT_CUDA(cudaEventSynchronize(ev_kern_[b]));
  if (slots_done_ + h_counts_[b] > capacity_pages)
    return fail(err, errlen, GRIT_EINVAL, "packed buffer too small: need more than %llu pages",
                (unsigned long long)capacity_pages);
  return collect(i, static_cast<char*>(h_dst_base) + slots_done_ * kPage, first_slot, n_out, err, errlen);
}

int PackStream::wait_chunk(uint64_t i, char* err, size_t errlen) {
  GRIT_CUDA(cudaEventSynchronize(ev_kern_[i % kPipeDepth]));
  return GRIT_OK;
}

int PackStream::wait_out(uint64_t i, char* err, size_t errlen) {
  GRIT_CUDA(cudaEventSynchronize(ev_d2h_[i % kPipeDepth]));
  return GRIT_OK;
}

int PackStream::finish(grit_pack_stats* stats, char* err, size_t errlen) {
  cudaStream_t s = ctx_->stream;
  if (timing_started_) GRIT_CUDA(cudaEventRecord(ctx_->ev1, s));
  if (int rc = read_counters(ctx_, s, err, errlen)) return rc;
  GRIT_CUDA(cudaStreamSynchronize(ctx_->d2h));
  GRIT_CUDA(cudaStreamSynchronize(ctx_->h2d));
  if (pages_done_ != total_pages_)
    return fail(err, errlen, GRIT_EINVAL, "pack stream saw %llu of %llu pages", (unsigned long long)pages_done_,
                (unsigned long long)total_pages_);
  const DevCounters& c = *ctx_->h_cnt;
  if (stats) {
    stats->n_pages = total_pages_;
    stats->n_zero = c.kind[0];
    stats->n_data = c.kind[1];
    stats->n_dup = c.kind[2];
    stats->n_parent = c.kind[3];
    stats->n_peer = c.kind[4];
    float ms = 0;
    if (timing_started_) cudaEventElapsedTime(&ms, ctx_->ev0, ctx_->ev1);
    stats->gpu_ms = ms;
    stats->stage_ms[0] = stats->stage_ms[1] = stats->stage_ms[2] = 0;
    stats->wrote_through = 0;
  }
  if (c.overflow) return fail(err, errlen, GRIT_EINVAL, "internal: chunk output overflow");
  return GRIT_OK;
}

// ------------------------------------------------------------------ restore
// The packed pages are uploaded once and stay resident; output chunks are restored and downloaded behind the
// upload.  Output chunks are taken in `lanes` interleaved runs of the page range (lane 0's first chunk, lane 1's
// first chunk, ...): with callbacks, several writer threads then have chunks of different destination files in
// hand at the same time (a file system takes a writer or two per file, a tree of files takes many).  The packed
// chunks each output chunk needs -- the slots of its stored pages and of the owners of its duplicates -- are
// uploaded just before it, each once, in the order that falls out of that.
int restore_stream(grit_ctx* ctx, const RestoreIO& io, uint64_t n_slots, const uint32_t* h_ref,
                   const uint32_t* h_crc, uint64_t n_pages, const void* parent_image, uint64_t parent_pages,
                   bool parent_on_device, int verify, uint64_t chunk_pages, grit_restore_stats* stats,
                   double* h2d_total, double* d2h_total, char* err, size_t errlen, const PeerImages* peers, int lanes,
                   int depth) {
  GRIT_CUDA(cudaSetDevice(ctx->device));
  if (n_pages > GRIT_MAX_PAGES) return fail(err, errlen, GRIT_EINVAL, "image exceeds GRIT_MAX_PAGES");
  chunk_pages = std::max<uint64_t>(chunk_pages, 256);
  const uint64_t n_chunks = (n_pages + chunk_pages - 1) / chunk_pages;
  const void* h_packed = io.packed_direct;
  void* h_out = io.out_direct;
  const bool out_pinned = h_out != nullptr, in_pinned = h_packed != nullptr;
  depth = std::min(std::max(depth, kPipeDepth), kRestoreMaxDepth);
  if (out_pinned) depth = kPipeDepth;
  lanes = (int)std::min<uint64_t>(std::max(lanes, 1), std::max<uint64_t>(n_chunks / 4, 1));

  // processing order of the output chunks
  std::vector<uint64_t> order;
  order.reserve(n_chunks);
  {
    const uint64_t per = (n_chunks + lanes - 1) / std::max(lanes, 1);
    for (uint64_t k = 0; k < per; k++)
      for (int l = 0; l < lanes; l++)
        if (l * per + k < std::min<uint64_t>((l + 1) * per, n_chunks)) order.push_back(l * per + k);
  }
  // upload order of the packed chunks, and how far into it each output chunk needs the upload to have come
  const uint64_t up_chunks = (n_slots + chunk_pages - 1) / chunk_pages;
  std::vector<uint64_t> up_order, need_upto(n_chunks, 0);
  {
    std::vector<char> scheduled(up_chunks, 0);
    std::vector<uint64_t> want;
    for (uint64_t i = 0; i < n_chunks; i++) {
      const uint64_t p0 = order[i] * chunk_pages, p1 = std::min(n_pages, p0 + chunk_pages);
      want.clear();
      for (uint64_t p = p0; p < p1; p++) {
        uint32_t x = h_ref[p];
        if (GRIT_REF_KIND(x) == GRIT_KIND_DUP && GRIT_REF_VALUE(x) < n_pages) x = h_ref[GRIT_REF_VALUE(x)];
        if (GRIT_REF_KIND(x) != GRIT_KIND_DATA) continue;
        const uint64_t c = GRIT_REF_VALUE(x) / chunk_pages;
        if (c < up_chunks && !scheduled[c]) {
          scheduled[c] = 1;
          want.push_back(c);
        }
      }
      std::sort(want.begin(), want.end());
      up_order.insert(up_order.end(), want.begin(), want.end());
      need_upto[i] = up_order.size();
    }
  }
  const uint64_t n_up = up_order.size();

  struct Res {
    void *d_packed = nullptr, *d_parent = nullptr, *d_out[kRestoreMaxDepth] = {}, *h_stage_out[kRestoreMaxDepth] = {};
    void* h_stage_in[3] = {};
    uint32_t *d_ref = nullptr, *d_crc = nullptr;
    cudaEvent_t ev_kern[kRestoreMaxDepth] = {}, ev_d2h[kRestoreMaxDepth] = {};
    cudaEvent_t ev_up = nullptr;  // the latest upload enqueued
    cudaEvent_t ev_stage[3] = {};
    ~Res() {
      cudaDeviceSynchronize();  // nothing of this call may still be running when the cached buffers are reused
      for (int b = 0; b < kRestoreMaxDepth; b++) {
        if (ev_kern[b]) cudaEventDestroy(ev_kern[b]);
        if (ev_d2h[b]) cudaEventDestroy(ev_d2h[b]);
      }
      for (int b = 0; b < 3; b++)
        if (ev_stage[b]) cudaEventDestroy(ev_stage[b]);
      if (ev_up) cudaEventDestroy(ev_up);
    }
  } r;

  cudaStream_t s = ctx->stream;
  const size_t chunk_bytes = chunk_pages * kPage;
  const double t_begin = wall();
  if (int rc = ctx_dev(ctx, kDevRPacked, std::max<uint64_t>(n_slots, 1) * kPage, &r.d_packed, err, errlen)) return rc;
  if (int rc = ctx_dev(ctx, kDevRRef, std::max<uint64_t>(n_pages, 1) * 4, (void**)&r.d_ref, err, errlen)) return rc;
  if (int rc = ctx_dev(ctx, kDevRCrc, std::max<uint64_t>(n_pages, 1) * 4, (void**)&r.d_crc, err, errlen)) return rc;
  GRIT_CUDA(cudaEventCreateWithFlags(&r.ev_up, cudaEventDisableTiming));
  for (int b = 0; b < depth; b++) {
    if (int rc = ctx_dev(ctx, b < 3 ? kDevROut0 + b : kDevROutX0 + b - 3, chunk_bytes, &r.d_out[b], err, errlen)) return rc;
    GRIT_CUDA(cudaEventCreateWithFlags(&r.ev_kern[b], cudaEventDisableTiming));
    GRIT_CUDA(cudaEventCreateWithFlags(&r.ev_d2h[b], cudaEventDisableTiming));
    if (!out_pinned)
      if (int rc = ctx_pin(ctx, b < 3 ? kPinRStageOut0 + b : kPinRStageOutX0 + b - 3, chunk_bytes, &r.h_stage_out[b], err, errlen))
        return rc;
  }
  double h2d = 0, d2h = 0;
  if (n_pages) {
    GRIT_CUDA(cudaMemcpyAsync(r.d_ref, h_ref, n_pages * 4, cudaMemcpyHostToDevice, s));
    if (h_crc) GRIT_CUDA(cudaMemcpyAsync(r.d_crc, h_crc, n_pages * 4, cudaMemcpyHostToDevice, s));
    h2d += n_pages * 8.0;
  }
  PeerImages imgs{};
  if (peers) imgs = *peers;
  imgs.parent = nullptr;
  imgs.parent_pages = 0;
  if (parent_image && parent_pages) {
    if (parent_on_device) {
      r.d_parent = const_cast<void*>(parent_image);
    } else {
      if (int rc = ctx_dev(ctx, kDevRParent, parent_pages * kPage, &r.d_parent, err, errlen)) return rc;
      GRIT_CUDA(cudaMemcpyAsync(r.d_parent, parent_image, parent_pages * kPage, cudaMemcpyHostToDevice, s));
      h2d += (double)parent_pages * kPage;
    }
    imgs.parent = static_cast<const unsigned char*>(r.d_parent);
    imgs.parent_pages = parent_pages;
  }
  if (int rc = reset_counters(ctx, s, err, errlen)) return rc;

  constexpr int kStageIn = 3;
  if (!in_pinned && n_up) {
    for (int b = 0; b < kStageIn; b++) {
      if (int rc = ctx_pin(ctx, kPinRStageIn0 + b, chunk_bytes, &r.h_stage_in[b], err, errlen)) return rc;
      GRIT_CUDA(cudaEventCreateWithFlags(&r.ev_stage[b], cudaEventDisableTiming));
    }
  }

  // Callback I/O runs on helper threads so that reading packed pages, the GPU pipeline and writing restored pages
  // overlap: the reader keeps the staging-in buffers full ahead of the upload, the writers drain staging-out
  // buffers as their D2H copies land.
  struct Helpers {
    std::mutex mu;
    std::condition_variable cv;
    uint64_t filled = 0, consumed = 0;  // reader: chunks read / chunks whose H2D is enqueued
    struct Job {
      uint64_t p0, p1;
      int b;
    };
    std::deque<Job> jobs;  // writers
    uint64_t submitted[kRestoreMaxDepth] = {}, finished[kRestoreMaxDepth] = {};  // per staging-out buffer
    int rc = 0;
    bool stop = false;
    double t_read = 0, t_write = 0, t_main_wait_read = 0, t_main_wait_write = 0;  // GRITPACK_DEBUG
    std::thread reader;
    std::vector<std::thread> writers;
    ~Helpers() {
      {
        std::lock_guard<std::mutex> l(mu);
        stop = true;
      }
      cv.notify_all();
      if (reader.joinable()) reader.join();
      for (auto& w : writers) w.join();
    }
  } hp;
  if (!in_pinned && n_up)
    hp.reader = std::thread([&] {
      cudaSetDevice(ctx->device);
      for (uint64_t k = 0; k < n_up; k++) {
        const int b = (int)(k % kStageIn);
        if (k >= (uint64_t)kStageIn) {  // the buffer's previous chunk must have left for the device
          std::unique_lock<std::mutex> l(hp.mu);
          hp.cv.wait(l, [&] { return hp.stop || hp.consumed > k - kStageIn; });
          if (hp.stop) return;
          l.unlock();
          cudaEventSynchronize(r.ev_stage[b]);
        }
        const uint64_t s0 = up_order[k] * chunk_pages, n = std::min(chunk_pages, n_slots - s0);
        const double t0 = wall();
        const int rc = io.read_packed(s0, n, r.h_stage_in[b]);
        std::lock_guard<std::mutex> l(hp.mu);
        hp.t_read += wall() - t0;
        if (rc && !hp.rc) hp.rc = rc;
        hp.filled = k + 1;
        hp.cv.notify_all();
        if (rc) return;
      }
    });
  if (!out_pinned && n_chunks)
    for (int w = 0; w < std::min(depth - 1, 4); w++)
      hp.writers.emplace_back([&] {
        cudaSetDevice(ctx->device);
        for (;;) {
          Helpers::Job j;
          {
            std::unique_lock<std::mutex> l(hp.mu);
            hp.cv.wait(l, [&] { return hp.stop || !hp.jobs.empty(); });
            if (hp.jobs.empty()) return;
            j = hp.jobs.front();
            hp.jobs.pop_front();
          }
          int rc = cudaEventSynchronize(r.ev_d2h[j.b]) == cudaSuccess ? 0 : GRIT_ECUDA;
          const double t1 = wall();
          if (!rc) rc = io.write_out(j.p0, j.p1, r.h_stage_out[j.b]);
          std::lock_guard<std::mutex> l(hp.mu);
          hp.t_write += wall() - t1;
          if (rc && !hp.rc) hp.rc = rc;
          hp.finished[j.b]++;
          hp.cv.notify_all();
        }
      });

  uint64_t uploaded = 0;  // positions of up_order enqueued so far
  auto upload_upto = [&](uint64_t pos) -> int {
    for (; uploaded < pos; uploaded++) {
      const uint64_t s0 = up_order[uploaded] * chunk_pages, n = std::min(chunk_pages, n_slots - s0);
      const char* src = static_cast<const char*>(h_packed) + s0 * kPage;
      const int b = (int)(uploaded % kStageIn);
      if (!in_pinned) {  // no pinned source: the reader thread fills pinned staging
        const double t0 = wall();
        std::unique_lock<std::mutex> l(hp.mu);
        hp.cv.wait(l, [&] { return hp.rc || hp.filled > uploaded; });
        hp.t_main_wait_read += wall() - t0;
        if (hp.rc) return fail(err, errlen, hp.rc, "reading packed pages failed");
        src = static_cast<const char*>(r.h_stage_in[b]);
      }
      GRIT_CUDA(cudaMemcpyAsync(static_cast<char*>(r.d_packed) + s0 * kPage, src, n * kPage, cudaMemcpyHostToDevice, ctx->h2d));
      if (!in_pinned) {
        GRIT_CUDA(cudaEventRecord(r.ev_stage[b], ctx->h2d));
        std::lock_guard<std::mutex> l(hp.mu);
        hp.consumed = uploaded + 1;
        hp.cv.notify_all();
      }
      h2d += (double)n * kPage;
    }
    return GRIT_OK;
  };

  GRIT_CUDA(cudaEventRecord(ctx->ev0, s));
  const double t_loop = wall();
  for (uint64_t i = 0; i < n_chunks; i++) {
    const int b = (int)(i % depth);
    const uint64_t p0 = order[i] * chunk_pages, p1 = std::min(n_pages, p0 + chunk_pages);
    if (uploaded < need_upto[i]) {
      if (int rc = upload_upto(need_upto[i])) return rc;
      GRIT_CUDA(cudaEventRecord(r.ev_up, ctx->h2d));  // one copy stream: the latest upload covers all before it
    }
    if (uploaded) GRIT_CUDA(cudaStreamWaitEvent(s, r.ev_up, 0));
    if (i >= (uint64_t)depth) GRIT_CUDA(cudaStreamWaitEvent(s, r.ev_d2h[b], 0));
    GRIT_CUDA(launch_restore(ctx->geom, ctx->d_tables, r.d_packed, n_slots, r.d_ref, r.d_crc, n_pages, p0, p1, imgs,
                             r.d_out[b], verify, ctx->d_cnt, s));
    ctx->launches++;
    GRIT_CUDA(cudaEventRecord(r.ev_kern[b], s));
    GRIT_CUDA(cudaStreamWaitEvent(ctx->d2h, r.ev_kern[b], 0));
    char* dst = static_cast<char*>(h_out) + p0 * kPage;
    if (!out_pinned) {
      {  // staging buffer b still holds an earlier chunk until a writer is done with it
        const double t0 = wall();
        std::unique_lock<std::mutex> l(hp.mu);
        hp.cv.wait(l, [&] { return hp.rc || hp.finished[b] == hp.submitted[b]; });
        hp.t_main_wait_write += wall() - t0;
        if (hp.rc) return fail(err, errlen, hp.rc, "writing restored pages failed");
      }
      dst = static_cast<char*>(r.h_stage_out[b]);
    }
    GRIT_CUDA(cudaMemcpyAsync(dst, r.d_out[b], (p1 - p0) * kPage, cudaMemcpyDeviceToHost, ctx->d2h));
    GRIT_CUDA(cudaEventRecord(r.ev_d2h[b], ctx->d2h));
    d2h += (double)(p1 - p0) * kPage;
    if (!out_pinned) {
      std::lock_guard<std::mutex> l(hp.mu);
      hp.submitted[b]++;
      hp.jobs.push_back({p0, p1, b});
      hp.cv.notify_all();
    }
  }
  GRIT_CUDA(cudaEventRecord(ctx->ev1, s));
  if (int rc = read_counters(ctx, s, err, errlen)) return rc;
  GRIT_CUDA(cudaStreamSynchronize(ctx->d2h));
  GRIT_CUDA(cudaStreamSynchronize(ctx->h2d));
  if (!out_pinned) {  // the writers finish the tail
    std::unique_lock<std::mutex> l(hp.mu);
    hp.cv.wait(l, [&] {
      if (hp.rc) return true;
      for (int b = 0; b < depth; b++)
        if (hp.finished[b] != hp.submitted[b]) return false;
      return true;
    });
    if (hp.rc) return fail(err, errlen, hp.rc, "writing restored pages failed");
  }
  if (h2d_total) *h2d_total += h2d;
  if (d2h_total) *d2h_total += d2h;
  if (debug_on())
    fprintf(stderr,
            "[gritpack] restore_stream: alloc %.1f ms, stream %.1f ms (%d lanes, depth %d; reader busy %.1f, writers busy %.1f "
            "in all, main thread waited %.1f for the reader and %.1f for the writers)\n",
            (t_loop - t_begin) * 1e3, (wall() - t_loop) * 1e3, lanes, depth, hp.t_read * 1e3, hp.t_write * 1e3,
            hp.t_main_wait_read * 1e3, hp.t_main_wait_write * 1e3);
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


// ------------------------------------------------------------------ restore into HBM, packed input through a ring
// BASELINE configs[4]: the destination -- the whole image, up to all of HBM -- is the one big allocation; the packed
// pages never are.  They arrive chunk by chunk in a small device ring; slots are numbered in page order, so packed
// chunk c serves a run of consecutive pages [first page whose slot is in c, first page whose slot is in c + 1), and
// one K4 launch restores that run straight into its place in d_out.  A duplicate of a page restored by an earlier
// launch is copied from d_out (its owner's slot may have left the ring); parent / peer pages come from their
// images as always.  Device memory beside d_out: ring (4 chunks) + 8 bytes per page.
int restore_to_device(grit_ctx* ctx, const void* h_packed, const std::function<int(uint64_t, uint64_t, void*)>& read_packed,
                      uint64_t n_slots, const uint32_t* h_ref, const uint32_t* h_crc, uint64_t n_pages,
                      const void* d_parent, uint64_t parent_pages, const PeerImages* peers, void* d_out, int verify,
                      uint64_t chunk_pages, grit_restore_stats* stats, double* h2d_total, char* err, size_t errlen) {
  GRIT_CUDA(cudaSetDevice(ctx->device));
  if (n_pages > GRIT_MAX_PAGES) return fail(err, errlen, GRIT_EINVAL, "image exceeds GRIT_MAX_PAGES");
  chunk_pages = std::max<uint64_t>(chunk_pages, 256);
  constexpr int kRing = 4, kStage = 3;
  const bool in_pinned = h_packed && is_pinned(h_packed);
  const uint64_t up_chunks = (n_slots + chunk_pages - 1) / chunk_pages;
  // run of pages each packed chunk serves: page_end[c] = first page whose stored slot lies beyond chunk c
  std::vector<uint64_t> page_end(std::max<uint64_t>(up_chunks, 1), n_pages);
  {
    uint64_t c = 0;
    for (uint64_t p = 0; p < n_pages && up_chunks; p++)
      if (GRIT_REF_KIND(h_ref[p]) == GRIT_KIND_DATA) {
        const uint64_t pc = GRIT_REF_VALUE(h_ref[p]) / chunk_pages;
        if (pc < c || pc >= up_chunks)
          return fail(err, errlen, GRIT_ECRC, "restore: page %llu stores slot %u out of order; the image cannot be streamed",
                      (unsigned long long)p, GRIT_REF_VALUE(h_ref[p]));
        while (c < pc) page_end[c++] = p;
      }
  }
  struct Res {
    cudaEvent_t ev_up[kRing] = {}, ev_kern[kRing] = {}, ev_stage[kStage] = {};
    ~Res() {
      cudaDeviceSynchronize();
      for (auto e : ev_up)
        if (e) cudaEventDestroy(e);
      for (auto e : ev_kern)
        if (e) cudaEventDestroy(e);
      for (auto e : ev_stage)
        if (e) cudaEventDestroy(e);
    }
  } r;
  void *d_ring[kRing] = {}, *h_stage[kStage] = {};
  uint32_t *d_ref = nullptr, *d_crc = nullptr;
  const size_t chunk_bytes = chunk_pages * kPage;
  for (int b = 0; b < kRing; b++) {
    if (int rc = ctx_dev(ctx, b < 3 ? kDevROut0 + b : kDevROutX0, chunk_bytes, &d_ring[b], err, errlen)) return rc;
    GRIT_CUDA(cudaEventCreateWithFlags(&r.ev_up[b], cudaEventDisableTiming));
    GRIT_CUDA(cudaEventCreateWithFlags(&r.ev_kern[b], cudaEventDisableTiming));
  }
  if (int rc = ctx_dev(ctx, kDevRRef, std::max<uint64_t>(n_pages, 1) * 4, (void**)&d_ref, err, errlen)) return rc;
  if (int rc = ctx_dev(ctx, kDevRCrc, std::max<uint64_t>(n_pages, 1) * 4, (void**)&d_crc, err, errlen)) return rc;
  if (!in_pinned && up_chunks)
    for (int b = 0; b < kStage; b++) {
      if (int rc = ctx_pin(ctx, kPinRStageIn0 + b, chunk_bytes, &h_stage[b], err, errlen)) return rc;
      GRIT_CUDA(cudaEventCreateWithFlags(&r.ev_stage[b], cudaEventDisableTiming));
    }
  cudaStream_t s = ctx->stream;
  double h2d = 0;
  if (n_pages) {
    GRIT_CUDA(cudaMemcpyAsync(d_ref, h_ref, n_pages * 4, cudaMemcpyHostToDevice, s));
    if (h_crc) GRIT_CUDA(cudaMemcpyAsync(d_crc, h_crc, n_pages * 4, cudaMemcpyHostToDevice, s));
    h2d += n_pages * 8.0;
  }
  PeerImages imgs{};
  if (peers) imgs = *peers;
  imgs.parent = static_cast<const unsigned char*>(d_parent);
  imgs.parent_pages = d_parent ? parent_pages : 0;
  if (int rc = reset_counters(ctx, s, err, errlen)) return rc;

  // without a pinned source a reader thread keeps pinned staging full ahead of the upload
  struct Reader {
    std::mutex mu;
    std::condition_variable cv;
    uint64_t filled = 0, consumed = 0;
    int rc = 0;
    bool stop = false;
    std::thread th;
    ~Reader() {
      {
        std::lock_guard<std::mutex> l(mu);
        stop = true;
      }
      cv.notify_all();
      if (th.joinable()) th.join();
    }
  } rd;
  if (!in_pinned && up_chunks)
    rd.th = std::thread([&] {
      cudaSetDevice(ctx->device);
      for (uint64_t c = 0; c < up_chunks; c++) {
        const int b = (int)(c % kStage);
        if (c >= (uint64_t)kStage) {
          std::unique_lock<std::mutex> l(rd.mu);
          rd.cv.wait(l, [&] { return rd.stop || rd.consumed > c - kStage; });
          if (rd.stop) return;
          l.unlock();
          cudaEventSynchronize(r.ev_stage[b]);
        }
        const uint64_t s0 = c * chunk_pages, n = std::min(chunk_pages, n_slots - s0);
        int rc = 0;
        if (read_packed) rc = read_packed(s0, n, h_stage[b]);
        else memcpy(h_stage[b], static_cast<const char*>(h_packed) + s0 * kPage, n * kPage);
        std::lock_guard<std::mutex> l(rd.mu);
        if (rc && !rd.rc) rd.rc = rc;
        rd.filled = c + 1;
        rd.cv.notify_all();
        if (rc) return;
      }
    });

  GRIT_CUDA(cudaEventRecord(ctx->ev0, s));
  uint64_t page0 = 0;
  for (uint64_t c = 0; c < std::max<uint64_t>(up_chunks, 1); c++) {
    const int b = (int)(c % kRing);
    uint64_t n = 0;
    if (up_chunks) {
      const uint64_t s0 = c * chunk_pages;
      n = std::min(chunk_pages, n_slots - s0);
      const char* src = static_cast<const char*>(h_packed) + s0 * kPage;
      if (!in_pinned) {
        std::unique_lock<std::mutex> l(rd.mu);
        rd.cv.wait(l, [&] { return rd.rc || rd.filled > c; });
        if (rd.rc) return fail(err, errlen, rd.rc, "reading packed pages failed");
        src = static_cast<const char*>(h_stage[c % kStage]);
      }
      if (c >= (uint64_t)kRing) GRIT_CUDA(cudaStreamWaitEvent(ctx->h2d, r.ev_kern[b], 0));  // ring slot still being read
      GRIT_CUDA(cudaMemcpyAsync(d_ring[b], src, n * kPage, cudaMemcpyHostToDevice, ctx->h2d));
      GRIT_CUDA(cudaEventRecord(r.ev_up[b], ctx->h2d));
      if (!in_pinned) {
        GRIT_CUDA(cudaEventRecord(r.ev_stage[c % kStage], ctx->h2d));
        std::lock_guard<std::mutex> l(rd.mu);
        rd.consumed = c + 1;
        rd.cv.notify_all();
      }
      h2d += (double)n * kPage;
      GRIT_CUDA(cudaStreamWaitEvent(s, r.ev_up[b], 0));
    }
    const uint64_t page1 = page_end[c];
    PackedWindow win;
    win.slot0 = c * chunk_pages;
    win.restored = static_cast<const uint4*>(d_out);
    GRIT_CUDA(launch_restore(ctx->geom, ctx->d_tables, d_ring[b], n, d_ref, d_crc, n_pages, page0, page1, imgs,
                             static_cast<char*>(d_out) + page0 * kPage, verify, ctx->d_cnt, s, win));
    ctx->launches++;
    GRIT_CUDA(cudaEventRecord(r.ev_kern[b], s));
    page0 = page1;
  }
  GRIT_CUDA(cudaEventRecord(ctx->ev1, s));
  if (int rc = read_counters(ctx, s, err, errlen)) return rc;
  GRIT_CUDA(cudaStreamSynchronize(ctx->h2d));
  if (h2d_total) *h2d_total += h2d;
  const DevCounters& k = *ctx->h_cnt;
  if (stats) {
    stats->n_pages = n_pages;
    stats->n_bad = k.n_bad;
    stats->first_bad = k.first_bad;
    float ms = 0;
    cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1);
    stats->gpu_ms = ms;
  }
  if (k.n_bad)
    return fail(err, errlen, GRIT_ECRC, "restore: %llu page(s) failed CRC-32C or had a bad reference, first page %llu",
                (unsigned long long)k.n_bad, (unsigned long long)k.first_bad);
  return GRIT_OK;
}

}  // namespace grit

extern "C" {

int grit_restore_to_device(grit_ctx* ctx, const void* h_packed, grit_read_packed_fn read, void* user, uint64_t n_slots,
                           const uint32_t* h_ref, const uint32_t* h_crc, uint64_t n_pages, const void* d_parent_image,
                           uint64_t parent_pages, const void* const* d_peer_images, const uint64_t* peer_pages,
                           uint32_t n_peers, void* d_out, int verify, uint64_t chunk_bytes, grit_restore_stats* stats,
                           char* err, size_t errlen) {
  if (!ctx) return fail(err, errlen, GRIT_EINVAL, "null context");
  if (n_peers > GRIT_MAX_PEERS) return fail(err, errlen, GRIT_EINVAL, "n_peers %u > %u", n_peers, GRIT_MAX_PEERS);
  if (n_pages && (!h_ref || !d_out || (verify && !h_crc) || (n_slots && !h_packed && !read)))
    return fail(err, errlen, GRIT_EINVAL, "null buffer");
  if (reinterpret_cast<uintptr_t>(d_out) & 15) return fail(err, errlen, GRIT_EINVAL, "buffer is not 16-byte aligned");
  PeerImages imgs{};
  imgs.n_peers = n_peers;
  for (uint32_t r = 0; r < n_peers; r++) {
    imgs.peer[r] = static_cast<const unsigned char*>(d_peer_images ? d_peer_images[r] : nullptr);
    imgs.peer_pages[r] = (imgs.peer[r] && peer_pages) ? peer_pages[r] : 0;
  }
  std::function<int(uint64_t, uint64_t, void*)> cb;
  if (read) cb = [read, user](uint64_t s0, uint64_t n, void* dst) { return read(user, s0, n, dst); };
  std::lock_guard<std::mutex> g(ctx->mu);
  const uint64_t chunk_pages = chunk_bytes >= (1u << 20) ? chunk_bytes / kPage : host_chunk_pages();
  return restore_to_device(ctx, read ? nullptr : h_packed, cb, n_slots, h_ref, h_crc, n_pages, d_parent_image, parent_pages,
                           &imgs, d_out, verify, chunk_pages, stats, nullptr, err, errlen);
}

int grit_pin_host(void* h_ptr, size_t bytes) {
  return cudaHostRegister(h_ptr, bytes, cudaHostRegisterDefault) == cudaSuccess ? GRIT_OK : GRIT_ECUDA;
}
int grit_unpin_host(void* h_ptr) { return cudaHostUnregister(h_ptr) == cudaSuccess ? GRIT_OK : GRIT_ECUDA; }

int grit_pack_host(grit_ctx* ctx, const void* h_image, uint64_t n_pages, const grit_table* parent, void* h_packed,
                   uint64_t capacity_pages, uint32_t* h_ref, uint32_t* h_crc, uint64_t* h_hash, grit_pack_stats* stats,
                   char* err, size_t errlen) {
  if (!ctx) return fail(err, errlen, GRIT_EINVAL, "null context");
  if (n_pages && (!h_image || !h_ref || !h_crc || !h_hash || (!h_packed && capacity_pages)))
    return fail(err, errlen, GRIT_EINVAL, "null buffer");
  std::lock_guard<std::mutex> g(ctx->mu);
  GRIT_CUDA(cudaSetDevice(ctx->device));
  const uint64_t chunk_pages = host_chunk_pages();
  const uint64_t n_chunks = (n_pages + chunk_pages - 1) / chunk_pages;
  const bool in_pinned = is_pinned(h_image), out_pinned = is_pinned(h_packed);
  const double t_begin = wall();
  PackStream ps(ctx, n_pages, chunk_pages);
  ps.set_outputs(h_ref, h_crc, h_hash);
  if (int rc = ps.begin(parent, err, errlen)) return rc;
  const double t_loop = wall();
  // pageable caller memory is bounced through pinned staging (one memcpy each way)
  if (n_pages)
    if (int rc = ps.alloc_staging(!in_pinned, !out_pinned, err, errlen)) return rc;
  struct Pending {
    uint64_t first_slot, n_out;
  } pend[kPipeDepth] = {};
  auto drain = [&](uint64_t i) -> int {  // chunk i's D2H -> caller memory (staged mode only)
    if (int rc = ps.wait_out(i, err, errlen)) return rc;
    if (!out_pinned) {
      const Pending& p = pend[i % kPipeDepth];
      memcpy(static_cast<char*>(h_packed) + p.first_slot * kPage, ps.out_buffer(i), p.n_out * kPage);
    }
    return GRIT_OK;
  };
  for (uint64_t i = 0; i <= n_chunks; i++) {
    if (i < n_chunks) {
      const uint64_t p0 = i * chunk_pages, n = std::min(chunk_pages, n_pages - p0);
      const void* src = static_cast<const char*>(h_image) + p0 * kPage;
      if (!in_pinned) {
        if (int rc = ps.wait_in_free(i, err, errlen)) return rc;
        memcpy(ps.in_buffer(i), src, n * kPage);
        src = ps.in_buffer(i);
      }
      if (i >= (uint64_t)kPipeDepth && !out_pinned)  // staging out buffer about to be reused
        if (int rc = drain(i - kPipeDepth)) return rc;
      if (int rc = ps.submit(i, src, n, err, errlen)) return rc;
    }
    if (i >= 1) {
      const uint64_t j = i - 1;
      Pending& p = pend[j % kPipeDepth];
      int rc;
      if (out_pinned) {
        rc = ps.collect_at(j, h_packed, capacity_pages, &p.first_slot, &p.n_out, err, errlen);
      } else {
        rc = ps.collect(j, ps.out_buffer(j), &p.first_slot, &p.n_out, err, errlen);
        if (!rc && p.first_slot + p.n_out > capacity_pages)
          rc = fail(err, errlen, GRIT_EINVAL, "packed buffer too small: need more than %llu pages",
                    (unsigned long long)capacity_pages);
      }
      if (rc) return rc;
    }
  }
  if (!out_pinned) {
    const uint64_t first = n_chunks > (uint64_t)kPipeDepth ? n_chunks - kPipeDepth : 0;
    for (uint64_t i = first; i < n_chunks; i++)
      if (int rc = drain(i)) return rc;
  }
  const double t_fin = wall();
  const int rc = ps.finish(stats, err, errlen);
  if (debug_on())
    fprintf(stderr, "[gritpack] pack_host: begin %.1f ms, stream %.1f ms, finish %.1f ms\n", (t_loop - t_begin) * 1e3,
            (t_fin - t_loop) * 1e3, (wall() - t_fin) * 1e3);
  return rc;
}

int grit_restore_host(grit_ctx* ctx, const void* h_packed, uint64_t n_slots, const uint32_t* h_ref,
                      const uint32_t* h_crc, uint64_t n_pages, const void* h_parent_image, uint64_t parent_pages,
                      void* h_out, int verify, grit_restore_stats* stats, char* err, size_t errlen) {
  if (!ctx) return fail(err, errlen, GRIT_EINVAL, "null context");
  if (n_pages && (!h_ref || !h_out || (verify && !h_crc) || (n_slots && !h_packed)))
    return fail(err, errlen, GRIT_EINVAL, "null buffer");
  std::lock_guard<std::mutex> g(ctx->mu);
  RestoreIO io;
  if (is_pinned(h_packed)) {
    io.packed_direct = h_packed;
  } else {
    io.read_packed = [&](uint64_t s0, uint64_t n, void* dst) {
      memcpy(dst, static_cast<const char*>(h_packed) + s0 * kPage, n * kPage);
      return 0;
    };
  }
  if (is_pinned(h_out)) {
    io.out_direct = h_out;
  } else {
    io.write_out = [&](uint64_t p0, uint64_t p1, const void* src) {
      memcpy(static_cast<char*>(h_out) + p0 * kPage, src, (p1 - p0) * kPage);
      return 0;
    };
  }
  return restore_stream(ctx, io, n_slots, h_ref, h_crc, n_pages, h_parent_image, parent_pages, false, verify,
                        host_chunk_pages(), stats, nullptr, nullptr, err, errlen);
}

}  // extern "C"
