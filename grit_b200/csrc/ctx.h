// grit_ctx: per-device state behind the C ABI (include/gritpack.h).
#pragma once
#include <cuda_runtime.h>

#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <mutex>

#include "../../include/gritpack.h"
#include "kernels.h"

struct grit_ctx {
  int device = -1;
  grit::LaunchGeom geom{};
  cudaStream_t stream = nullptr;  // kernels
  cudaStream_t h2d = nullptr;     // host pipeline copy-in
  cudaStream_t d2h = nullptr;     // host pipeline copy-out
  cudaEvent_t ev0 = nullptr, ev1 = nullptr, evA = nullptr, evB = nullptr;
  grit::Gf2Tables* d_tables = nullptr;
  grit::DevCounters* d_cnt = nullptr;
  grit::DevCounters* h_cnt = nullptr;  // pinned mirror
  // scratch, grown on demand and kept for the life of the context
  grit::TableSlot* d_slots = nullptr;
  uint64_t table_cap = 0;
  unsigned long long* d_status = nullptr;
  uint64_t status_cap = 0;
  uint64_t launches = 0;
  bool verify_dup_bytes = false;  // grit_ctx_set_option
  std::mutex mu;
  // the last write-through K1 on this context: K3 on the same image and packed buffer leaves pages in place
  const void* through_image = nullptr;
  const void* through_packed = nullptr;
  uint64_t through_pages = 0;
  // Working buffers of the host pipelines, kept between calls (grown on demand, freed with the
  // context) so that a steady stream of packs/restores does no cudaMalloc/cudaFree at all.
  struct Buf {
    void* p = nullptr;
    size_t cap = 0;
  };
  Buf dev[48], pin[32];
};

namespace grit {

inline int fail(char* err, size_t errlen, int code, const char* fmt, ...) {
  if (err && errlen) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(err, errlen, fmt, ap);
    va_end(ap);
  }
  return code;
}

#define GRIT_CUDA(call)                                                                         \
  do {                                                                                          \
    cudaError_t e_ = (call);                                                                    \
    if (e_ != cudaSuccess)                                                                      \
      return grit::fail(err, errlen, GRIT_ECUDA, "%s: %s (%s:%d)", #call, cudaGetErrorString(e_), \
                        __FILE__, __LINE__);                                                    \
  } while (0)

// internal stage helpers shared by api.cu and hostpipe.cu (all asynchronous on s)
int ensure_table(grit_ctx* ctx, uint64_t n_items, DedupTable* t, char* err, size_t errlen);
int ensure_status(grit_ctx* ctx, uint64_t n_pages, char* err, size_t errlen);
int reset_counters(grit_ctx* ctx, cudaStream_t s, char* err, size_t errlen);
int read_counters(grit_ctx* ctx, cudaStream_t s, char* err, size_t errlen);  // syncs s
// K2 with the other images given as views (contiguous arrays or slices gathered by comm.cu); lists_ready, if set,
// is waited for on s before the lists are read
int dedup_views_async(grit_ctx* ctx, uint64_t n_pages, const uint32_t* d_crc, const uint64_t* d_hash, uint32_t* d_ref,
                      const PeerCrcs& crcs, const ProbeLists& lists, cudaStream_t s, char* err, size_t errlen,
                      cudaEvent_t lists_ready = nullptr);
// K3 into d_packed (status words sized and cleared here)
int compact_async(grit_ctx* ctx, const void* d_image, uint64_t n_pages, uint32_t* d_ref, void* d_packed,
                  uint64_t capacity_pages, uint64_t slot_base, cudaStream_t s, char* err, size_t errlen);
// cached device / pinned-host buffer number `id` of at least `bytes` bytes (contents undefined)
int ctx_dev(grit_ctx* ctx, int id, size_t bytes, void** out, char* err, size_t errlen);
int ctx_pin(grit_ctx* ctx, int id, size_t bytes, void** out, char* err, size_t errlen);
enum DevBuf { kDevIn0 = 0, kDevOut0 = 3, kDevRef = 6, kDevCrc, kDevHash, kDevParentHash, kDevParentCrc,
              kDevRPacked, kDevRRef, kDevRCrc, kDevROut0 = 14, kDevRParent = 17,
              kDevPeerHash0 = 18 /* ..25 */, kDevPeerCrc0 = 26 /* ..33 */, kDevCommHash = 34, kDevCommCrc, kDevCommPad, kDevROutX0 = 37 /* ..41 */ };
enum PinBuf { kPinIn0 = 0, kPinOut0 = 3, kPinCounts = 6, kPinMeta0 = 7, kPinRStageOut0 = 10, kPinRStageIn0 = 13 /* ..15 */,
              kPinRStageOutX0 = 16 /* ..20 */ };
// comm.cu, for the directory packer: the context a communicator lives on, its rank, and a whole-table exchange --
// every rank's (hash64, crc) arrays gathered into device buffers of the context; lower[r] (r < rank) receives DEVICE
// pointers to rank r's tables (valid until the next exchange on this context).  Collective; synchronises.
grit_ctx* comm_context(grit_comm* c);
int comm_rank_of(const grit_comm* c);
int comm_gather_tables(grit_comm* c, const uint64_t* d_hash, const uint32_t* d_crc, uint64_t n_pages, grit_table* lower,
                       char* err, size_t errlen);
bool write_through_pack();  // false with GRITPACK_WRITE_THROUGH=0
bool onepass_pack();  // true with GRITPACK_ONEPASS=1: grit_pack_buffer packs non-dense images with the single-pass kernel

}  // namespace grit
