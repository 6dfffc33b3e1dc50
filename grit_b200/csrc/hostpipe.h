// Streaming pack / restore of an image that lives in host memory or in files:
// pinned buffers, H2D on one stream, kernels on a second, D2H on a third,
// chained by events so that chunk i+1 uploads while chunk i is analysed and
// chunk i-1 downloads.  Used by grit_pack_host / grit_restore_host (hostpipe.cu)
// and by the directory packer (dirpack.cpp).
//
// Reference counterpart of the whole pipeline: the io.Copy inside copyFile,
// pkg/gritagent/copy/copy.go:79.
#pragma once
#include <cstdint>
#include <functional>
#include <vector>

#include "ctx.h"

namespace grit {

constexpr int kPipeDepth = 3;  // chunks in flight
constexpr int kRestoreMaxDepth = 8;  // output chunks in flight of a restore that writes through callbacks

// Packs a stream of page chunks.  Call order: begin, then for each chunk
// submit(i) ... collect(i) (collect may lag submit by up to kPipeDepth-1), finish.
class PackStream {
 public:
  PackStream(grit_ctx* ctx, uint64_t total_pages, uint64_t chunk_pages);
  ~PackStream();
  // parent: HOST arrays of the previous checkpoint's per-page hashes, or null
  // peers_host[0..n_peers): HOST tables of the ranks below this one
  // peers_on_device: the peers' tables are DEVICE arrays already (gathered by a grit_comm), nothing is uploaded
  int begin(const grit_table* parent_host, char* err, size_t errlen, const grit_table* peers_host = nullptr,
            uint32_t n_peers = 0, bool peers_on_device = false);
  // A stream that only analyses (K1 of every chunk into the whole-image arrays, nothing stored or downloaded): the
  // first pass of a ranked directory pack, whose tables the ranks exchange before the second, packing pass.
  void set_analyze_only() { analyze_only_ = true; }
  int wait_chunk(uint64_t i, char* err, size_t errlen);  // the kernels of chunk i are done
  const uint64_t* device_hash() const { return d_hash_; }
  const uint32_t* device_crc() const { return d_crc_; }
  // pinned staging buffers, one per chunk in flight, for callers without pinned memory
  int alloc_staging(bool in, bool out, char* err, size_t errlen);
  // pinned staging buffer for chunk i (valid between submit-ready and the H2D of chunk i)
  void* in_buffer(uint64_t i) { return h_in_[i % kPipeDepth]; }
  void* out_buffer(uint64_t i) { return h_out_[i % kPipeDepth]; }
  // blocks until in_buffer(i) may be overwritten (the H2D of chunk i - depth is done)
  int wait_in_free(uint64_t i, char* err, size_t errlen);
  // enqueue H2D + kernels for chunk i; h_src must be pinned (in_buffer(i) or caller memory)
  int submit(uint64_t i, const void* h_src, uint64_t n_pages, char* err, size_t errlen);
  // wait for chunk i's kernels, learn how many pages it stored, enqueue the D2H into h_dst
  // (pinned: out_buffer(i) or caller memory).  *first_slot = global slot of its first page.
  int collect(uint64_t i, void* h_dst, uint64_t* first_slot, uint64_t* n_out, char* err, size_t errlen);
  // same, when the destination depends on first_slot: dst = base + first_slot * 4096
  int collect_at(uint64_t i, void* h_dst_base, uint64_t capacity_pages, uint64_t* first_slot, uint64_t* n_out,
                 char* err, size_t errlen);
  // blocks until chunk i's D2H has landed
  int wait_out(uint64_t i, char* err, size_t errlen);
  // host arrays (any memory) that collect() fills with each chunk's ref / crc / hash
  void set_outputs(uint32_t* h_ref, uint32_t* h_crc, uint64_t* h_hash) {
    out_ref_ = h_ref;
    out_crc_ = h_crc;
    out_hash_ = h_hash;
  }
  // statistics; syncs everything.  Every chunk must have been collected.
  int finish(grit_pack_stats* stats, char* err, size_t errlen);
  uint64_t chunk_pages() const { return chunk_pages_; }
  double h2d_bytes = 0, d2h_bytes = 0;

 private:
  grit_ctx* ctx_;
  uint64_t total_pages_, chunk_pages_, pages_done_ = 0, slots_done_ = 0;
  void* d_in_[kPipeDepth] = {};
  void* d_out_[kPipeDepth] = {};
  void* h_in_[kPipeDepth] = {};
  void* h_out_[kPipeDepth] = {};
  cudaEvent_t ev_h2d_[kPipeDepth] = {}, ev_kern_[kPipeDepth] = {}, ev_d2h_[kPipeDepth] = {};
  unsigned long long* h_counts_ = nullptr;  // pinned: n_stored per in-flight chunk
  void* h_meta_[kPipeDepth] = {};           // pinned: ref | crc | hash of the in-flight chunks
  uint64_t chunk_pages_n_[kPipeDepth] = {};
  uint32_t *out_ref_ = nullptr, *out_crc_ = nullptr;
  uint64_t* out_hash_ = nullptr;
  uint32_t *d_ref_ = nullptr, *d_crc_ = nullptr;
  uint64_t* d_hash_ = nullptr;
  uint64_t* d_parent_hash_ = nullptr;
  uint32_t* d_parent_crc_ = nullptr;
  uint64_t chunk_page0_[kPipeDepth] = {};
  DedupTable table_{};
  PeerCrcs crcs_{};
  bool timing_started_ = false, analyze_only_ = false;
};

// Where a streamed restore reads the packed pages from and delivers the image to.
// A *_direct pointer is pinned host memory holding everything; otherwise the
// callback moves one chunk through a pinned staging buffer owned by the stream.
struct RestoreIO {
  const void* packed_direct = nullptr;
  std::function<int(uint64_t slot0, uint64_t n_slots, void* dst_pinned)> read_packed;
  void* out_direct = nullptr;
  // called once per output chunk, from writer threads (several at a time, any order), when the chunk has landed in
  // host memory
  std::function<int(uint64_t page0, uint64_t page1, const void* src_pinned)> write_out;
};

// parent_image: the parent's fully restored image, in host memory (uploaded) or,
// with parent_on_device, already on the context's device.
// peers: restored images of lower ranks, already on the device (may be null)
int restore_stream(grit_ctx* ctx, const RestoreIO& io, uint64_t n_slots, const uint32_t* h_ref,
                   const uint32_t* h_crc, uint64_t n_pages, const void* parent_image, uint64_t parent_pages,
                   bool parent_on_device, int verify, uint64_t chunk_pages, grit_restore_stats* stats,
                   double* h2d_bytes, double* d2h_bytes, char* err, size_t errlen, const PeerImages* peers = nullptr,
                   int lanes = 1, int depth = kPipeDepth);

// restore into a device buffer with the packed pages streaming through a small device ring (configs[4])
int restore_to_device(grit_ctx* ctx, const void* h_packed, const std::function<int(uint64_t, uint64_t, void*)>& read_packed,
                      uint64_t n_slots, const uint32_t* h_ref, const uint32_t* h_crc, uint64_t n_pages,
                      const void* d_parent, uint64_t parent_pages, const PeerImages* peers, void* d_out, int verify,
                      uint64_t chunk_pages, grit_restore_stats* stats, double* h2d_total, char* err, size_t errlen);

bool is_pinned(const void* p);

}  // namespace grit
