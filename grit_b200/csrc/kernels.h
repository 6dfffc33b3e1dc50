// Launchers of the sm_100a page kernels (kernels.cu).  Internal to libgritpack;
// the C ABI in api.cu / hostpipe.cu / dirpack.cpp is the public surface.
#pragma once
#include <cuda_runtime.h>

#include <cstdint>

#include "gf2_tables.h"

namespace grit {

constexpr uint32_t kPage = 4096;
constexpr uint32_t kRefPending = 0xFFFFFFFFu;  // K1: non-zero page, not yet resolved
constexpr uint32_t kRefKeep = 0xFFFFFFFEu;     // K2: must be stored, slot not yet assigned
constexpr uint32_t kOriginSelf = 15;
constexpr uint32_t kCompactTile = 256;  // pages per compaction tile (k_compact_wide: four times that)

// table[hash64] = lowest (origin << 26 | page index); origin 0 parent, 1+r peer r, 15 self
// key and owner share a 16-byte slot: one DRAM sector per insert, probe or look-up instead of two
struct TableSlot {
  unsigned long long key;  // 0 = empty
  uint32_t val;            // lowest (origin << 26 | page index) seen for the key
  uint32_t pad;
};
struct DedupTable {
  TableSlot* slots;
  uint64_t mask;  // capacity - 1, capacity a power of two
};

// A per-page array of another image.  Plain arrays are contiguous; arrays that arrived through the cross-rank
// exchange (comm.cu) come in slices of 2^shift pages, every rank's slice s side by side: element i lives at
// base[(i >> shift) * stride + (i & (2^shift - 1))].  shift = 26 (GRIT_MAX_PAGES) makes that base[i].
template <class T>
struct ListView {
  const T* base = nullptr;
  uint64_t stride = 0;
  uint32_t shift = 26;
  __host__ __device__ const T& operator[](uint64_t i) const {
    return base[(i >> shift) * stride + (i & ((1ull << shift) - 1))];
  }
  __host__ __device__ explicit operator bool() const { return base != nullptr; }
};

struct PeerCrcs {
  ListView<uint32_t> parent;
  ListView<uint32_t> peer[8];
};

// the hash lists of the images a pack may refer to, probed against this image's table in one launch
struct ProbeLists {
  ListView<unsigned long long> hash[9];  // [0] parent, [1 + r] rank r
  uint64_t n[9];
};

struct PeerImages {
  const unsigned char* parent;
  uint64_t parent_pages;
  const unsigned char* peer[8];
  uint64_t peer_pages[8];
  uint32_t n_peers;
};

// K4 with only part of the packed image on the device (a restore whose packed input streams through a ring):
// the packed pointer holds slots [slot0, slot0 + n_slots); a duplicate whose owner page lies in front of the launch's
// first page is copied from the restored image itself (`restored`, the whole image on the device) instead of from
// the owner's slot, which may have left the ring.  {0, nullptr}: the whole packed image is there.
struct PackedWindow {
  uint64_t slot0 = 0;
  const uint4* restored = nullptr;
};

// counters written by the kernels, read back by the host after a sync
struct DevCounters {
  unsigned long long kind[5];    // pages per GRIT_KIND_*, filled by compact
  unsigned long long n_stored;   // pages given a slot by the last compact
  unsigned long long overflow;   // pages that did not fit capacity_pages
  unsigned long long n_bad;      // restore: pages that failed
  unsigned long long first_bad;  // restore: lowest such page
  unsigned long long slot_base;       // compact, chained mode: first slot of this launch
  unsigned long long slot_base_next;  // compact: slot_base + pages stored by this launch
  unsigned int ticket;                // compact / fused pack: next tile to analyse
  unsigned int copy_ticket;           // fused pack: next tile whose pages are decided and copied
  unsigned int watermark;             // fused pack: tiles [0, watermark) have entered their hashes in the table
  unsigned int pad;
  // --- owned by the write-through K1 (grit_analyze_pages_through); not touched by the per-call reset,
  // so that a later K3 on the same buffers can see how K1 went
  unsigned long long first_drop;      // lowest zero page K1 has seen (UINT64_MAX: none); storing stops above it
  unsigned int dense_hint;            // 1: the sampled pages held no zero page and K1 wrote the pages through
  unsigned int n_nonkeep;             // K2 since that K1: pages that became DUP / PARENT / PEER references
  unsigned long long through_from;    // first page K1 wrote through (pages in front of it were only analysed)
  unsigned long long dup_mismatch;    // verify_dup_bytes: duplicates by key whose bytes differed (since the context was made)
};

struct LaunchGeom {
  int sm_count;
};

cudaError_t launch_analyze(const LaunchGeom& g, const Gf2Tables* d_tab, const void* d_image,
                           uint64_t n_pages, uint32_t* d_crc, uint64_t* d_hash, uint32_t* d_ref,
                           cudaStream_t s);
cudaError_t launch_table_clear(const LaunchGeom& g, DedupTable t, cudaStream_t s);  // every key 0, every owner 0xFFFFFFFF
cudaError_t launch_table_insert(const LaunchGeom& g, DedupTable t, const uint64_t* d_hash,
                                const uint32_t* d_ref_or_null, uint64_t n, uint32_t origin,
                                uint64_t index_base, cudaStream_t s, bool pending_only = false);
// incremental K2 over a table that holds the parent alone: PARENT references at once, the rest left kRefPending
cudaError_t launch_resolve_parent(const LaunchGeom& g, DedupTable t, uint64_t n_pages, const uint32_t* d_crc,
                                  const uint64_t* d_hash, uint32_t* d_ref, ListView<uint32_t> parent_crc,
                                  DevCounters* d_cnt, cudaStream_t s);
cudaError_t launch_table_probe_lists(const LaunchGeom& g, DedupTable t, const ProbeLists& lists, cudaStream_t s);
cudaError_t launch_resolve(const LaunchGeom& g, DedupTable t, uint64_t page0, uint64_t page1, const uint32_t* d_crc,
                           const uint64_t* d_hash, uint32_t* d_ref, PeerCrcs crcs, DevCounters* d_cnt, cudaStream_t s,
                           bool pending_only = false);
cudaError_t launch_verify_dups(const LaunchGeom& g, const void* d_image, uint64_t n_pages, uint32_t* d_ref, DevCounters* d_cnt,
                               cudaStream_t s);
cudaError_t launch_compact(const LaunchGeom& g, const void* d_image, uint64_t n_pages, uint32_t* d_ref,
                           void* d_packed, uint64_t capacity_pages, uint64_t slot_base, bool chained,
                           unsigned long long* d_tile_status, DevCounters* d_cnt, cudaStream_t s,
                           bool in_place = false);
// Ranks above 0 of a ranked pack: after the first slice's tables have been exchanged, compare this rank's hashes with
// the lower ranks' at the same page index.  Replicated state (weights) sits at the same offsets on every rank: many
// matches mean the image will mostly turn into PEER references and writing it through would be wasted
// (dense_hint := 0); none mean the ranks hold different data (sharded state) and the local sample's verdict stands.
cudaError_t launch_peer_overlap_hint(const unsigned long long* d_hash, const ProbeLists& lists, uint64_t n, DevCounters* d_cnt,
                                     cudaStream_t s);
// sample of the image -> d_cnt->dense_hint: is writing the pages through to their own slots worth it
cudaError_t launch_probe_dense(const void* d_image, uint64_t n_pages, DevCounters* d_cnt, cudaStream_t s);
// K1 for a full image: write-through variant when the hint says dense, plain K1 otherwise.  d_image, the per-page
// arrays and d_packed start at page `page_base` of the image; capacity_pages counts from the image's first page.
cudaError_t launch_analyze_auto(const LaunchGeom& g, const Gf2Tables* d_tab, const void* d_image, uint64_t n_pages,
                                uint32_t* d_crc, uint64_t* d_hash, uint32_t* d_ref, void* d_packed,
                                uint64_t capacity_pages, uint64_t page_base, DevCounters* d_cnt, cudaStream_t s);
// K1+K2+K3 in one pass over a whole device-resident image, pages parked in TMEM between analysis and store.  The
// table must already hold the parent's and the lower ranks' hashes; d_tile_status: one zeroed word per 32-page tile.
cudaError_t launch_pack_onepass(const LaunchGeom& g, const Gf2Tables* d_tab, const void* d_image, uint64_t n_pages,
                                uint32_t* d_crc, uint64_t* d_hash, uint32_t* d_ref, DedupTable t, PeerCrcs crcs, void* d_packed,
                                uint64_t capacity_pages, unsigned long long* d_tile_status, DevCounters* d_cnt, cudaStream_t s);
uint32_t onepass_tile_pages();
cudaError_t launch_restore(const LaunchGeom& g, const Gf2Tables* d_tab, const void* d_packed,
                           uint64_t n_slots, const uint32_t* d_ref, const uint32_t* d_crc,
                           uint64_t n_pages, uint64_t page0, uint64_t page1, PeerImages imgs, void* d_out,
                           int verify, DevCounters* d_cnt, cudaStream_t s, PackedWindow win = PackedWindow{});
cudaError_t launch_synth(const LaunchGeom& g, void* d_image, uint64_t n_pages, uint64_t page_base,
                         uint64_t seed, uint32_t pct_zero, uint32_t pct_dup, uint32_t dup_pool,
                         cudaStream_t s);

uint32_t page_kernel_smem_bytes();
cudaError_t configure_kernels();  // opt in to the large dynamic shared memory, once per device

}  // namespace grit
