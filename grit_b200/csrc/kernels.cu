// Page kernels of gritpack for sm_100a (B200).  HBM-bound byte/integer work:
// 128-bit coalesced loads, one persistent CTA per SM for the CRC kernels (the
// lane-private GF(2) tables fill most of the SM's shared memory), grids sized in
// multiples of the SM count for the rest.  No tensor-core path: nothing here is
// GEMM shaped.
//
// None of this has a counterpart under /root/reference: the reference's data
// mover is io.Copy (pkg/gritagent/copy/copy.go:79).  BASELINE.json north_star
// defines the work; gf2_tables.h explains the CRC algebra.
#include "kernels.h"

#include <type_traits>

namespace grit {

extern __shared__ __align__(1024) unsigned char smem[];

namespace {

constexpr int kWarps = 16;
constexpr int kThreads = kWarps * 32;

// shared memory map of the CRC kernels (bytes)
constexpr uint32_t kSmA = 0;             // A tables, one copy per lane: 128 KiB
constexpr uint32_t kSmT = 128 * 1024;    // T tables, shared: 4 KiB
constexpr uint32_t kSmW = kSmT + 4096;   // W[l] tables, shared: 32 KiB
constexpr uint32_t kSmBytes = kSmW + 32768;

constexpr unsigned long long kGold = 0x9E3779B97F4A7C15ull;
constexpr unsigned long long kFmul = 0xD6E8FEB86659FD93ull;

__host__ __device__ constexpr unsigned long long sm64(unsigned long long x) {
  x += 0x9E3779B97F4A7C15ull;
  unsigned long long z = x;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}

// NH keys: K[n] = hi32(sm64^(n+1)("gritpak1")) | 1, n = 4*jj + i
__host__ __device__ constexpr uint32_t nh_key(int n) {
  unsigned long long s = 0x6772697470616B31ull;
  for (int k = 0; k <= n; k++) s = sm64(s);
  return (uint32_t)(s >> 32) | 1u;
}
template <int N>
struct NhKey {
  static constexpr uint32_t v = nh_key(N);
};

__device__ __forceinline__ unsigned long long fmix(unsigned long long x) {
  x ^= x >> 32;
  x *= kFmul;
  x ^= x >> 32;
  x *= kFmul;
  x ^= x >> 32;
  return x;
}

__device__ __forceinline__ uint4 ldg_stream(const uint4* p) {
  uint4 v;
  // .L2::256B: fetch the 256-byte aligned block around the line into L2 -- the other half is what this lane group
  // reads one step later (or read one step earlier), and DRAM serves 256 contiguous bytes better than 2 x 128
  // (k_analyze 2.81 -> 2.755 ms on the 16 GB image; DRAM traffic unchanged, every line is still fetched once)
  asm volatile("ld.global.nc.L1::no_allocate.L2::256B.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
               : "l"(p));
  return v;
}
__device__ __forceinline__ void stg_stream(uint4* p, const uint4& v) {
  asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y),
               "r"(v.z), "r"(v.w)
               : "memory");
}

__device__ __forceinline__ uint32_t lds32(uint32_t byte_off) {
  return *reinterpret_cast<const uint32_t*>(smem + byte_off);
}

// Expand the GF(2) tables into shared memory.  A[k][e] gets one copy per lane so
// that bank == lane for every lookup: entry e of tables 2r and 2r+1 live in the
// 256-byte row e of 64 KiB region r, table 2r in the first 128 bytes.
__device__ __forceinline__ void load_tables(const Gf2Tables* __restrict__ g) {
  uint32_t* s = reinterpret_cast<uint32_t*>(smem);
  for (uint32_t idx = threadIdx.x; idx < 4u * 256u * 32u; idx += blockDim.x) {
    uint32_t lane = idx & 31, e = (idx >> 5) & 255, k = idx >> 13;
    s[(k >> 1) * 16384 + e * 64 + (k & 1) * 32 + lane] = g->A[k][e];
  }
  const uint32_t* t = &g->T[0][0];
  for (uint32_t idx = threadIdx.x; idx < 1024; idx += blockDim.x) s[kSmT / 4 + idx] = t[idx];
  const uint32_t* w = &g->W[0][0][0];
  for (uint32_t idx = threadIdx.x; idx < 8192; idx += blockDim.x) s[kSmW / 4 + idx] = w[idx];
  __syncthreads();
}

// adv(128, c): one PRMT builds (byte << 8) | lane*4, the table select is the
// LDS immediate, so a lookup costs one ALU op and one conflict-free LDS.
__device__ __forceinline__ uint32_t adv128(uint32_t c, uint32_t lane4) {
  uint32_t a0 = __byte_perm(c, lane4, 0x5504);
  uint32_t a1 = __byte_perm(c, lane4, 0x5514);
  uint32_t a2 = __byte_perm(c, lane4, 0x5524);
  uint32_t a3 = __byte_perm(c, lane4, 0x5534);
  return lds32(kSmA + a0) ^ lds32(kSmA + 128 + a1) ^ lds32(kSmA + 65536 + a2) ^
         lds32(kSmA + 65536 + 128 + a3);
}

__device__ __forceinline__ uint32_t adv_shared(uint32_t d, uint32_t base) {
  return lds32(base + ((d & 0xFF) << 2)) ^ lds32(base + 1024 + ((d >> 6) & 0x3FC)) ^
         lds32(base + 2048 + ((d >> 14) & 0x3FC)) ^ lds32(base + 3072 + ((d >> 22) & 0x3FC));
}

// Per-lane running state over one page: four CRC chains (one per word slot of
// the lane's 16-byte segments), the OR of all words, the NH accumulator.
struct PageAcc {
  uint32_t c0, c1, c2, c3;
  uint32_t z;
  unsigned long long h;
};

template <int JJ, bool kFirst, bool kHash>
__device__ __forceinline__ void absorb(PageAcc& a, const uint4& v, uint32_t lane4) {
  if (kFirst) {
    a.c0 = v.x;
    a.c1 = v.y;
    a.c2 = v.z;
    a.c3 = v.w;
  } else {
    a.c0 = adv128(a.c0, lane4) ^ v.x;
    a.c1 = adv128(a.c1, lane4) ^ v.y;
    a.c2 = adv128(a.c2, lane4) ^ v.z;
    a.c3 = adv128(a.c3, lane4) ^ v.w;
  }
  if (kHash) {
    a.z |= v.x | v.y | v.z | v.w;
    a.h += (unsigned long long)(v.x + NhKey<4 * JJ + 0>::v) * (unsigned long long)(v.y + NhKey<4 * JJ + 1>::v);
    a.h += (unsigned long long)(v.z + NhKey<4 * JJ + 2>::v) * (unsigned long long)(v.w + NhKey<4 * JJ + 3>::v);
  }
}

// compile-time loop: f(std::integral_constant<int, J>) for J = A .. B-1
template <int A, int B, class F>
__device__ __forceinline__ void static_for(F&& f) {
  if constexpr (A < B) {
    f(std::integral_constant<int, A>{});
    static_for<A + 1, B>(f);
  }
}

constexpr int kRing = 16;  // 16-byte loads each lane keeps in flight (half a page per 8-lane group)

template <bool kHint>
__device__ __forceinline__ uint4 ldg_or_zero(const uint4* p, bool on, uint64_t) {
  if (!on) return make_uint4(0, 0, 0, 0);
  return ldg_stream(p);
}

template <bool kHint = false, int KR>
__device__ __forceinline__ void ring_fill(uint4 (&ring)[KR], const uint4* src, bool on, uint64_t pol = 0) {
#pragma unroll
  for (int j = 0; j < KR; j++) ring[j] = ldg_or_zero<kHint>(src + 8 * j, on, pol);
}

// One page per 8-lane group, 32 steps of 16 bytes per lane.  On entry `ring` holds
// steps 0..15 of the current page; every step consumes one slot and immediately
// refills it with the load 16 steps ahead (the second half of this page, then the
// first half of `next`), so 15 loads per lane stay in flight throughout.  On exit
// `ring` holds steps 0..15 of `next`.  sink(j, v) sees every 16-byte value in order.
template <bool kCrc, bool kHash, bool kHint = false, int KR, class Sink>
__device__ __forceinline__ void page_pass(PageAcc& a, uint4 (&ring)[KR], const uint4* cur, bool cur_on,
                                          const uint4* next, bool next_on, uint32_t lane4, Sink&& sink,
                                          uint64_t pol = 0) {
  static_for<0, 32>([&](auto jc) {
    constexpr int j = decltype(jc)::value;
    const uint4 v = ring[j % KR];
    if (j < 32 - KR) ring[j % KR] = ldg_or_zero<kHint>(cur + 8 * (j + KR), cur_on, pol);
    else ring[j % KR] = ldg_or_zero<kHint>(next + 8 * (j + KR - 32), next_on, pol);
    sink(jc, v);
    if (kHash && j % 8 == 0 && j > 0) a.h *= kGold;
    if (kCrc) absorb<j % 8, j == 0, kHash>(a, v, lane4);
  });
}

// Eight 16-byte steps at a time, for kernels that also store (K4): loads and stores go out in
// bursts of eight per lane, which DRAM serves better than a load/store pair per step.
template <bool kFirstBatch, bool kHash>
__device__ __forceinline__ void absorb_batch(PageAcc& a, const uint4 (&v)[8], uint32_t lane4) {
  if (kHash && !kFirstBatch) a.h *= kGold;
  absorb<0, kFirstBatch, kHash>(a, v[0], lane4);
  absorb<1, false, kHash>(a, v[1], lane4);
  absorb<2, false, kHash>(a, v[2], lane4);
  absorb<3, false, kHash>(a, v[3], lane4);
  absorb<4, false, kHash>(a, v[4], lane4);
  absorb<5, false, kHash>(a, v[5], lane4);
  absorb<6, false, kHash>(a, v[6], lane4);
  absorb<7, false, kHash>(a, v[7], lane4);
}

__device__ __forceinline__ void load_batch(uint4 (&v)[8], const uint4* src, bool on) {
#pragma unroll
  for (int j = 0; j < 8; j++) v[j] = on ? ldg_stream(src + 8 * j) : make_uint4(0, 0, 0, 0);
}

// Fold the lane's four chains, weight by the lane's position, XOR over the
// 8-lane group: the CRC-32C of the page (valid in every lane of the group).
__device__ __forceinline__ uint32_t finish_crc(const PageAcc& a, uint32_t l) {
  uint32_t d = a.c0;
  d = adv_shared(d, kSmT) ^ a.c1;
  d = adv_shared(d, kSmT) ^ a.c2;
  d = adv_shared(d, kSmT) ^ a.c3;
  uint32_t r = adv_shared(d, kSmW + l * 4096);
  r ^= __shfl_xor_sync(0xFFFFFFFFu, r, 1);
  r ^= __shfl_xor_sync(0xFFFFFFFFu, r, 2);
  r ^= __shfl_xor_sync(0xFFFFFFFFu, r, 4);
  return r ^ kCrcZeroPage;
}

__device__ __forceinline__ unsigned long long finish_hash(const PageAcc& a, uint32_t l) {
  unsigned long long v = fmix(a.h + (unsigned long long)(l + 1) * kFmul);
  v += __shfl_xor_sync(0xFFFFFFFFu, v, 1);
  v += __shfl_xor_sync(0xFFFFFFFFu, v, 2);
  v += __shfl_xor_sync(0xFFFFFFFFu, v, 4);
  v = fmix(v ^ kGold);
  return v ? v : 1ull;
}

#ifndef GRIT_K1_WARPS
#define GRIT_K1_WARPS 16
#endif
#ifndef GRIT_K1_RING
#define GRIT_K1_RING 16
#endif
constexpr int kK1Warps = GRIT_K1_WARPS, kK1Threads = kK1Warps * 32, kK1Ring = GRIT_K1_RING;
// ------------------------------------------------------------------ K1 analyze
// A warp takes four consecutive pages per step; lanes 8g..8g+7 own page g.
__global__ void __launch_bounds__(kK1Threads, 1)
k_analyze(const Gf2Tables* __restrict__ tabs, const uint4* __restrict__ image, uint64_t n_pages,
          uint32_t* __restrict__ crc, unsigned long long* __restrict__ hash, uint32_t* __restrict__ ref,
          const DevCounters* __restrict__ cnt, int want_hint) {
  if (want_hint >= 0 && (int)cnt->dense_hint != want_hint) return;  // the write-through variant runs instead
  load_tables(tabs);
  const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const uint32_t g = lane >> 3, l = lane & 7, lane4 = lane * 4;
  const uint64_t n_groups = (n_pages + 3) / 4;
  const uint64_t stride = (uint64_t)gridDim.x * kK1Warps;
  uint64_t grp = (uint64_t)blockIdx.x * kK1Warps + warp;
  if (grp >= n_groups) return;

  uint4 ring[kK1Ring];
  uint64_t p = 4 * grp + g;
  const uint4* src = image + p * 256 + l;
  ring_fill(ring, src, p < n_pages);
  for (; grp < n_groups; grp += stride) {
    PageAcc a;
    a.z = 0;
    a.h = 0;
    const bool on = p < n_pages;
    const uint64_t p_next = 4 * (grp + stride) + g;  // this warp's next group: prefetched across the boundary
    const uint4* src_next = image + p_next * 256 + l;
    const bool next_on = (grp + stride < n_groups) && p_next < n_pages;
    page_pass<true, true>(a, ring, src, on, src_next, next_on, lane4, [](auto, const uint4&) {});

    const uint32_t page_crc = finish_crc(a, l);
    uint32_t z = a.z;
    z |= __shfl_xor_sync(0xFFFFFFFFu, z, 1);
    z |= __shfl_xor_sync(0xFFFFFFFFu, z, 2);
    z |= __shfl_xor_sync(0xFFFFFFFFu, z, 4);
    const unsigned long long page_hash = finish_hash(a, l);
    if (l == 0 && on) {
      crc[p] = page_crc;
      hash[p] = page_hash;
      ref[p] = z ? kRefPending : 0u;
    }
    p = p_next;
    src = src_next;
  }
}

// ------------------------------------------------------------------ K1 with write-through
// For images that need (almost) no compaction -- model weights and optimizer state have no zero
// pages and no duplicates, so slot(p) == p for every page up to the first page that is dropped --
// every page is stored to packed[p] while it is analysed, and K3 finds those pages in place and
// only moves what lies behind the first dropped page: N read + N written of DRAM traffic, the
// algorithmic minimum, instead of N + 2N.
// The stores go out in bursts: a page's values are parked in tensor memory (tcgen05.st) while it is analysed and written
// afterwards, eight 16-byte steps per lane back to back (tcgen05.ld + STG), so that DRAM sees 1 KiB of writes per page
// at a time instead of a 128-byte store between every two 128-byte loads: 5.36 -> 5.19 ms on the 16 GB dense image
// (92 -> 95 % of the measured HBM peak; -DGRIT_K1T_DIRECT builds the direct-store form).  Burst 4 / 8 / 16: 5.21 / 5.19-5.22
// / 5.21-5.26 ms.
// CRC + hash64 + a register load ring + stores do not fit 128 registers, so here the load ring
// lives in shared memory: each lane's 16-byte loads are cp.async copies (global -> shared, no
// register) into a private slot that the same lane reads back with one LDS.128 kSmRing steps later.
// Depth measured on the 16 GB dense image (pack, ms): 2: 7.49, 3: 6.40, 4: 5.94-6.10, 5: 5.79-5.83, 6: 5.93-5.97, 7: 6.10-6.13,
// 8: slower again.  Five copies per lane (40 KB per SM) cover the read latency; more only crowd the stores out.
#ifndef GRIT_K1T_RING
#define GRIT_K1T_RING 5
#endif
#ifndef GRIT_K1T_BURST
#define GRIT_K1T_BURST 8
#endif
constexpr int kSmRing = GRIT_K1T_RING;            // cp.async copies each lane keeps in flight
constexpr uint32_t kSmRingBase = kSmBytes;       // behind the tables
constexpr uint32_t kSmBytesWT = kSmBytes + kSmRing * kThreads * 16;  // 208 896 B of the 232 448 B limit

__device__ __forceinline__ void cp_async16(uint32_t smem_addr, const void* gptr) {
  asm volatile("cp.async.cg.shared.global.L2::256B [%0], [%1], 16;" ::"r"(smem_addr), "l"(gptr) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}
__device__ __forceinline__ uint4 lds128(uint32_t smem_addr) {
  uint4 v;
  asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(smem_addr) : "memory");
  return v;
}

__global__ void __launch_bounds__(kThreads, 1)
k_analyze_through(const Gf2Tables* __restrict__ tabs, const uint4* __restrict__ image, uint64_t n_pages,
                  uint32_t* __restrict__ crc, unsigned long long* __restrict__ hash, uint32_t* __restrict__ ref,
                  uint4* __restrict__ packed, uint64_t capacity_pages, uint64_t page_base, DevCounters* __restrict__ cnt,
                  int want_hint) {
  // image, crc, hash, ref and packed all start at page `page_base` of the whole image (a slice of a ranked pack);
  // first_drop and capacity_pages count pages of the whole image
  if (want_hint >= 0 && (int)cnt->dense_hint != want_hint) return;  // the other K1 variant runs instead
  load_tables(tabs);
  const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const uint32_t g = lane >> 3, l = lane & 7, lane4 = lane * 4;
  const uint64_t n_groups = (n_pages + 3) / 4;
  const uint64_t stride = (uint64_t)gridDim.x * kWarps;
  uint64_t grp = (uint64_t)blockIdx.x * kWarps + warp;
#ifndef GRIT_K1T_DIRECT
  // Write bursts: the values of a page are parked in tensor memory while the page is analysed and stored afterwards,
  // eight 16-byte steps per lane back to back, so that DRAM sees 1 KiB write bursts per page instead of a 128-byte
  // store between every two 128-byte loads.  Warp w parks its four pages in quadrant w % 4, columns (w / 4) * 128.
  __shared__ uint32_t s_tmem;
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"((uint32_t)__cvta_generic_to_shared(&s_tmem))
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem_all = s_tmem;
  const uint32_t tcol = tmem_all + (((warp & 3) * 32) << 16) + (warp >> 2) * 128;
  if (grp < n_groups) {
#else
  if (grp >= n_groups) return;
#endif
  const uint32_t ring0 = (uint32_t)__cvta_generic_to_shared(smem) + kSmRingBase + threadIdx.x * 16;  // slot s at + s * 8192
  const uint64_t last = n_pages - 1;

  // every step issues exactly one copy and one commit, so "at most kSmRing-1 groups pending" always
  // means that the oldest slot has landed; pages past the end re-read the last page (result dropped)
  uint64_t p = 4 * grp + g;
  const uint4* src = image + (p < last ? p : last) * 256 + l;
#pragma unroll
  for (int j = 0; j < kSmRing; j++) {
    cp_async16(ring0 + j * (kThreads * 16), src + 8 * j);
    cp_async_commit();
  }
  uint32_t slot = 0;
  for (; grp < n_groups; grp += stride) {
    PageAcc a;
    a.z = 0;
    a.h = 0;
    const bool on = p < n_pages;
    const unsigned long long first_drop = __ldcg(&cnt->first_drop);
    const bool through = on && page_base + p < capacity_pages && page_base + p < first_drop;
    uint4* d = packed + p * 256 + l;
    const uint64_t p_next = 4 * (grp + stride) + g;
    const uint64_t p_load = grp + stride < n_groups ? (p_next < last ? p_next : last) : (p < last ? p : last);
    const uint4* src_next = image + p_load * 256 + l;
    static_for<0, 32>([&](auto jc) {
      constexpr int j = decltype(jc)::value;
      cp_async_wait<kSmRing - 1>();
      const uint32_t sa = ring0 + slot * (kThreads * 16);
      const uint4 v = lds128(sa);
#ifndef GRIT_K1T_DIRECT
      asm volatile("tcgen05.st.sync.aligned.32x32b.x4.b32 [%0], {%1, %2, %3, %4};" ::"r"(tcol + 4 * j), "r"(v.x), "r"(v.y),
                   "r"(v.z), "r"(v.w)
                   : "memory");
#else
      if (through) stg_stream(d + 8 * j, v);
#endif
      cp_async16(sa, j + kSmRing < 32 ? src + 8 * (j + kSmRing) : src_next + 8 * (j + kSmRing - 32));
      cp_async_commit();
      slot = slot + 1 == kSmRing ? 0 : slot + 1;
      if (j % 8 == 0 && j > 0) a.h *= kGold;
      absorb<j % 8, j == 0, true>(a, v, lane4);
    });
#ifndef GRIT_K1T_DIRECT
    asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
    if (__any_sync(0xFFFFFFFFu, through)) {
#pragma unroll 1
      for (int jb = 0; jb < 32; jb += GRIT_K1T_BURST) {
        uint4 w[GRIT_K1T_BURST];
#pragma unroll
        for (int j = 0; j < GRIT_K1T_BURST; j++)
          asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0, %1, %2, %3}, [%4];"
                       : "=r"(w[j].x), "=r"(w[j].y), "=r"(w[j].z), "=r"(w[j].w)
                       : "r"(tcol + 4 * (jb + j))
                       : "memory");
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
        if (through) {
#pragma unroll
          for (int j = 0; j < GRIT_K1T_BURST; j++) stg_stream(d + 8 * (jb + j), w[j]);
        }
      }
    }
#endif
    const uint32_t page_crc = finish_crc(a, l);
    uint32_t z = a.z;
    z |= __shfl_xor_sync(0xFFFFFFFFu, z, 1);
    z |= __shfl_xor_sync(0xFFFFFFFFu, z, 2);
    z |= __shfl_xor_sync(0xFFFFFFFFu, z, 4);
    const unsigned long long page_hash = finish_hash(a, l);
    if (l == 0 && on) {
      crc[p] = page_crc;
      hash[p] = page_hash;
      ref[p] = z ? kRefPending : 0u;
      if (!z && page_base + p < first_drop) atomicMin(&cnt->first_drop, (unsigned long long)(page_base + p));
    }
    p = p_next;
    src = src_next;
  }
  cp_async_wait<0>();
#ifndef GRIT_K1T_DIRECT
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem_all) : "memory");
#endif
}

// Decides, from 4096 evenly spread pages (128 bytes of each), whether K1 should write the pages through to their own
// slots.  That pays when a long prefix of the image stays in place: no duplicate anywhere near the front (a page that
// is dropped shifts every later slot, and K3 then moves what K1' already wrote) and no zero page in the first quarter
// (K1' stops storing at the first zero page it meets; the rest is K3's).  So: the lowest sampled zero page must lie
// beyond a quarter of the image, and fewer than four samples may repeat another sample's 128 bytes.  One block; sets
// cnt->dense_hint for the two K1 variants and K3.
constexpr uint32_t kProbeSamples = 4096, kProbeSlots = 8192;
__global__ void __launch_bounds__(1024)
k_probe_dense(const uint4* __restrict__ image, uint64_t n_pages, DevCounters* __restrict__ cnt) {
  __shared__ uint32_t s_sig[kProbeSlots];
  __shared__ unsigned long long s_first_zero;
  __shared__ uint32_t s_repeats;
  for (uint32_t i = threadIdx.x; i < kProbeSlots; i += blockDim.x) s_sig[i] = 0;
  if (threadIdx.x == 0) {
    s_first_zero = ~0ull;
    s_repeats = 0;
  }
  __syncthreads();
  // four samples per thread, their 32 loads issued before any is used (one block: latency is all this kernel costs)
  constexpr int kPer = kProbeSamples / 1024;
  uint4 v[kPer][8];
  uint64_t pg[kPer];
#pragma unroll
  for (int q = 0; q < kPer; q++) {
    const uint32_t i = threadIdx.x + 1024 * q;
    pg[q] = (uint64_t)i * n_pages / kProbeSamples;
    const uint4* src = image + pg[q] * 256 + (i & 31);  // eight 16-byte pieces, 512 bytes apart; the offset varies by sample
#pragma unroll
    for (int k = 0; k < 8; k++) v[q][k] = __ldg(src + 32 * k);
  }
#pragma unroll
  for (int q = 0; q < kPer; q++) {
    const uint32_t i = threadIdx.x + 1024 * q;
    if (i && pg[q] == (uint64_t)(i - 1) * n_pages / kProbeSamples) continue;  // fewer pages than samples
    uint32_t o = 0;
    unsigned long long sig = kGold;
#pragma unroll
    for (int k = 0; k < 8; k++) {
      o |= v[q][k].x | v[q][k].y | v[q][k].z | v[q][k].w;
      sig = (sig ^ ((unsigned long long)v[q][k].y << 32 | v[q][k].x)) * kFmul;
      sig = (sig ^ ((unsigned long long)v[q][k].w << 32 | v[q][k].z)) * kFmul;
    }
    if (o == 0) {  // reads as a zero page
      atomicMin(&s_first_zero, (unsigned long long)pg[q]);
      continue;
    }
    // the sampled bytes differ with the offset, so two samples only compare equal when they read the same offset of
    // equal pages: signature plus offset
    uint32_t g = (uint32_t)(sig >> 32) ^ ((i & 31) * 0x9E3779B9u);
    g = g ? g : 1u;
    for (uint32_t sl = g & (kProbeSlots - 1);; sl = (sl + 1) & (kProbeSlots - 1)) {
      const uint32_t was = atomicCAS(&s_sig[sl], 0u, g);
      if (was == 0u) break;
      if (was == g) {
        atomicAdd(&s_repeats, 1u);
        break;
      }
    }
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    const bool dense = s_first_zero >= n_pages / 4 && s_repeats < 4;
    cnt->dense_hint = dense ? 1u : 0u;
  }
}

// ------------------------------------------------------------------ K2 dedup
__device__ __forceinline__ uint64_t table_home(unsigned long long h, uint64_t mask) {
  return ((h * kGold) >> 20) & mask;
}

__global__ void __launch_bounds__(256)
k_table_clear(DedupTable t) {
  uint4* w = reinterpret_cast<uint4*>(t.slots);
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i <= t.mask; i += (uint64_t)gridDim.x * blockDim.x)
    stg_stream(w + i, make_uint4(0u, 0u, 0xFFFFFFFFu, 0u));
}

// One item per thread and step: several chains in flight per thread (batches of 2, 4, 8 were tried) only slow these
// kernels down -- at full occupancy the memory system is already saturated with scattered 32-byte accesses.
__global__ void __launch_bounds__(256)
k_table_insert(DedupTable t, const unsigned long long* __restrict__ hash, const uint32_t* __restrict__ ref,
               uint64_t n, uint32_t origin, uint64_t index_base, bool pending_only) {
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (uint64_t)gridDim.x * blockDim.x) {
    if (ref && (pending_only ? ref[i] != kRefPending : ref[i] == 0u)) continue;  // zero pages never enter the table
    const unsigned long long h = hash[i];
    if (h == 0ull) continue;  // hash64 is never 0: a damaged hashes.bin must not claim the empty key
    const uint32_t v = origin << 26 | (uint32_t)(index_base + i);
    for (uint64_t s = table_home(h, t.mask);; s = (s + 1) & t.mask) {
      unsigned long long k = t.slots[s].key;
      if (k == 0ull) k = atomicCAS(&t.slots[s].key, 0ull, h);
      if (k == 0ull || k == h) {
        // a much-repeated page (a pool of identical buffers) sends hundreds of pages to one slot: only a value
        // that lowers the owner is worth an atomic, and the plain read costs a hit in L2
        if (k == 0ull || __ldcg(&t.slots[s].val) > v) atomicMin(&t.slots[s].val, v);
        break;
      }
    }
  }
}

// Lowers the owner of every key of THIS image's table that one of the other images (parent, lower ranks) holds too.
__global__ void __launch_bounds__(256)
k_table_probe_lists(DedupTable t, ProbeLists lists) {
  const uint32_t origin = blockIdx.y;
  const ListView<unsigned long long> hash = lists.hash[origin];
  const uint64_t n = lists.n[origin];
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (uint64_t)gridDim.x * blockDim.x) {
    const unsigned long long h = hash[i];
    if (h == 0ull) continue;  // hash64 is never 0: padding of a gathered slice, or a damaged hashes.bin
    for (uint64_t s = table_home(h, t.mask);; s = (s + 1) & t.mask) {
      const unsigned long long k = t.slots[s].key;
      if (k == h) {
        const uint32_t v = origin << 26 | (uint32_t)i;
        if (__ldcg(&t.slots[s].val) > v) atomicMin(&t.slots[s].val, v);
        break;
      }
      if (k == 0ull) break;
    }
  }
}

// K2 of an incremental pack (a parent and no peers).  The table holds the PARENT's pages alone, and a page of this
// image that finds its hash there is settled after that one read: the parent is origin 0, so its page owns the hash,
// and the page here is a PARENT reference when the CRCs agree and a stored page when they do not (pages never refer
// to each other past a differing owner).  Only the pages whose hash the parent does not hold are left kRefPending;
// they go into a second table (k_table_insert pending_only) and through k_resolve(pending_only), which is the
// general K2 on the few dirtied pages.  An unchanged page costs one table read instead of an insert, a probe and a
// lookup.
__global__ void __launch_bounds__(256)
k_resolve_parent(DedupTable t, uint64_t n_pages, const uint32_t* __restrict__ crc,
                 const unsigned long long* __restrict__ hash, uint32_t* __restrict__ ref, ListView<uint32_t> parent_crc,
                 DevCounters* __restrict__ cnt) {
  for (uint64_t p = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; p < n_pages; p += (uint64_t)gridDim.x * blockDim.x) {
    if (ref[p] == 0u) continue;
    const unsigned long long h = hash[p];
    uint32_t out = h ? kRefPending : kRefKeep;  // K1 never gives hash 0 (the empty key); a caller's own 0 is stored
    for (uint64_t s = table_home(h, t.mask); h; s = (s + 1) & t.mask) {
      const ulonglong2 slot = *reinterpret_cast<const ulonglong2*>(&t.slots[s]);  // key and owner in one access
      if (slot.x == h) {
        const uint32_t idx = (uint32_t)slot.y & ((1u << 26) - 1);
        out = parent_crc[idx] == crc[p] ? (3u << 29 | idx) : kRefKeep;
        break;
      }
      if (slot.x == 0ull) break;
    }
    ref[p] = out;
    if (out >> 29 == 3u) atomicAdd(&cnt->n_nonkeep, 1u);
  }
}

__global__ void __launch_bounds__(256)
k_resolve(DedupTable t, uint64_t page0, uint64_t page1, const uint32_t* __restrict__ crc,
          const unsigned long long* __restrict__ hash, uint32_t* __restrict__ ref, PeerCrcs crcs, DevCounters* __restrict__ cnt,
          bool pending_only) {
  for (uint64_t p = page0 + (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; p < page1;
       p += (uint64_t)gridDim.x * blockDim.x) {
    if (pending_only ? ref[p] != kRefPending : ref[p] == 0u) continue;
    const unsigned long long h = hash[p];
    if (h == 0ull) {  // K1 never gives hash 0 (the empty key, not inserted); a caller's own 0 is stored
      ref[p] = kRefKeep;
      continue;
    }
    uint64_t s = table_home(h, t.mask);
    while (t.slots[s].key != h) s = (s + 1) & t.mask;  // inserted by k_table_insert
    const uint32_t v = t.slots[s].val, origin = v >> 26, idx = v & ((1u << 26) - 1);
    uint32_t out = kRefKeep;
    if (!(origin == kOriginSelf && idx == (uint32_t)p)) {
      const uint32_t ocrc = origin == 0             ? crcs.parent[idx]
                            : origin == kOriginSelf ? crc[idx]
                                                    : crcs.peer[origin - 1][idx];
      if (ocrc == crc[p]) {
        out = origin == 0             ? (3u << 29 | idx)
              : origin == kOriginSelf ? (2u << 29 | idx)
                                      : (4u << 29 | (origin - 1) << 26 | idx);
      }
    }
    ref[p] = out;
    if (out != kRefKeep) atomicAdd(&cnt->n_nonkeep, 1u);  // rare in the images where K3 asks (dense ones)
  }
}

// Optional check between K2 and K3 (grit_ctx_set_option "verify_dup_bytes"): a page that K2 turned into a reference to
// an earlier page of the same image on the strength of its 96-bit key (hash64, CRC-32C) is compared with that page
// byte for byte; if they differ it is stored after all.  One warp per duplicate, 2 x 4 KiB read.
__global__ void __launch_bounds__(256)
k_verify_dups(const uint4* __restrict__ image, uint64_t n_pages, uint32_t* __restrict__ ref, DevCounters* __restrict__ cnt) {
  const uint32_t lane = threadIdx.x & 31;
  const uint64_t warps = (uint64_t)gridDim.x * (blockDim.x >> 5);
  for (uint64_t p = (uint64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); p < n_pages; p += warps) {
    const uint32_t r = ref[p];
    if ((r >> 29) != 2u) continue;
    const uint64_t q = r & 0x1FFFFFFFu;
    const uint4 *a = image + p * 256 + lane, *b = image + q * 256 + lane;
    uint32_t diff = 0;
#pragma unroll
    for (int j = 0; j < 8; j++) {
      const uint4 x = ldg_stream(a + 32 * j), y = ldg_stream(b + 32 * j);
      diff |= (x.x ^ y.x) | (x.y ^ y.y) | (x.z ^ y.z) | (x.w ^ y.w);
    }
    if (__any_sync(0xFFFFFFFFu, diff != 0) && lane == 0) {
      ref[p] = kRefKeep;
      atomicAdd(&cnt->dup_mismatch, 1ull);
    }
  }
}

// ------------------------------------------------------------------ K3 compact
// Single pass: per 256-page tile, ballot-count the pages to keep, get the tile's
// exclusive prefix by decoupled look-back over the tile status words, assign
// slots in page order, then all warps copy the kept pages with 128-bit accesses.
constexpr unsigned long long kFlagAgg = 1ull << 62, kFlagIncl = 2ull << 62;

__device__ __forceinline__ unsigned long long ld_status(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_status(unsigned long long* p, unsigned long long v) {
  asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

// kPer pages per thread, 256 * kPer pages per tile.  One page per thread suits an image that stores most of its pages:
// a tile is then hundreds of page copies, and short tiles keep the last blocks' tail short.  In an image of mostly
// references (an incremental checkpoint) a 256-page tile is a dozen copies behind a fixed chain of L2 round trips
// (ticket, ref words, look-back), and four pages per thread cut the number of such chains by four.
template <int kPer>
__device__ __forceinline__ void
compact_body(const uint4* __restrict__ image, uint64_t n_pages, uint32_t* __restrict__ ref,
             uint4* __restrict__ packed, uint64_t capacity_pages, uint64_t slot_base, int chained, int in_place,
             unsigned long long* __restrict__ status, DevCounters* __restrict__ cnt) {
  constexpr uint32_t kTile = 256u * kPer;
  __shared__ uint32_t s_tile, s_warp_cnt[8], s_excl;
  __shared__ uint2 s_kind[8];
  in_place = in_place && cnt->dense_hint;  // the write-through K1 ran: pages in front of the first drop are in place
  // Nothing was dropped at all (no zero page met, no reference made): every page is stored, in its own slot, where the
  // write-through K1 already put it.  No scan, no look-back: one streaming pass over ref.
  if (in_place && cnt->through_from == 0 && cnt->first_drop == ~0ull && cnt->n_nonkeep == 0u) {
    for (uint64_t p = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; p < n_pages; p += (uint64_t)gridDim.x * blockDim.x)
      ref[p] = 1u << 29 | (uint32_t)p;
    if (blockIdx.x == 0 && threadIdx.x == 0) {
      cnt->kind[1] = n_pages;
      cnt->n_stored = n_pages;
      cnt->slot_base_next = n_pages;
      if (n_pages > capacity_pages) cnt->overflow = n_pages - capacity_pages;
    }
    return;
  }
  __shared__ uint16_t s_list[kTile];
  const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const uint32_t n_tiles = (uint32_t)((n_pages + kTile - 1) / kTile);
  // chained (streaming) mode: slots continue from the previous launch's total, which only
  // the device knows; the packed buffer then holds just this launch's pages, from index 0.
  const uint64_t ref_base = chained ? cnt->slot_base : slot_base;
  const uint64_t out_base = chained ? 0 : slot_base;

  // The page statistics are summed per block and added once at its end: the counters share a line with the ticket,
  // and a handful of atomics per tile on that one line would pace an image of short tiles (mostly references).
  uint32_t sum_zero = 0, sum_keep = 0, sum_dup = 0, sum_par = 0, sum_peer = 0;
  for (;;) {
    if (threadIdx.x == 0) s_tile = atomicAdd(&cnt->ticket, 1u);
    __syncthreads();
    const uint32_t tile = s_tile;
    if (tile >= n_tiles) {
      if (threadIdx.x == 0) {
        if (sum_zero) atomicAdd(&cnt->kind[0], (unsigned long long)sum_zero);
        if (sum_keep) atomicAdd(&cnt->kind[1], (unsigned long long)sum_keep);
        if (sum_dup) atomicAdd(&cnt->kind[2], (unsigned long long)sum_dup);
        if (sum_par) atomicAdd(&cnt->kind[3], (unsigned long long)sum_par);
        if (sum_peer) atomicAdd(&cnt->kind[4], (unsigned long long)sum_peer);
      }
      return;
    }
    // kPer consecutive pages per thread
    const uint64_t p0 = (uint64_t)tile * kTile + (uint32_t)kPer * threadIdx.x;
    uint32_t keepm = 0, pack_a = 0, pack_b = 0;  // bit j: page p0 + j is stored; zero | dup << 16; parent | peer << 16
#pragma unroll
    for (int j = 0; j < kPer; j++) {
      const bool in = p0 + j < n_pages;
      const uint32_t r = in ? ref[p0 + j] : 0u;
      const bool keep = r == kRefKeep;
      keepm |= keep ? 1u << j : 0u;
      const uint32_t kind = keep ? 1u : (in ? r >> 29 : 7u);
      pack_a += (kind == 0u ? 1u : 0u) + (kind == 2u ? 1u << 16 : 0u);
      pack_b += (kind == 3u ? 1u : 0u) + (kind == 4u ? 1u << 16 : 0u);
    }
    const uint32_t mine = __popc(keepm);
    uint32_t inc = mine;  // stored pages of this warp up to and including this thread
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t t = __shfl_up_sync(0xFFFFFFFFu, inc, o);
      inc += lane >= (uint32_t)o ? t : 0u;
    }
    pack_a = __reduce_add_sync(0xFFFFFFFFu, pack_a);
    pack_b = __reduce_add_sync(0xFFFFFFFFu, pack_b);
    if (lane == 31) s_warp_cnt[warp] = inc;
    if (lane == 0) s_kind[warp] = make_uint2(pack_a, pack_b);
    __syncthreads();

    uint32_t before = 0, total = 0;
#pragma unroll
    for (int w = 0; w < 8; w++) {
      const uint32_t c = s_warp_cnt[w];
      before += w < (int)warp ? c : 0u;
      total += c;
    }
    if (threadIdx.x == 0) {
      uint32_t ka = 0, kb = 0;  // a block's sums stay below 2^16 per field
#pragma unroll
      for (int w = 0; w < 8; w++) {
        ka += s_kind[w].x;
        kb += s_kind[w].y;
      }
      sum_zero += ka & 0xFFFFu;
      sum_dup += ka >> 16;
      sum_par += kb & 0xFFFFu;
      sum_peer += kb >> 16;
      sum_keep += total;
    }

    if (warp == 0) {
      if (lane == 0) st_status(&status[tile], (tile == 0 ? kFlagIncl : kFlagAgg) | total);
      uint32_t excl = 0;
      if (tile > 0) {
        int64_t look = (int64_t)tile - 1;
        for (;;) {
          const int64_t idx = look - lane;
          unsigned long long st;
          do {
            st = idx >= 0 ? ld_status(&status[idx]) : kFlagIncl;
          } while (__any_sync(0xFFFFFFFFu, (st >> 62) == 0));
          const uint32_t incl = __ballot_sync(0xFFFFFFFFu, (st >> 62) == 2);
          const uint32_t upto = incl ? (uint32_t)__ffs(incl) - 1u : 31u;
          uint32_t v = lane <= upto ? (uint32_t)st : 0u;
#pragma unroll
          for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xFFFFFFFFu, v, o);
          excl += v;
          if (incl) break;
          look -= 32;
        }
        if (lane == 0) st_status(&status[tile], kFlagIncl | (excl + total));
      }
      if (lane == 0) {
        s_excl = excl;
        if (tile == n_tiles - 1) {
          cnt->n_stored = excl + total;
          cnt->slot_base_next = ref_base + excl + total;
        }
      }
    }
    const uint32_t rank0 = before + inc - mine;
    {
      uint32_t k = rank0;
#pragma unroll
      for (int j = 0; j < kPer; j++)
        if (keepm >> j & 1u) s_list[k++] = (uint16_t)((uint32_t)kPer * threadIdx.x + j);
    }
    __syncthreads();
    const uint64_t first_slot = out_base + s_excl;
    {
      uint32_t k = rank0;
#pragma unroll
      for (int j = 0; j < kPer; j++)
        if (keepm >> j & 1u) {
          ref[p0 + j] = 1u << 29 | (uint32_t)(ref_base + s_excl + k);
          if (first_slot + k >= capacity_pages) atomicAdd(&cnt->overflow, 1ull);
          k++;
        }
    }
    // after a write-through K1 a tile whose pages all keep their own index is already in place
    const uint64_t tile_base = (uint64_t)tile * kTile;
    const uint32_t tile_pages = (uint32_t)(n_pages - tile_base < kTile ? n_pages - tile_base : kTile);
    const bool tile_in_place = in_place && first_slot == tile_base && total == tile_pages && tile_base >= cnt->through_from;
    // copy: one page per warp per step, two steps in flight
    for (uint32_t k = tile_in_place ? total : warp; k < total; k += 16) {
      const uint32_t k2 = k + 8;
      const bool two = k2 < total;
      const uint64_t sa = first_slot + k, sb = first_slot + k2;
      const uint4* a = image + ((uint64_t)tile * kTile + s_list[k]) * 256 + lane;
      const uint4* b = image + ((uint64_t)tile * kTile + s_list[two ? k2 : k]) * 256 + lane;
      uint4 va[8], vb[8];
#pragma unroll
      for (int j = 0; j < 8; j++) va[j] = ldg_stream(a + 32 * j);
      if (two) {
#pragma unroll
        for (int j = 0; j < 8; j++) vb[j] = ldg_stream(b + 32 * j);
      }
      if (sa < capacity_pages) {
#pragma unroll
        for (int j = 0; j < 8; j++) stg_stream(packed + sa * 256 + lane + 32 * j, va[j]);
      }
      if (two && sb < capacity_pages) {
#pragma unroll
        for (int j = 0; j < 8; j++) stg_stream(packed + sb * 256 + lane + 32 * j, vb[j]);
      }
    }
    __syncthreads();  // s_tile, s_list reused by the next tile
  }
}

// K2 counted the references it made.  More than half of the pages: the wide tiles take the image (k_compact_wide),
// else the short ones; the device knows, so both kernels are launched and one of them returns at once.
__device__ __forceinline__ bool compact_is_sparse(const DevCounters* cnt, uint64_t n_pages) {
  return (uint64_t)cnt->n_nonkeep * 2 > n_pages;
}

__global__ void __launch_bounds__(256)
k_compact(const uint4* __restrict__ image, uint64_t n_pages, uint32_t* __restrict__ ref,
          uint4* __restrict__ packed, uint64_t capacity_pages, uint64_t slot_base, int chained, int in_place,
          unsigned long long* __restrict__ status, DevCounters* __restrict__ cnt, int paired) {
  if (paired && compact_is_sparse(cnt, n_pages)) return;
  compact_body<1>(image, n_pages, ref, packed, capacity_pages, slot_base, chained, in_place, status, cnt);
}

__global__ void __launch_bounds__(256)
k_compact_wide(const uint4* __restrict__ image, uint64_t n_pages, uint32_t* __restrict__ ref,
               uint4* __restrict__ packed, uint64_t capacity_pages, uint64_t slot_base, int in_place,
               unsigned long long* __restrict__ status, DevCounters* __restrict__ cnt) {
  if (!compact_is_sparse(cnt, n_pages)) return;
  compact_body<4>(image, n_pages, ref, packed, capacity_pages, slot_base, 0, in_place, status, cnt);
}

// ------------------------------------------------------------------ K1+K2+K3 in one pass, pages parked in TMEM
// The two-pass pack reads every stored page twice (K1, then K3): N + 2U of DRAM traffic for N + U of work.  Here a
// page is read once, and while its CRC / hash64 are computed its 16-byte values are parked in tensor memory -- the
// 256 KiB per SM that only tcgen05 instructions reach and that nothing else in this library uses (the lane-private
// CRC tables leave no room in shared memory).  Once the page's fate is known it is written out of TMEM to its slot,
// or dropped.  DRAM sees N read + U written.
//
// A CTA (one per SM) owns tiles t = k * gridDim + blockIdx of 32 pages.  Eight compute warps analyse a tile exactly
// like k_analyze (8 lanes per page, 4 pages per warp) and store every value they consume to their TMEM lanes
// (tcgen05.st 32x32b.x4: thread i -> its quadrant's lane i, four columns per step); two TMEM halves alternate, so
// tile k+1 is analysed while tile k waits for its verdict.  Two control warps (even / odd tiles) do everything that
// needs the rest of the GPU, off the compute warps' critical path:
//   insert the tile's hashes in the table, publish INSERTED;
//   wait until every earlier tile has inserted (the owner of a hash is the lowest page holding it, so a verdict
//   needs all earlier pages in the table), then read each page's owner: keep / DUP / PARENT / PEER;
//   publish the tile's count of kept pages, decoupled look-back over the earlier tiles for the first slot;
//   hand the 32 slots (or "drop") to the compute warps, which move their pages TMEM -> registers -> packed image
//   (tcgen05.ld + 128-bit stores) before they reuse the half.
// Every wait is for a lower-numbered tile, every CTA is resident, each takes its tiles in increasing order: the
// lowest unfinished tile can always proceed.
constexpr int kOpWarps = 8;                       // compute warps
constexpr int kOpDrainWarps = 4;                  // one per TMEM quadrant
// four warpgroups: compute 0-3 | compute 4-7 | control 8, 9 (10, 11 idle) | drain 12-15; setmaxnreg moves registers
// from the last two to the first two
constexpr int kOpThreads = 512;
constexpr int kOpComputeRegs = 192, kOpOtherRegs = 64;
constexpr uint32_t kOpNoTile = 0xFFFFFFFFu;
constexpr int kOpTile = kOpWarps * 4;             // 32 pages, one per control lane
constexpr unsigned long long kStInserted = 1ull << 62, kStAgg = 2ull << 62, kStPrefix = 3ull << 62;
constexpr int kBarComputed = 1, kBarDecided = 3, kBarDrained = 5, kBarCompute = 7;  // named barriers: + half
constexpr int kCntComputed = kOpWarps * 32 + 32, kCntDecided = 32 + kOpDrainWarps * 32, kCntDrained = (kOpDrainWarps + kOpWarps) * 32;
__device__ __forceinline__ void tmem_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tmem_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
constexpr int kOpSlotDrop = -1;
#ifndef GRIT_OP_RING
#define GRIT_OP_RING 16
#endif
// 16-byte loads per lane in flight (8 warps x 32 lanes x 16 x 16 B = 64 KB per SM); measured per 32-page tile:
// 4: 8.2 us, 8: 7.8, 12: 6.3, 16: 7.1, 24: 8.9
constexpr int kOpRing = GRIT_OP_RING;

__device__ __forceinline__ void bar_sync(int id, int count) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(count) : "memory"); }
__device__ __forceinline__ void bar_arrive(int id, int count) { asm volatile("bar.arrive %0, %1;" ::"r"(id), "r"(count) : "memory"); }

__device__ __forceinline__ void tmem_st4(uint32_t taddr, const uint4& v) {
  asm volatile("tcgen05.st.sync.aligned.32x32b.x4.b32 [%0], {%1, %2, %3, %4};" ::"r"(taddr), "r"(v.x), "r"(v.y), "r"(v.z),
               "r"(v.w)
               : "memory");
}
__device__ __forceinline__ uint4 tmem_ld4(uint32_t taddr) {
  uint4 v;
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0, %1, %2, %3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
               : "r"(taddr)
               : "memory");
  return v;
}
__device__ __forceinline__ void tmem_wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tmem_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

#ifdef GRIT_OP_PROF
__device__ unsigned long long g_op_prof[16];  // ns summed over CTAs: compute, wait verdict, drain | wait computed, insert, wait inserted, verdict, look-back
__device__ __forceinline__ unsigned long long op_now() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
#define OP_T(var) const unsigned long long var = op_now()
// a barrier does not block where it is issued but at the first use of what it protects: read the clock after such a use
__device__ __forceinline__ unsigned long long op_now_after(uint32_t dep) {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t) : "r"(dep));
  return t;
}
#define OP_TD(var, dep) const unsigned long long var = op_now_after((uint32_t)(dep))
#define OP_ADD(slot, a, b) \
  if (lane == 0 && (warp == 0 || warp == kOpWarps)) atomicAdd(&g_op_prof[slot], (b) - (a))
#define OP_ADD_RAW(slot, a, b) atomicAdd(&g_op_prof[slot], (b) - (a))
#else
#define OP_T(var)
#define OP_TD(var, dep)
#define OP_ADD(slot, a, b)
#define OP_ADD_RAW(slot, a, b)
#endif

struct OnepassShared {
  uint32_t tmem_base;
  uint32_t tile_of_half[2];      // the tile parked in each TMEM half (kOpNoTile: the CTA has run out of tiles)
  uint32_t next_id;              // hand-over of the next ticket between the compute warps
  unsigned long long hash[2][kOpTile];
  uint32_t crc[2][kOpTile];
  uint32_t nonzero[2][kOpTile];  // 1: page exists and is not all zero; 0: zero page; 2: no such page
  int slot[2][kOpTile];          // verdict: slot in the packed image, or kOpSlotDrop
};

__global__ void __launch_bounds__(kOpThreads, 1)
k_pack_onepass(const Gf2Tables* __restrict__ tabs, const uint4* __restrict__ image, uint64_t n_pages,
               uint32_t* __restrict__ crc, unsigned long long* __restrict__ hash, uint32_t* __restrict__ ref, DedupTable t,
               PeerCrcs crcs, uint4* __restrict__ packed, uint64_t capacity_pages, unsigned long long* __restrict__ status,
               DevCounters* __restrict__ cnt) {
  __shared__ OnepassShared sh;
  const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const uint32_t n_tiles = (uint32_t)((n_pages + kOpTile - 1) / kOpTile);
  if (blockIdx.x >= n_tiles) return;
  load_tables(tabs);  // all warps fill the tables; ends with __syncthreads
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(
                     (uint32_t)__cvta_generic_to_shared(&sh.tmem_base))
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tmem_fence_before();
  __syncthreads();
  tmem_fence_after();
  const uint32_t tmem_base = sh.tmem_base;

  if (warp < kOpWarps) {
    // ---------------------------------------------------------------- compute warps (two warpgroups, most registers)
    asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(kOpComputeRegs));
    const uint32_t g = lane >> 3, l = lane & 7, lane4 = lane * 4;
    // this warp's TMEM window: its quadrant's 32 lanes, 256 columns (two halves of 128) at (warp / 4) * 256
    const uint32_t tm = tmem_base + (((warp & 3) * 32) << 16) + (warp >> 2) * 256;
    // Tiles k * grid + blockIdx: no CTA's parked tile ever waits for a tile that has not been started (tickets taken
    // ahead of time did: measured 23 ms against 11)
    uint32_t id = blockIdx.x, id_next = blockIdx.x + gridDim.x;
    uint4 ring[kOpRing];
    uint64_t p = (uint64_t)id * kOpTile + warp * 4 + g;
    const uint4* src = image + p * 256 + l;
    ring_fill(ring, src, p < n_pages);
    uint32_t k = 0;
    for (;; k++) {
      const uint32_t half = k & 1;
      if (id >= n_tiles) break;
      OP_T(t0);
      if (k >= 2) {  // the half still holds tile k - 2 until the drain warps are done with it
        bar_sync(kBarDrained + half, kCntDrained);
        tmem_fence_after();
      }
      OP_TD(t1, sh.slot[half][0]);
      if (threadIdx.x == 0) sh.tile_of_half[half] = id;
      PageAcc a;
      a.z = 0;
      a.h = 0;
      const bool on = p < n_pages;
      const uint64_t p_next = (uint64_t)id_next * kOpTile + warp * 4 + g;
      const uint4* src_next = image + p_next * 256 + l;
      const bool next_on = id_next < n_tiles && p_next < n_pages;
      const uint32_t tcol = tm + half * 128;
      page_pass<true, true>(a, ring, src, on, src_next, next_on, lane4, [&](auto jc, const uint4& v) {
        constexpr int j = decltype(jc)::value;
        tmem_st4(tcol + 4 * j, v);
      });
      const uint32_t page_crc = finish_crc(a, l);
      uint32_t z = a.z;
      z |= __shfl_xor_sync(0xFFFFFFFFu, z, 1);
      z |= __shfl_xor_sync(0xFFFFFFFFu, z, 2);
      z |= __shfl_xor_sync(0xFFFFFFFFu, z, 4);
      const unsigned long long page_hash = finish_hash(a, l);
      if (l == 0) {
        const uint32_t pl = warp * 4 + g;
        sh.hash[half][pl] = page_hash;
        sh.crc[half][pl] = page_crc;
        sh.nonzero[half][pl] = on ? (z ? 1u : 0u) : 2u;
        if (on) {
          crc[p] = page_crc;
          hash[p] = page_hash;
          if (z) {  // the control warp will claim this page's table slot: have the lines in L2 by then
            const uint64_t home = table_home(page_hash, t.mask);
            asm volatile("prefetch.global.L2 [%0];" ::"l"(t.slots + home));
          }
        }
      }
      tmem_wait_st();
      tmem_fence_before();
      bar_sync(kBarCompute, kOpWarps * 32);  // thread 0's tile_of_half before anyone's arrival
      bar_arrive(kBarComputed + half, kCntComputed);  // the control warp of this parity takes it from here
      OP_T(t2);
      OP_ADD(0, t1, t2);
      OP_ADD(1, t0, t1);
      id = id_next;
      id_next += gridDim.x;
      p = p_next;
      src = src_next;
    }
    // out of tiles: tell both control warps (and through them the drain warps), meeting the drain warps' last arrivals
    for (uint32_t e = 0; e < 2; e++, k++) {
      const uint32_t half = k & 1;
      if (k >= 2) bar_sync(kBarDrained + half, kCntDrained);
      if (threadIdx.x == 0) sh.tile_of_half[half] = kOpNoTile;
      bar_sync(kBarCompute, kOpWarps * 32);
      bar_arrive(kBarComputed + half, kCntComputed);
    }
  } else if (warp >= 12) {
    // ---------------------------------------------------------------- drain warps: TMEM quadrant warp % 4
    asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(kOpOtherRegs));
    const uint32_t g = lane >> 3, l = lane & 7, q = warp & 3;
    for (uint32_t k = 0;; k++) {
      const uint32_t half = k & 1;
      OP_T(d0);
      bar_sync(kBarDecided + half, kCntDecided);
      tmem_fence_after();
      if (sh.tile_of_half[half] == kOpNoTile) {  // out of tiles; the other half's notice is the next thing to come
        bar_sync(kBarDecided + (half ^ 1), kCntDecided);
        break;
      }
      OP_TD(d1, sh.slot[half][q * 4 + g]);
#pragma unroll 1
      for (uint32_t cw = q; cw < (uint32_t)kOpWarps; cw += 4) {  // the compute warps that share this quadrant
        const uint32_t tm = tmem_base + ((q * 32) << 16) + (cw >> 2) * 256 + half * 128;
        const int s = sh.slot[half][cw * 4 + g];
        uint4* d = packed + (uint64_t)(s < 0 ? 0 : s) * 256 + l;
        if (!__any_sync(0xFFFFFFFFu, s >= 0)) continue;
#pragma unroll 1
        for (int jb = 0; jb < 32; jb += 8) {  // eight steps per round trip to TMEM, then their eight stores
          uint4 v[8];
#pragma unroll
          for (int j = 0; j < 8; j++) v[j] = tmem_ld4(tm + 4 * (jb + j));
          tmem_wait_ld();
          if (s >= 0) {
#pragma unroll
            for (int j = 0; j < 8; j++) stg_stream(d + 8 * (jb + j), v[j]);
          }
        }
      }
      tmem_fence_before();
      bar_arrive(kBarDrained + half, kCntDrained);
      OP_T(d2);
      if (warp == 12 && lane == 0) {
        OP_ADD_RAW(2, d1, d2);
        OP_ADD_RAW(3, d0, d1);
      }
    }
  } else if (warp < kOpWarps + 2) {
    // ---------------------------------------------------------------- control warps: lane i <-> page i of the tile
    asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(kOpOtherRegs));
    const uint32_t half = warp - kOpWarps;  // control warp 0: the tiles parked in TMEM half 0, control warp 1: half 1
    const uint32_t lt_mask = (1u << lane) - 1;
    uint32_t n_kind[5] = {0, 0, 0, 0, 0};
    uint32_t wm = 0;  // every tile below wm is known to have inserted
    for (;;) {
      OP_T(c0);
      bar_sync(kBarComputed + half, kCntComputed);
      const uint32_t tile = sh.tile_of_half[half];
      if (tile == kOpNoTile) {
        bar_arrive(kBarDecided + half, kCntDecided);
        break;
      }
      const uint64_t p = (uint64_t)tile * kOpTile + lane;
      const uint32_t nz = sh.nonzero[half][lane];
      OP_TD(c1, nz);
      const unsigned long long h = sh.hash[half][lane];
      const uint32_t my_crc = sh.crc[half][lane];
      uint64_t sl = 0;
      if (nz == 1u) {
        const uint32_t v = kOriginSelf << 26 | (uint32_t)p;
        for (sl = table_home(h, t.mask);; sl = (sl + 1) & t.mask) {
          // one round trip: the CAS claims an empty slot or tells whose it is (most slots are met empty)
          const unsigned long long key = atomicCAS(&t.slots[sl].key, 0ull, h);
          if (key == 0ull || key == h) {
            if (key == 0ull || __ldcg(&t.slots[sl].val) > v) atomicMin(&t.slots[sl].val, v);  // no result used: a RED
            break;
          }
        }
      }
      __threadfence();  // the tile's crc / hash stores (seen through the barrier) and table atomics before the flag
      __syncwarp();
      if (lane == 0) st_status(&status[tile], kStInserted);
      OP_T(c2);
      // every earlier tile must have inserted before an owner can be read
      for (uint32_t b0 = wm; b0 < tile; b0 += 128) {
        for (;;) {
          unsigned long long s0 = kStInserted, s1 = kStInserted, s2 = kStInserted, s3 = kStInserted;
          const uint32_t i0 = b0 + lane;
          if (i0 < tile) s0 = ld_status(&status[i0]);
          if (i0 + 32 < tile) s1 = ld_status(&status[i0 + 32]);
          if (i0 + 64 < tile) s2 = ld_status(&status[i0 + 64]);
          if (i0 + 96 < tile) s3 = ld_status(&status[i0 + 96]);
          const bool ok = (s0 >> 62) && (s1 >> 62) && (s2 >> 62) && (s3 >> 62);
          if (__all_sync(0xFFFFFFFFu, ok)) break;
          __nanosleep(100);
        }
      }
      wm = tile;
      __threadfence();
      OP_T(c3);
      // verdict
      uint32_t out = 0u;  // zero page
      bool keep = false;
      if (nz == 1u) {
        const uint32_t v = __ldcg(&t.slots[sl].val), origin = v >> 26, idx = v & ((1u << 26) - 1);
        out = kRefKeep;
        if (!(origin == kOriginSelf && idx == (uint32_t)p)) {
          const uint32_t ocrc = origin == 0             ? crcs.parent[idx]
                                : origin == kOriginSelf ? __ldcg(&crc[idx])
                                                        : crcs.peer[origin - 1][idx];
          if (ocrc == my_crc)
            out = origin == 0             ? (3u << 29 | idx)
                  : origin == kOriginSelf ? (2u << 29 | idx)
                                          : (4u << 29 | (origin - 1) << 26 | idx);
        }
        keep = out == kRefKeep;
      }
      const uint32_t bal = __ballot_sync(0xFFFFFFFFu, keep);
      const uint32_t total = __popc(bal), rank = __popc(bal & lt_mask);
      if (lane == 0) st_status(&status[tile], (tile == 0 ? kStPrefix : kStAgg) | total);
      OP_T(c4);
      uint32_t excl = 0;
      if (tile > 0) {
        // decoupled look-back, 128 tiles per round trip (in step with each other, a whole round of CTAs publishes
        // its counts at about the same time: the nearest prefix is usually a round away)
        int64_t look = (int64_t)tile - 1;
        for (;;) {
          unsigned long long st[4];
          for (;;) {
            bool ready = true;
#pragma unroll
            for (int b = 0; b < 4; b++) {
              const int64_t idx = look - 32 * b - lane;
              st[b] = idx >= 0 ? ld_status(&status[idx]) : kStPrefix;
              ready = ready && (st[b] >> 62) >= 2;
            }
            // a batch beyond the nearest prefix need not be ready: only what lies in front of it counts
            uint32_t need = 0xFFFFFFFFu;
            bool done = false, ok = true;
#pragma unroll
            for (int b = 0; b < 4; b++) {
              const uint32_t pre = __ballot_sync(0xFFFFFFFFu, (st[b] >> 62) == 3);
              const uint32_t unr = __ballot_sync(0xFFFFFFFFu, (st[b] >> 62) < 2);
              if (!done) {
                const uint32_t upto = pre ? ((2u << (__ffs(pre) - 1)) - 1u) : 0xFFFFFFFFu;  // lanes up to the first prefix
                if (unr & upto) ok = false;
                if (pre) done = true;
              }
            }
            (void)need;
            (void)ready;
            if (ok) break;
            __nanosleep(40);
          }
          bool found = false;
#pragma unroll
          for (int b = 0; b < 4; b++) {
            if (found) break;
            const uint32_t incl = __ballot_sync(0xFFFFFFFFu, (st[b] >> 62) == 3);
            const uint32_t upto = incl ? (uint32_t)__ffs(incl) - 1u : 31u;
            uint32_t v = lane <= upto ? (uint32_t)st[b] : 0u;
#pragma unroll
            for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xFFFFFFFFu, v, o);
            excl += v;
            found = incl != 0;
          }
          if (found) break;
          look -= 128;
        }
        if (lane == 0) st_status(&status[tile], kStPrefix | (excl + total));
      }
      if (lane == 0 && tile == n_tiles - 1) {
        cnt->n_stored = excl + total;
        cnt->slot_base_next = excl + total;
      }
      int sv = kOpSlotDrop;
      if (keep) {
        const uint32_t slot = excl + rank;
        out = 1u << 29 | slot;
        if (slot < capacity_pages) sv = (int)slot;
        else atomicAdd(&cnt->overflow, 1ull);
      }
      if (nz != 2u) {
        ref[p] = out;
        n_kind[keep ? 1u : out >> 29]++;
      }
      sh.slot[half][lane] = sv;
      __threadfence_block();
      bar_arrive(kBarDecided + half, kCntDecided);
      OP_T(c5);
      OP_ADD(8, c0, c1);
      OP_ADD(9, c1, c2);
      OP_ADD(10, c2, c3);
      OP_ADD(11, c3, c4);
      OP_ADD(12, c4, c5);
    }
#pragma unroll
    for (int q = 0; q < 5; q++) {
      uint32_t v = n_kind[q];
#pragma unroll
      for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xFFFFFFFFu, v, o);
      if (lane == 0 && v) atomicAdd(&cnt->kind[q], (unsigned long long)v);
    }
  } else {
    asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(kOpOtherRegs));  // warps 10, 11: only fill the warpgroup
  }
  tmem_fence_before();
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem_base) : "memory");
}

// ------------------------------------------------------------------ K4 restore
// Where the bytes of page p live: nullptr for a zero page or a reference that cannot be followed.
struct PageSrc {
  const uint4* src;
  bool bad;
};

__device__ __forceinline__ PageSrc resolve_source(uint64_t p, bool on, const uint4* __restrict__ packed,
                                                  uint64_t n_slots, const uint32_t* __restrict__ ref,
                                                  uint64_t n_pages, const PeerImages& imgs, const PackedWindow& win,
                                                  uint64_t page0) {
  PageSrc r{nullptr, false};
  if (!on) return r;
  uint32_t x = ref[p], kind = x >> 29, v = x & 0x1FFFFFFFu;
  if (kind == 2u) {  // DUP -> the owner's slot, or (windowed restore) the owner's page where an earlier launch put it
    const uint32_t x2 = v < n_pages ? ref[v] : 0u;
    r.bad = (x2 >> 29) != 1u;
    if (win.restored && v < page0) {
      if (!r.bad) r.src = win.restored + (uint64_t)v * 256;
      return r;
    }
    kind = 1u;
    v = x2 & 0x1FFFFFFFu;
  }
  if (kind == 1u) {  // `packed` holds slots [win.slot0, win.slot0 + n_slots)
    if (!r.bad && v >= win.slot0 && v - win.slot0 < n_slots) r.src = packed + (uint64_t)(v - win.slot0) * 256;
    else r.bad = true;
  } else if (kind == 3u) {
    if (imgs.parent && v < imgs.parent_pages) r.src = reinterpret_cast<const uint4*>(imgs.parent) + (uint64_t)v * 256;
    else r.bad = true;
  } else if (kind == 4u) {
    const uint32_t rk = v >> 26, idx = v & ((1u << 26) - 1);
    if (rk < imgs.n_peers && imgs.peer[rk] && idx < imgs.peer_pages[rk])
      r.src = reinterpret_cast<const uint4*>(imgs.peer[rk]) + (uint64_t)idx * 256;
    else r.bad = true;
  } else if (kind != 0u) {
    r.bad = true;
  }
  return r;
}

// restores pages [page0, page1) of the image into out[0 .. page1-page0)
__global__ void __launch_bounds__(kThreads, 1)
k_restore(const Gf2Tables* __restrict__ tabs, const uint4* __restrict__ packed, uint64_t n_slots,
          const uint32_t* __restrict__ ref, const uint32_t* __restrict__ crc, uint64_t n_pages,
          uint64_t page0, uint64_t page1, PeerImages imgs, PackedWindow win, uint4* __restrict__ out, int verify,
          DevCounters* __restrict__ cnt) {
  if (verify) load_tables(tabs);
  const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const uint32_t g = lane >> 3, l = lane & 7, lane4 = lane * 4;
  const uint64_t n_groups = (page1 - page0 + 3) / 4;
  const uint64_t stride = (uint64_t)gridDim.x * kWarps;
  uint64_t grp = (uint64_t)blockIdx.x * kWarps + warp;
  if (grp >= n_groups) return;

  for (; grp < n_groups; grp += stride) {
    const uint64_t p = page0 + 4 * grp + g;
    const bool on = p < page1;
    const PageSrc ps = resolve_source(p, on, packed, n_slots, ref, n_pages, imgs, win, page0);
    const bool have = ps.src != nullptr;
    bool bad = ps.bad;
    const uint4* s = have ? ps.src + l : packed;  // never dereferenced without a source
    uint4* d = out + (p - page0) * 256 + l;
    PageAcc a;
    a.z = 0;
    a.h = 0;
    uint4 cur[8], nxt[8];
    load_batch(cur, s, have);
    load_batch(nxt, s + 64, have);
    if (on) {
#pragma unroll
      for (int j = 0; j < 8; j++) stg_stream(d + 8 * j, cur[j]);
    }
    if (verify) absorb_batch<true, false>(a, cur, lane4);
    load_batch(cur, s + 128, have);
    if (on) {
#pragma unroll
      for (int j = 0; j < 8; j++) stg_stream(d + 64 + 8 * j, nxt[j]);
    }
    if (verify) absorb_batch<false, false>(a, nxt, lane4);
    load_batch(nxt, s + 192, have);
    if (on) {
#pragma unroll
      for (int j = 0; j < 8; j++) stg_stream(d + 128 + 8 * j, cur[j]);
    }
    if (verify) absorb_batch<false, false>(a, cur, lane4);
    if (on) {
#pragma unroll
      for (int j = 0; j < 8; j++) stg_stream(d + 192 + 8 * j, nxt[j]);
    }
    if (verify) {
      absorb_batch<false, false>(a, nxt, lane4);
      const uint32_t page_crc = finish_crc(a, l);
      if (have && page_crc != crc[p]) bad = true;
    }
    if (bad && l == 0) {
      atomicAdd(&cnt->n_bad, 1ull);
      atomicMin(&cnt->first_bad, (unsigned long long)p);
    }
  }
}

// ------------------------------------------------------------------ K5 synth
__global__ void __launch_bounds__(256)
k_synth(uint4* __restrict__ image, uint64_t n_pages, uint64_t page_base, unsigned long long seed,
        uint32_t pct_zero, uint32_t pct_dup, uint32_t dup_pool) {
  const uint64_t n = n_pages * 256;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (uint64_t)gridDim.x * blockDim.x) {
    const uint64_t gp = page_base + (i >> 8);
    const uint32_t seg = (uint32_t)i & 255;
    const unsigned long long r = sm64(seed ^ (gp * kGold));
    const uint32_t cls = (uint32_t)(r % 100);
    uint4 v = make_uint4(0, 0, 0, 0);
    if (cls >= pct_zero) {
      unsigned long long key = gp;
      if (cls < pct_zero + pct_dup && dup_pool) key = 1ull << 63 | ((r >> 32) % dup_pool);
      const unsigned long long base = sm64(seed ^ sm64(key));
      const unsigned long long lo = sm64(base + 2ull * seg), hi = sm64(base + 2ull * seg + 1);
      v = make_uint4((uint32_t)lo, (uint32_t)(lo >> 32), (uint32_t)hi, (uint32_t)(hi >> 32));
    }
    image[i] = v;
  }
}

int grid_for(const LaunchGeom& g, uint64_t n_items, int per_block, int blocks_per_sm) {
  const uint64_t need = (n_items + per_block - 1) / per_block;
  const uint64_t cap = (uint64_t)g.sm_count * blocks_per_sm;
  return (int)(need < cap ? (need ? need : 1) : cap);
}

}  // namespace

uint32_t page_kernel_smem_bytes() { return kSmBytes; }

cudaError_t configure_kernels() {
  cudaError_t e = cudaFuncSetAttribute(k_analyze, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmBytes);
  if (e != cudaSuccess) return e;
  e = cudaFuncSetAttribute(k_analyze_through, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmBytesWT);
  if (e != cudaSuccess) return e;
  e = cudaFuncSetAttribute(k_pack_onepass, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmBytes);
  if (e != cudaSuccess) return e;
  return cudaFuncSetAttribute(k_restore, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmBytes);
}

uint32_t onepass_tile_pages() { return kOpTile; }

#ifdef GRIT_OP_PROF
extern "C" void grit_op_prof(unsigned long long* out, int reset) {  // development aid, not in the ABI
  cudaMemcpyFromSymbol(out, g_op_prof, sizeof(g_op_prof));
  if (reset) {
    unsigned long long z[16] = {};
    cudaMemcpyToSymbol(g_op_prof, z, sizeof z);
  }
}
#endif

cudaError_t launch_pack_onepass(const LaunchGeom& g, const Gf2Tables* d_tab, const void* d_image, uint64_t n_pages,
                                uint32_t* d_crc, uint64_t* d_hash, uint32_t* d_ref, DedupTable t, PeerCrcs crcs, void* d_packed,
                                uint64_t capacity_pages, unsigned long long* d_tile_status, DevCounters* d_cnt, cudaStream_t s) {
  if (!n_pages) return cudaSuccess;
  const int grid = grid_for(g, n_pages, kOpTile, 1);  // every CTA must be resident: at most one per SM
  k_pack_onepass<<<grid, kOpThreads, kSmBytes, s>>>(d_tab, static_cast<const uint4*>(d_image), n_pages, d_crc,
                                                    reinterpret_cast<unsigned long long*>(d_hash), d_ref, t, crcs,
                                                    static_cast<uint4*>(d_packed), capacity_pages, d_tile_status, d_cnt);
  return cudaGetLastError();
}

cudaError_t launch_analyze(const LaunchGeom& g, const Gf2Tables* d_tab, const void* d_image, uint64_t n_pages,
                           uint32_t* d_crc, uint64_t* d_hash, uint32_t* d_ref, cudaStream_t s) {
  if (!n_pages) return cudaSuccess;
  const int grid = grid_for(g, (n_pages + 3) / 4, kK1Warps, 1);
  k_analyze<<<grid, kK1Threads, kSmBytes, s>>>(d_tab, static_cast<const uint4*>(d_image), n_pages, d_crc,
                                               reinterpret_cast<unsigned long long*>(d_hash), d_ref, nullptr, -1);
  return cudaGetLastError();
}

// K1 for a full image: runs the write-through variant if the sample (launch_probe_dense) found the image dense and
// plain K1 otherwise -- decided on the device, no host sync.  K3 reads the same hint to skip pages in place.
// The arrays may start at page `page_base` of the whole image (sliced K1 of a ranked pack).
cudaError_t launch_probe_dense(const void* d_image, uint64_t n_pages, DevCounters* d_cnt, cudaStream_t s) {
  if (!n_pages) return cudaSuccess;
  k_probe_dense<<<1, 1024, 0, s>>>(static_cast<const uint4*>(d_image), n_pages, d_cnt);
  return cudaGetLastError();
}

__global__ void __launch_bounds__(1024)
k_peer_overlap_hint(const unsigned long long* __restrict__ hash, ProbeLists lists, uint64_t n, DevCounters* __restrict__ cnt) {
  __shared__ uint32_t s_match;
  if (threadIdx.x == 0) s_match = 0;
  __syncthreads();
  const uint64_t samples = n < 8192 ? n : 8192;
  for (uint64_t i = threadIdx.x; i < samples; i += blockDim.x) {
    const uint64_t p = i * n / samples;
    const unsigned long long h = hash[p];
    bool m = false;
    for (int o = 1; o < 9; o++)
      if (lists.hash[o] && p < lists.n[o] && lists.hash[o][p] == h) m = true;
    if (m) atomicAdd(&s_match, 1u);
  }
  __syncthreads();
  if (threadIdx.x == 0 && s_match * 16 > samples) cnt->dense_hint = 0u;
}

cudaError_t launch_peer_overlap_hint(const unsigned long long* d_hash, const ProbeLists& lists, uint64_t n, DevCounters* d_cnt,
                                     cudaStream_t s) {
  if (!n) return cudaSuccess;
  k_peer_overlap_hint<<<1, 1024, 0, s>>>(d_hash, lists, n, d_cnt);
  return cudaGetLastError();
}

cudaError_t launch_analyze_auto(const LaunchGeom& g, const Gf2Tables* d_tab, const void* d_image, uint64_t n_pages,
                                uint32_t* d_crc, uint64_t* d_hash, uint32_t* d_ref, void* d_packed,
                                uint64_t capacity_pages, uint64_t page_base, DevCounters* d_cnt, cudaStream_t s) {
  if (!n_pages) return cudaSuccess;
  const int grid = grid_for(g, (n_pages + 3) / 4, kWarps, 1);
  k_analyze_through<<<grid, kThreads, kSmBytesWT, s>>>(d_tab, static_cast<const uint4*>(d_image), n_pages, d_crc,
                                                       reinterpret_cast<unsigned long long*>(d_hash), d_ref,
                                                       static_cast<uint4*>(d_packed), capacity_pages, page_base, d_cnt, 1);
  k_analyze<<<grid_for(g, (n_pages + 3) / 4, kK1Warps, 1), kK1Threads, kSmBytes, s>>>(
      d_tab, static_cast<const uint4*>(d_image), n_pages, d_crc, reinterpret_cast<unsigned long long*>(d_hash), d_ref, d_cnt, 0);
  return cudaGetLastError();
}

cudaError_t launch_table_clear(const LaunchGeom& g, DedupTable t, cudaStream_t s) {
  k_table_clear<<<grid_for(g, t.mask + 1, 256, 8), 256, 0, s>>>(t);
  return cudaGetLastError();
}

cudaError_t launch_table_insert(const LaunchGeom& g, DedupTable t, const uint64_t* d_hash,
                                const uint32_t* d_ref_or_null, uint64_t n, uint32_t origin, uint64_t index_base,
                                cudaStream_t s, bool pending_only) {
  if (!n) return cudaSuccess;
  k_table_insert<<<grid_for(g, n, 256, 8), 256, 0, s>>>(
      t, reinterpret_cast<const unsigned long long*>(d_hash), d_ref_or_null, n, origin, index_base, pending_only);
  return cudaGetLastError();
}

cudaError_t launch_resolve_parent(const LaunchGeom& g, DedupTable t, uint64_t n_pages, const uint32_t* d_crc,
                                  const uint64_t* d_hash, uint32_t* d_ref, ListView<uint32_t> parent_crc,
                                  DevCounters* d_cnt, cudaStream_t s) {
  if (!n_pages) return cudaSuccess;
  k_resolve_parent<<<grid_for(g, n_pages, 256, 8), 256, 0, s>>>(
      t, n_pages, d_crc, reinterpret_cast<const unsigned long long*>(d_hash), d_ref, parent_crc, d_cnt);
  return cudaGetLastError();
}

cudaError_t launch_table_probe_lists(const LaunchGeom& g, DedupTable t, const ProbeLists& lists, cudaStream_t s) {
  uint64_t longest = 0;
  for (int o = 0; o < 9; o++)
    if (lists.hash[o]) longest = std::max(longest, lists.n[o]);
  if (!longest) return cudaSuccess;
  dim3 grid(grid_for(g, longest, 256, 8), 9);  // an origin without a list costs a few empty blocks
  k_table_probe_lists<<<grid, 256, 0, s>>>(t, lists);
  return cudaGetLastError();
}

cudaError_t launch_resolve(const LaunchGeom& g, DedupTable t, uint64_t page0, uint64_t page1, const uint32_t* d_crc,
                           const uint64_t* d_hash, uint32_t* d_ref, PeerCrcs crcs, DevCounters* d_cnt, cudaStream_t s,
                           bool pending_only) {
  if (page1 <= page0) return cudaSuccess;
  k_resolve<<<grid_for(g, page1 - page0, 256, 8), 256, 0, s>>>(
      t, page0, page1, d_crc, reinterpret_cast<const unsigned long long*>(d_hash), d_ref, crcs, d_cnt, pending_only);
  return cudaGetLastError();
}

cudaError_t launch_verify_dups(const LaunchGeom& g, const void* d_image, uint64_t n_pages, uint32_t* d_ref, DevCounters* d_cnt,
                               cudaStream_t s) {
  if (!n_pages) return cudaSuccess;
  k_verify_dups<<<grid_for(g, n_pages, 64, 8), 256, 0, s>>>(static_cast<const uint4*>(d_image), n_pages, d_ref, d_cnt);
  return cudaGetLastError();
}

cudaError_t launch_compact(const LaunchGeom& g, const void* d_image, uint64_t n_pages, uint32_t* d_ref,
                           void* d_packed, uint64_t capacity_pages, uint64_t slot_base, bool chained,
                           unsigned long long* d_tile_status, DevCounters* d_cnt, cudaStream_t s, bool in_place) {
  if (!n_pages) return cudaSuccess;
  const int place = (in_place && !chained && slot_base == 0) ? 1 : 0;
  k_compact<<<grid_for(g, n_pages, kCompactTile, 8), 256, 0, s>>>(
      static_cast<const uint4*>(d_image), n_pages, d_ref, static_cast<uint4*>(d_packed), capacity_pages,
      slot_base, chained ? 1 : 0, place, d_tile_status, d_cnt, chained ? 0 : 1);
  if (!chained)  // a chained launch's slot base is the device's; its chunks are short, the plain tiles do
    k_compact_wide<<<grid_for(g, n_pages, 4 * kCompactTile, 8), 256, 0, s>>>(
        static_cast<const uint4*>(d_image), n_pages, d_ref, static_cast<uint4*>(d_packed), capacity_pages,
        slot_base, place, d_tile_status, d_cnt);
  return cudaGetLastError();
}

cudaError_t launch_restore(const LaunchGeom& g, const Gf2Tables* d_tab, const void* d_packed, uint64_t n_slots,
                           const uint32_t* d_ref, const uint32_t* d_crc, uint64_t n_pages, uint64_t page0,
                           uint64_t page1, PeerImages imgs, void* d_out, int verify, DevCounters* d_cnt,
                           cudaStream_t s, PackedWindow win) {
  if (page1 <= page0) return cudaSuccess;
  const int grid = grid_for(g, (page1 - page0 + 3) / 4, kWarps, 1);
  k_restore<<<grid, kThreads, verify ? kSmBytes : 0, s>>>(d_tab, static_cast<const uint4*>(d_packed), n_slots,
                                                          d_ref, d_crc, n_pages, page0, page1, imgs, win,
                                                          static_cast<uint4*>(d_out), verify, d_cnt);
  return cudaGetLastError();
}

cudaError_t launch_synth(const LaunchGeom& g, void* d_image, uint64_t n_pages, uint64_t page_base, uint64_t seed,
                         uint32_t pct_zero, uint32_t pct_dup, uint32_t dup_pool, cudaStream_t s) {
  if (!n_pages) return cudaSuccess;
  k_synth<<<grid_for(g, n_pages * 256, 256, 8), 256, 0, s>>>(static_cast<uint4*>(d_image), n_pages, page_base,
                                                             seed, pct_zero, pct_dup, dup_pool);
  return cudaGetLastError();
}

}  // namespace grit
