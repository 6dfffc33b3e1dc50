/*
 * gritpack oracle -- TEST INFRASTRUCTURE ONLY (see gritpack_oracle.h).
 *
 * Plain scalar C.  Deliberately shares no code with grit_b200/csrc: every
 * routine here is the slow, obvious form of what the CUDA kernels compute.
 */
#include "gritpack_oracle.h"

#include <pthread.h>
#include <stdlib.h>
#include <string.h>
#if defined(__SSE4_2__)
#include <nmmintrin.h>
#endif

/* ------------------------------------------------------------------ CRC-32C
 * Go: hash/crc32, crc32.Castagnoli = 0x82f63b78 (reflected), simpleUpdate():
 *   crc = ^crc; for each byte v: crc = tab[byte(crc)^v] ^ (crc >> 8); return ^crc
 * The reference itself has no call site; north_star fixes this definition.
 */
static uint32_t crc_tab[256];
static pthread_once_t crc_once = PTHREAD_ONCE_INIT;

static void crc_init(void) {
  for (uint32_t i = 0; i < 256; i++) {
    uint32_t c = i;
    for (int k = 0; k < 8; k++) c = (c & 1) ? (c >> 1) ^ 0x82F63B78u : c >> 1;
    crc_tab[i] = c;
  }
}

uint32_t orc_crc32c(uint32_t crc, const void *data, size_t len) {
  pthread_once(&crc_once, crc_init);
  const uint8_t *p = (const uint8_t *)data;
  crc = ~crc;
  for (size_t i = 0; i < len; i++) crc = crc_tab[(uint8_t)crc ^ p[i]] ^ (crc >> 8);
  return ~crc;
}

uint32_t orc_crc32c_hw(uint32_t crc, const void *data, size_t len) {
#if defined(__SSE4_2__)
  const uint8_t *p = (const uint8_t *)data;
  uint64_t c = (uint32_t)~crc;
  while (len >= 8) {
    uint64_t v;
    memcpy(&v, p, 8);
    c = _mm_crc32_u64(c, v);
    p += 8;
    len -= 8;
  }
  while (len--) c = _mm_crc32_u8((uint32_t)c, *p++);
  return ~(uint32_t)c;
#else
  return orc_crc32c(crc, data, len);
#endif
}

/* ------------------------------------------------------------------ hash64
 * Definition (new in this repository; no reference counterpart):
 * the page is 256 segments of 16 bytes, each four little-endian u32 words.
 * Eight interleaved accumulators l = 0..7 take the segments q with q % 8 == l,
 * in four blocks of eight.  Each segment adds an NH term
 *   (w0+K[jj][0])*(w1+K[jj][1]) + (w2+K[jj][2])*(w3+K[jj][3])   (adds mod 2^32,
 * products and sums mod 2^64), the accumulator is multiplied by GOLD between
 * blocks, finalised with its index, summed over l and finalised again.
 * 0 is reserved (empty table slot) and is mapped to 1.
 */
#define GOLD 0x9E3779B97F4A7C15ull
#define FMUL 0xD6E8FEB86659FD93ull

static uint64_t sm64(uint64_t x) {
  x += 0x9E3779B97F4A7C15ull;
  uint64_t z = x;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}

static uint64_t fmix(uint64_t x) {
  x ^= x >> 32;
  x *= FMUL;
  x ^= x >> 32;
  x *= FMUL;
  x ^= x >> 32;
  return x;
}

static uint32_t nh_key[8][4];
static pthread_once_t nh_once = PTHREAD_ONCE_INIT;

static void nh_init(void) {
  uint64_t s = 0x6772697470616B31ull; /* "gritpak1" */
  for (int j = 0; j < 8; j++)
    for (int i = 0; i < 4; i++) {
      s = sm64(s);
      nh_key[j][i] = (uint32_t)(s >> 32) | 1u;
    }
}

static uint32_t ld32(const uint8_t *p) {
  return (uint32_t)p[0] | (uint32_t)p[1] << 8 | (uint32_t)p[2] << 16 | (uint32_t)p[3] << 24;
}

uint64_t orc_hash64_page(const uint8_t *page) {
  pthread_once(&nh_once, nh_init);
  uint64_t sum = 0;
  for (int l = 0; l < 8; l++) {
    uint64_t acc = 0;
    for (int b = 0; b < 4; b++) {
      if (b) acc *= GOLD;
      for (int jj = 0; jj < 8; jj++) {
        int q = 8 * (8 * b + jj) + l;
        const uint8_t *s = page + 16 * q;
        uint32_t w0 = ld32(s) + nh_key[jj][0], w1 = ld32(s + 4) + nh_key[jj][1];
        uint32_t w2 = ld32(s + 8) + nh_key[jj][2], w3 = ld32(s + 12) + nh_key[jj][3];
        acc += (uint64_t)w0 * w1;
        acc += (uint64_t)w2 * w3;
      }
    }
    sum += fmix(acc + (uint64_t)(l + 1) * FMUL);
  }
  uint64_t h = fmix(sum ^ GOLD);
  return h ? h : 1;
}

int orc_page_is_zero(const uint8_t *page) {
  for (unsigned i = 0; i < ORC_PAGE; i++)
    if (page[i]) return 0;
  return 1;
}

void orc_analyze(const uint8_t *image, uint64_t n_pages, uint32_t *crc,
                 uint64_t *hash, uint8_t *zero) {
  for (uint64_t p = 0; p < n_pages; p++) {
    const uint8_t *pg = image + p * ORC_PAGE;
    if (crc) crc[p] = orc_crc32c(0, pg, ORC_PAGE);
    if (hash) hash[p] = orc_hash64_page(pg);
    if (zero) zero[p] = (uint8_t)orc_page_is_zero(pg);
  }
}

/* ------------------------------------------------------------- dedup table
 * hash64 -> lowest (origin, index): origin 0 = parent, 1+r = peer rank r,
 * 15 = this image.  Open addressing, linear probing.
 */
typedef struct {
  uint64_t *key;
  uint32_t *val;
  uint64_t mask;
} dtab;

static int dtab_init(dtab *t, uint64_t n_items) {
  uint64_t cap = 16;
  while (cap < 2 * n_items + 2) cap <<= 1;
  t->key = (uint64_t *)calloc(cap, 8);
  t->val = (uint32_t *)malloc(cap * 4);
  t->mask = cap - 1;
  return (t->key && t->val) ? 0 : -1;
}
static void dtab_free(dtab *t) {
  free(t->key);
  free(t->val);
}
static void dtab_min(dtab *t, uint64_t h, uint32_t v) {
  uint64_t i = (h * GOLD) >> 20 & t->mask;
  for (;; i = (i + 1) & t->mask) {
    if (t->key[i] == 0) {
      t->key[i] = h;
      t->val[i] = v;
      return;
    }
    if (t->key[i] == h) {
      if (v < t->val[i]) t->val[i] = v;
      return;
    }
  }
}
static int dtab_get(const dtab *t, uint64_t h, uint32_t *v) {
  uint64_t i = (h * GOLD) >> 20 & t->mask;
  for (;; i = (i + 1) & t->mask) {
    if (t->key[i] == 0) return 0;
    if (t->key[i] == h) {
      *v = t->val[i];
      return 1;
    }
  }
}

#define ORIGIN_SELF 15u
#define MKVAL(o, i) ((uint32_t)(o) << 26 | (uint32_t)(i))

/* decide every page's reference from the per-page hashes; shared by both packs */
static int resolve_refs(const uint8_t *zero, const uint32_t *crc, const uint64_t *hash,
                        uint64_t n_pages, const orc_table *parent, const orc_table *peers,
                        uint32_t n_peers, uint32_t *ref, orc_pack_stats *st) {
  uint64_t items = n_pages + (parent ? parent->n_pages : 0);
  for (uint32_t r = 0; r < n_peers; r++) items += peers[r].n_pages;
  dtab t;
  if (dtab_init(&t, items)) return -1;
  if (parent)
    for (uint64_t i = 0; i < parent->n_pages; i++) dtab_min(&t, parent->hash[i], MKVAL(0, i));
  for (uint32_t r = 0; r < n_peers; r++)
    for (uint64_t i = 0; i < peers[r].n_pages; i++) dtab_min(&t, peers[r].hash[i], MKVAL(1 + r, i));
  memset(st, 0, sizeof *st);
  st->n_pages = n_pages;
  uint32_t slot = 0;
  for (uint64_t p = 0; p < n_pages; p++) {
    if (zero[p]) {
      ref[p] = ORC_KIND_ZERO << 29;
      st->n_zero++;
      continue;
    }
    uint32_t v;
    if (dtab_get(&t, hash[p], &v)) {
      uint32_t origin = v >> 26, idx = v & (ORC_MAX_PAGES - 1);
      uint32_t ocrc = origin == 0             ? parent->crc[idx]
                      : origin == ORIGIN_SELF ? crc[idx]
                                              : peers[origin - 1].crc[idx];
      if (ocrc == crc[p]) {
        if (origin == 0) {
          ref[p] = ORC_KIND_PARENT << 29 | idx;
          st->n_parent++;
        } else if (origin == ORIGIN_SELF) {
          ref[p] = ORC_KIND_DUP << 29 | idx;
          st->n_dup++;
        } else {
          ref[p] = ORC_KIND_PEER << 29 | (origin - 1) << 26 | idx;
          st->n_peer++;
        }
        continue;
      }
      /* same hash64, different CRC: keep the page, leave the table alone */
    } else {
      dtab_min(&t, hash[p], MKVAL(ORIGIN_SELF, p));
    }
    ref[p] = ORC_KIND_DATA << 29 | slot++;
    st->n_data++;
  }
  dtab_free(&t);
  return 0;
}

int orc_pack(const uint8_t *image, uint64_t n_pages, const orc_table *parent,
             const orc_table *peers, uint32_t n_peers, uint32_t *ref, uint32_t *crc,
             uint64_t *hash, uint8_t *packed, orc_pack_stats *stats) {
  if (n_pages > ORC_MAX_PAGES || n_peers > 8) return -1;
  uint8_t *zero = (uint8_t *)malloc(n_pages ? n_pages : 1);
  if (!zero) return -1;
  orc_analyze(image, n_pages, crc, hash, zero);
  orc_pack_stats st;
  int rc = resolve_refs(zero, crc, hash, n_pages, parent, peers, n_peers, ref, &st);
  free(zero);
  if (rc) return rc;
  for (uint64_t p = 0; p < n_pages; p++)
    if (ref[p] >> 29 == ORC_KIND_DATA)
      memcpy(packed + (uint64_t)(ref[p] & 0x1FFFFFFFu) * ORC_PAGE, image + p * ORC_PAGE, ORC_PAGE);
  if (stats) *stats = st;
  return 0;
}

/* source bytes of page p, or NULL when the reference cannot be followed */
static const uint8_t *page_source(uint64_t p, const uint8_t *packed, uint64_t n_slots,
                                  const uint32_t *ref, uint64_t n_pages,
                                  const uint8_t *parent_image, uint64_t parent_pages,
                                  const uint8_t *const *peer_images, const uint64_t *peer_pages,
                                  uint32_t n_peers) {
  uint32_t kind = ref[p] >> 29, v = ref[p] & 0x1FFFFFFFu;
  if (kind == ORC_KIND_DUP) {
    if (v >= n_pages || ref[v] >> 29 != ORC_KIND_DATA) return NULL;
    kind = ORC_KIND_DATA;
    v = ref[v] & 0x1FFFFFFFu;
  }
  if (kind == ORC_KIND_DATA) return v < n_slots ? packed + (uint64_t)v * ORC_PAGE : NULL;
  if (kind == ORC_KIND_PARENT)
    return (parent_image && v < parent_pages) ? parent_image + (uint64_t)v * ORC_PAGE : NULL;
  if (kind == ORC_KIND_PEER) {
    uint32_t r = v >> 26, i = v & (ORC_MAX_PAGES - 1);
    if (r >= n_peers || !peer_images || !peer_images[r] || i >= peer_pages[r]) return NULL;
    return peer_images[r] + (uint64_t)i * ORC_PAGE;
  }
  return NULL;
}

uint64_t orc_restore(const uint8_t *packed, uint64_t n_slots, const uint32_t *ref,
                     const uint32_t *crc, uint64_t n_pages, const uint8_t *parent_image,
                     uint64_t parent_pages, const uint8_t *const *peer_images,
                     const uint64_t *peer_pages, uint32_t n_peers, uint8_t *out, int verify,
                     uint64_t *first_bad) {
  uint64_t bad = 0, fb = UINT64_MAX;
  for (uint64_t p = 0; p < n_pages; p++) {
    uint8_t *dst = out + p * ORC_PAGE;
    if (ref[p] >> 29 == ORC_KIND_ZERO) {
      memset(dst, 0, ORC_PAGE);
      continue;
    }
    const uint8_t *src = page_source(p, packed, n_slots, ref, n_pages, parent_image, parent_pages,
                                     peer_images, peer_pages, n_peers);
    int ok = src != NULL;
    if (ok) {
      memcpy(dst, src, ORC_PAGE);
      if (verify && orc_crc32c(0, dst, ORC_PAGE) != crc[p]) ok = 0;
    } else {
      memset(dst, 0, ORC_PAGE);
    }
    if (!ok) {
      bad++;
      if (p < fb) fb = p;
    }
  }
  if (first_bad) *first_bad = fb;
  return bad;
}

/* ------------------------------------------------- multithreaded baseline
 * Same work as the GPU path with all host threads: analyze in parallel (hardware
 * CRC where the CPU has it), resolve sequentially, copy in parallel.  Timed by
 * bench.py as a CPU baseline; never used as a checker.
 */
typedef struct {
  const uint8_t *image;
  uint64_t lo, hi;
  uint32_t *crc;
  uint64_t *hash;
  uint8_t *zero;
  const uint32_t *ref;
  uint8_t *packed;
  const uint8_t *src_packed;
  uint8_t *out;
  uint64_t n_pages, n_slots, bad;
  int verify;
} mt_job;

static void *mt_analyze(void *a) {
  mt_job *j = (mt_job *)a;
  for (uint64_t p = j->lo; p < j->hi; p++) {
    const uint8_t *pg = j->image + p * ORC_PAGE;
    j->crc[p] = orc_crc32c_hw(0, pg, ORC_PAGE);
    j->hash[p] = orc_hash64_page(pg);
    j->zero[p] = (uint8_t)orc_page_is_zero(pg);
  }
  return NULL;
}
static void *mt_copy(void *a) {
  mt_job *j = (mt_job *)a;
  for (uint64_t p = j->lo; p < j->hi; p++)
    if (j->ref[p] >> 29 == ORC_KIND_DATA)
      memcpy(j->packed + (uint64_t)(j->ref[p] & 0x1FFFFFFFu) * ORC_PAGE, j->image + p * ORC_PAGE,
             ORC_PAGE);
  return NULL;
}
static void *mt_restore(void *a) {
  mt_job *j = (mt_job *)a;
  for (uint64_t p = j->lo; p < j->hi; p++) {
    uint8_t *dst = j->out + p * ORC_PAGE;
    const uint8_t *src = page_source(p, j->src_packed, j->n_slots, j->ref, j->n_pages, NULL, 0,
                                     NULL, NULL, 0);
    if (j->ref[p] >> 29 == ORC_KIND_ZERO || !src) {
      memset(dst, 0, ORC_PAGE);
      if (j->ref[p] >> 29 != ORC_KIND_ZERO) j->bad++;
      continue;
    }
    memcpy(dst, src, ORC_PAGE);
    if (j->verify && orc_crc32c_hw(0, dst, ORC_PAGE) != j->crc[p]) j->bad++;
  }
  return NULL;
}
static void mt_run(void *(*fn)(void *), mt_job *proto, uint64_t n_pages, int nt) {
  if (nt < 1) nt = 1;
  if (nt > 256) nt = 256;
  pthread_t th[256];
  mt_job jobs[256];
  for (int i = 0; i < nt; i++) {
    jobs[i] = *proto;
    jobs[i].lo = n_pages * i / nt;
    jobs[i].hi = n_pages * (i + 1) / nt;
    jobs[i].bad = 0;
    pthread_create(&th[i], NULL, fn, &jobs[i]);
  }
  proto->bad = 0;
  for (int i = 0; i < nt; i++) {
    pthread_join(th[i], NULL);
    proto->bad += jobs[i].bad;
  }
}

int orc_pack_mt(const uint8_t *image, uint64_t n_pages, int n_threads, uint32_t *ref,
                uint32_t *crc, uint64_t *hash, uint8_t *packed, orc_pack_stats *stats) {
  if (n_pages > ORC_MAX_PAGES) return -1;
  pthread_once(&nh_once, nh_init);
  uint8_t *zero = (uint8_t *)malloc(n_pages ? n_pages : 1);
  if (!zero) return -1;
  mt_job j;
  memset(&j, 0, sizeof j);
  j.image = image;
  j.crc = crc;
  j.hash = hash;
  j.zero = zero;
  mt_run(mt_analyze, &j, n_pages, n_threads);
  orc_pack_stats st;
  int rc = resolve_refs(zero, crc, hash, n_pages, NULL, NULL, 0, ref, &st);
  free(zero);
  if (rc) return rc;
  j.ref = ref;
  j.packed = packed;
  mt_run(mt_copy, &j, n_pages, n_threads);
  if (stats) *stats = st;
  return 0;
}

uint64_t orc_restore_mt(const uint8_t *packed, uint64_t n_slots, const uint32_t *ref,
                        const uint32_t *crc, uint64_t n_pages, int n_threads, uint8_t *out,
                        int verify) {
  mt_job j;
  memset(&j, 0, sizeof j);
  j.src_packed = packed;
  j.n_slots = n_slots;
  j.ref = ref;
  j.crc = (uint32_t *)crc;
  j.n_pages = n_pages;
  j.out = out;
  j.verify = verify;
  mt_run(mt_restore, &j, n_pages, n_threads);
  return j.bad;
}

/* ------------------------------------------------------------ synthetic image
 * Page g (global index) draws r = sm64(seed ^ g*GOLD):
 *   r % 100 <  pct_zero            -> zero page
 *   r % 100 <  pct_zero + pct_dup  -> copy d = (r >> 32) % dup_pool of the pool
 *   otherwise                      -> unique page
 * A non-zero page with key k holds the 512 little-endian u64 values
 * sm64(base + i), base = sm64(seed ^ sm64(k)); k = g for unique pages and
 * 2^63 | d for pool pages.
 */
void orc_synth_image(uint8_t *image, uint64_t n_pages, uint64_t page_base, uint64_t seed,
                     uint32_t pct_zero, uint32_t pct_dup, uint32_t dup_pool) {
  for (uint64_t p = 0; p < n_pages; p++) {
    uint64_t g = page_base + p, r = sm64(seed ^ (g * GOLD));
    uint32_t cls = (uint32_t)(r % 100);
    uint8_t *pg = image + p * ORC_PAGE;
    if (cls < pct_zero) {
      memset(pg, 0, ORC_PAGE);
      continue;
    }
    uint64_t key = g;
    if (cls < pct_zero + pct_dup && dup_pool) key = 1ull << 63 | ((r >> 32) % dup_pool);
    uint64_t base = sm64(seed ^ sm64(key));
    for (int i = 0; i < 512; i++) {
      uint64_t v = sm64(base + (uint64_t)i);
      for (int b = 0; b < 8; b++) pg[8 * i + b] = (uint8_t)(v >> (8 * b));
    }
  }
}
