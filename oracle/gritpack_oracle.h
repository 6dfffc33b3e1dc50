/*
 * gritpack oracle -- TEST INFRASTRUCTURE ONLY.
 *
 * A scalar CPU restatement of the checkpoint-image pack/restore path, used as
 * the checker for the CUDA implementation.  Nothing under grit_b200/ may
 * include, link or call this; only tests/, __graft_entry__.smoke() and the
 * cpu_baseline / --impl reference legs of bench.py do.
 *
 * What it restates, and what pins it:
 *   - orc_transfer_data / orc_create_sentinel_file (oracle_copy.cpp) restate
 *     /root/reference/pkg/gritagent/copy/copy.go:17-102.  The reference has no
 *     test, fixture or golden vector for that file and no Go toolchain exists
 *     here, so for the tree copy parity is UNPINNED BY THE REFERENCE; the
 *     contract is byte identity + mode bits + the 19-byte sentinel, which
 *     tests/ check directly against the source tree.
 *   - orc_crc32c restates Go's hash/crc32 with crc32.Castagnoli (Go 1.23.4,
 *     go.mod:3; the stdlib is not under /root/reference).  Pinned by the
 *     known-answer vectors of RFC 3720 B.4 and Go's crc32 golden table
 *     (tests/golden/crc32c_kat.json) and by the CPU's own SSE4.2 CRC32
 *     instruction (orc_crc32c_hw), an independent implementation of the same
 *     polynomial.
 *   - hash64 / the packed-image format are new in this repository (the
 *     reference computes nothing on this path); the oracle is their
 *     definition, frozen by tests/golden/pack_golden.json.
 */
#ifndef GRITPACK_ORACLE_H
#define GRITPACK_ORACLE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ORC_PAGE 4096u

/* page reference encoding: kind in bits 31..29, value in bits 28..0 */
#define ORC_KIND_ZERO 0u   /* page is all zero, nothing stored               */
#define ORC_KIND_DATA 1u   /* value = slot in this image's packed pages      */
#define ORC_KIND_DUP 2u    /* value = earlier page index of this image       */
#define ORC_KIND_PARENT 3u /* value = page index in the parent image         */
#define ORC_KIND_PEER 4u   /* value = rank << 26 | page index on that rank   */
#define ORC_MAX_PAGES (1u << 26)

/* Go: crc32.Update(crc, crc32.MakeTable(crc32.Castagnoli), data) */
uint32_t orc_crc32c(uint32_t crc, const void *data, size_t len);
/* same polynomial through the SSE4.2 instruction; second opinion */
uint32_t orc_crc32c_hw(uint32_t crc, const void *data, size_t len);

uint64_t orc_hash64_page(const uint8_t *page);
int orc_page_is_zero(const uint8_t *page);

void orc_analyze(const uint8_t *image, uint64_t n_pages, uint32_t *crc,
                 uint64_t *hash, uint8_t *zero);

typedef struct {
  const uint64_t *hash; /* per logical page of that image */
  const uint32_t *crc;
  uint64_t n_pages;
} orc_table;

typedef struct {
  uint64_t n_pages, n_zero, n_data, n_dup, n_parent, n_peer;
} orc_pack_stats;

/*
 * Canonical sequential pack.  parent may be NULL; peers[0..n_peers) are the
 * hash tables of ranks 0..n_peers-1 (all lower than this rank).  packed must
 * hold n_pages*4096 bytes in the worst case.  Returns 0, or -1 on bad args.
 */
int orc_pack(const uint8_t *image, uint64_t n_pages, const orc_table *parent,
             const orc_table *peers, uint32_t n_peers, uint32_t *ref,
             uint32_t *crc, uint64_t *hash, uint8_t *packed,
             orc_pack_stats *stats);

/*
 * Inverse.  peer_images[r] is rank r's restored image (may be NULL when no
 * page refers to it).  Returns the number of pages that could not be restored
 * or whose CRC did not verify; *first_bad gets the lowest such page.
 */
uint64_t orc_restore(const uint8_t *packed, uint64_t n_slots,
                     const uint32_t *ref, const uint32_t *crc,
                     uint64_t n_pages, const uint8_t *parent_image,
                     uint64_t parent_pages, const uint8_t *const *peer_images,
                     const uint64_t *peer_pages, uint32_t n_peers,
                     uint8_t *out, int verify, uint64_t *first_bad);

/* multithreaded same-work port, used only as a timed CPU baseline */
int orc_pack_mt(const uint8_t *image, uint64_t n_pages, int n_threads,
                uint32_t *ref, uint32_t *crc, uint64_t *hash, uint8_t *packed,
                orc_pack_stats *stats);
uint64_t orc_restore_mt(const uint8_t *packed, uint64_t n_slots,
                        const uint32_t *ref, const uint32_t *crc,
                        uint64_t n_pages, int n_threads, uint8_t *out,
                        int verify);

/* seeded synthetic image, identical to the GPU generator (grit_synth_image) */
void orc_synth_image(uint8_t *image, uint64_t n_pages, uint64_t page_base,
                     uint64_t seed, uint32_t pct_zero, uint32_t pct_dup,
                     uint32_t dup_pool);

/* copy.go restatement (oracle_copy.cpp) */
int orc_transfer_data(const char *src_dir, const char *dst_dir, char *err,
                      size_t errlen);
int orc_create_sentinel_file(const char *dir, const char *file_name);

#ifdef __cplusplus
}
#endif
#endif
