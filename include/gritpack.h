/*
 * gritpack -- C ABI of libgritpack.so, the B200 (sm_100a) replacement for the
 * grit-agent data mover.
 *
 * The reference moves a pod's checkpoint image directory between the node and
 * the checkpoint PV with a plain recursive copy:
 *     pkg/gritagent/copy/copy.go:17   TransferData(ctx, srcDir, dstDir) error
 *     pkg/gritagent/copy/copy.go:92   CreateSentinelFile(dir, fileName) error
 * called from pkg/gritagent/checkpoint/checkpoint.go:20 (upload) and
 * pkg/gritagent/restore/restore.go:16,20 (download, then the sentinel).
 * It has no cgo, no FFI and no plugin registry: a maintainer swaps the two
 * calls for the entry points below through the cgo stub in INTEGRATION.md.
 *
 * Conventions: every call returns 0 on success and a negative GRIT_E* code on
 * failure, with a human-readable message copied into err[errlen] when err is
 * not NULL (the bridge wraps it in errors.New).  Plain pointers and sizes only.
 * "d_" pointers are CUDA device pointers on the context's device, "h_" are
 * host pointers.  `stream` is a cudaStream_t passed as void* (NULL = the CUDA
 * default stream): the kernels are ordered after the work already enqueued on
 * it; calls that return host-visible results synchronise it before returning.
 */
#ifndef GRITPACK_H
#define GRITPACK_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GRIT_ABI_VERSION 1
#define GRIT_PAGE_SIZE 4096u
#define GRIT_MAX_PAGES (1u << 26) /* 256 GiB per image; a B200 holds 180 GB */
#define GRIT_MAX_PEERS 8u

#define GRIT_OK 0
#define GRIT_EINVAL (-1)  /* bad argument                                      */
#define GRIT_ECUDA (-2)   /* CUDA runtime error, or no usable sm_100 device    */
#define GRIT_EIO (-3)     /* file system error                                 */
#define GRIT_ECRC (-4)    /* restore: a page failed CRC-32C or a bad reference */
#define GRIT_EFORMAT (-5) /* packed image is not one this library wrote        */
#define GRIT_ENOMEM (-6)

/* Page reference: kind in bits 31..29, value in bits 28..0. */
#define GRIT_KIND_ZERO 0u   /* all-zero page, nothing stored                   */
#define GRIT_KIND_DATA 1u   /* value = slot (4 KiB units) in the packed pages  */
#define GRIT_KIND_DUP 2u    /* value = earlier page of this image (a DATA page)*/
#define GRIT_KIND_PARENT 3u /* value = page index in the parent image          */
#define GRIT_KIND_PEER 4u   /* value = rank << 26 | page index on that rank    */
#define GRIT_REF_KIND(r) ((uint32_t)(r) >> 29)
#define GRIT_REF_VALUE(r) ((uint32_t)(r) & 0x1FFFFFFFu)

typedef struct grit_ctx grit_ctx;

/* Per-page hashes of another image this one may refer to. */
typedef struct grit_table {
  const uint64_t *hash; /* hash64 per logical page */
  const uint32_t *crc;  /* CRC-32C per logical page */
  uint64_t n_pages;
} grit_table;

typedef struct grit_pack_stats {
  uint64_t n_pages, n_zero, n_data, n_dup, n_parent, n_peer;
  double gpu_ms;      /* device time of the kernels of this call (CUDA events) */
  double stage_ms[3]; /* grit_pack_buffer: K1, K2, K3 separately (0 elsewhere)  */
  uint64_t wrote_through; /* 1: K1 stored the pages to their own slots (dense image) */
} grit_pack_stats;

typedef struct grit_restore_stats {
  uint64_t n_pages, n_bad, first_bad; /* first_bad = UINT64_MAX when n_bad == 0 */
  double gpu_ms;
} grit_restore_stats;

/* Options of the directory-level calls.  Zero-initialise, then set fields. */
typedef struct grit_opts {
  uint32_t struct_size; /* sizeof(grit_opts), for ABI growth                   */
  int32_t device;       /* CUDA device ordinal                                 */
  uint32_t verify;      /* restore: recompute and compare every page's CRC-32C */
  uint32_t io_threads;  /* file reader/writer threads, 0 = default (8)         */
  uint64_t chunk_bytes; /* pipeline chunk, 0 = by image size (4..16 MiB)      */
  const char *parent_dir; /* previous packed checkpoint for incremental, or NULL */
  /* Cross-rank dedup of a multi-GPU job (grit_pack_dir only): the per-page tables
   * of the ranks below this one, HOST pointers, as an all-gather of the ranks'
   * K1 results or the lower ranks' hashes.bin / pagemap.bin provide them, and the
   * directories those ranks pack into on the shared PV (recorded in the manifest
   * so that grit_restore_dir can follow PEER references).  Lowest rank owns a page. */
  uint32_t n_peers;              /* 0..GRIT_MAX_PEERS                              */
  const grit_table *peer_tables; /* [n_peers]                                      */
  const char *const *peer_dirs;  /* [n_peers]                                      */
  /* grit_pack_dir of one rank of a job, all ranks at once (collective): instead of
   * peer_tables, a communicator made with grit_comm_create.  The call runs on the
   * communicator's context (`device` is ignored); the tree goes through the GPU
   * twice -- analysed, the tables exchanged over NCCL, then packed against the lower
   * ranks' tables, which never leave the device -- so the host link carries it
   * twice and the PV stores what no lower rank stores.  n_peers / peer_dirs still
   * name the packed directories of the ranks below this one (n_peers = rank).   */
  struct grit_comm *comm;
} grit_opts;

typedef struct grit_stats {
  uint64_t bytes_logical; /* bytes of the file tree                            */
  uint64_t bytes_stored;  /* bytes written to (pack) or read from (restore) the PV */
  uint64_t n_files, n_dirs;
  uint64_t n_pages, n_zero, n_data, n_dup, n_parent, n_peer, n_bad;
  double seconds;     /* wall clock of the call                                */
  double gpu_ms;      /* device time in kernels                                */
  double h2d_bytes, d2h_bytes;
} grit_stats;

/* ---- library ---------------------------------------------------------- */

/* GRIT_ABI_VERSION of the loaded library. */
int grit_version(void);

/* Go: crc32.Update(crc, crc32.MakeTable(crc32.Castagnoli), data).  Host code,
 * used for manifest checksums; page CRCs are computed on the GPU. */
uint32_t grit_crc32c(uint32_t crc, const void *data, size_t len);

/* Creates the per-device state (streams, GF(2) tables, scratch).  Fails with
 * GRIT_ECUDA when the device is missing: there is no CPU fallback.  The
 * reference keeps no state between calls (copy.go:17-64); a context is the
 * explicit create/destroy pair SURVEY 8(b) "Ownership" allows. */
int grit_ctx_create(int device, grit_ctx **out, char *err, size_t errlen);
void grit_ctx_destroy(grit_ctx *ctx);

/* Brings the CUDA runtime up on `device` (primary context, this library's device
 * code) and returns; safe to call from any thread, any number of times.  A
 * process's first CUDA call costs 0.3-0.5 s on a B200 box whatever it does: an
 * agent calls this on a thread of its own at start-up -- before
 * RuntimeCheckpointPod on upload (checkpoint.go:14), so that the cost hides behind
 * the CRIU dump -- and the directory calls find the device ready.  The reference
 * has no such start-up cost and no counterpart. */
int grit_warmup(int device);

/* Number of kernels this context has launched (bench.py's gpu_launches). */
uint64_t grit_ctx_launches(const grit_ctx *ctx);

/* Options of a context, by name.  "verify_dup_bytes" (0 / 1, default 0): grit_pack_buffer compares every page that
 * became a reference to an earlier page of the same image -- on the strength of the 96-bit key (hash64, CRC-32C) -- with
 * that page byte for byte before it drops it, and stores it after all if the bytes differ; grit_ctx_dup_mismatches
 * counts such pages since the context was made.  References to a parent or to a lower rank cannot be checked this way
 * (only their tables are there at pack time).  Unknown names return GRIT_EINVAL.  No reference counterpart. */
int grit_ctx_set_option(grit_ctx *ctx, const char *name, uint64_t value);
uint64_t grit_ctx_dup_mismatches(grit_ctx *ctx);

/* ---- page kernels on device buffers ----------------------------------
 * These have no reference counterpart: copy.go:79 (io.Copy) moves opaque
 * bytes.  They are the new work north_star places behind TransferData. */

/* K1: per 4 KiB page CRC-32C (== Go hash/crc32 Castagnoli of the page),
 * hash64, and d_ref[p] = 0 (GRIT_KIND_ZERO) for an all-zero page, 0xFFFFFFFF
 * otherwise.  d_image must be 16-byte aligned. */
int grit_analyze_pages(grit_ctx *ctx, const void *d_image, uint64_t n_pages,
                       uint32_t *d_crc, uint64_t *d_hash, uint32_t *d_ref,
                       void *stream, char *err, size_t errlen);

/* K1 for a caller that will run K3 next on the same image with d_packed as the
 * target (grit_compact_pages, slot_base 0): same results as grit_analyze_pages,
 * and if a 4096-page sample of the image holds no zero page, every page is also
 * stored to d_packed at its own index, up to the first zero page met
 * (capacity_pages bounds the stores).  The context remembers the two buffers;
 * the next grit_compact_pages on them leaves in place every page whose slot is
 * its own index, which for a dense image (weights, optimizer state) is all of
 * them: the image is then read once and written once.  Any other analyze, pack
 * or compact call on the context forgets the buffers.  No reference counterpart. */
int grit_analyze_pages_through(grit_ctx *ctx, const void *d_image,
                               uint64_t n_pages, uint32_t *d_crc,
                               uint64_t *d_hash, uint32_t *d_ref,
                               void *d_packed, uint64_t capacity_pages,
                               void *stream, char *err, size_t errlen);

/* K2: dedup.  Input: the arrays K1 wrote.  Output: d_ref[p] becomes ZERO, DUP,
 * PARENT, PEER, or 0xFFFFFFFE for a page that must be stored.  The owner of a
 * hash64 is the lowest (parent, peer rank 0.., this image) x page index; a
 * page only refers to its owner when the CRC-32C matches too.  parent and
 * peers[0..n_peers) hold DEVICE pointers; either may be NULL/0. */
int grit_dedup_pages(grit_ctx *ctx, uint64_t n_pages, const uint32_t *d_crc,
                     const uint64_t *d_hash, uint32_t *d_ref,
                     const grit_table *parent, const grit_table *peers,
                     uint32_t n_peers, void *stream, char *err, size_t errlen);

/* Between K2 and K3, for callers who want more than the 96-bit key: every page K2 turned into a reference to an earlier
 * page of the same image (GRIT_KIND_DUP) is compared with that page byte for byte; where the bytes differ d_ref[p] goes
 * back to 0xFFFFFFFE (store it) and *n_mismatch counts it.  grit_pack_buffer does this itself when the context option
 * "verify_dup_bytes" is set.  No reference counterpart. */
int grit_verify_dup_pages(grit_ctx *ctx, const void *d_image, uint64_t n_pages,
                          uint32_t *d_ref, uint64_t *n_mismatch, void *stream,
                          char *err, size_t errlen);

/* K3: stream compaction.  Pages marked 0xFFFFFFFE get consecutive slots
 * slot_base, slot_base+1, ... in page order (warp ballot + decoupled look-back
 * prefix scan) and are copied to d_packed + slot * 4096; d_ref[p] becomes
 * DATA|slot.  *n_stored receives the number of pages stored. */
int grit_compact_pages(grit_ctx *ctx, const void *d_image, uint64_t n_pages,
                       uint32_t *d_ref, void *d_packed, uint64_t capacity_pages,
                       uint64_t slot_base, uint64_t *n_stored, void *stream,
                       char *err, size_t errlen);

/* K3 for callers that must not allocate.  It has no counterpart in the reference.
 * Same inputs, outputs and results as grit_compact_pages, but no grit_ctx and no
 * cudaMalloc, cudaMallocHost or cudaFree, on any call: the scan's working memory
 * (one status word per 256-page tile plus a counter block, 137 KiB in all) is a
 * fixed array inside the library image.  Meant for a packer that runs while the
 * checkpointed job's allocator holds nearly all of HBM (an allocation could
 * fail, or force that allocator to give memory back), for callers that cannot
 * take the device-wide synchronisation cudaMalloc/cudaFree imply, and for host
 * runtimes that hand out a fixed set of buffers and forbid more.
 *   - n_pages above GRIT_STATIC_MAX_PAGES (16 GiB of image) is refused with
 *     GRIT_EINVAL; use grit_compact_pages for larger images.
 *   - The first call on a device looks up the array's device address
 *     (cudaGetSymbolAddress) and reads two device attributes; make that first
 *     call before starting any stream capture.  The CUDA runtime loads device
 *     code lazily: the array comes into being on a GPU at that first look-up and
 *     the kernel at its first launch, and the driver allocates device memory
 *     for both then -- so the first call is also the one that can still run out
 *     of memory.  Call grit_compact_pages_static_warmup(device) once at
 *     start-up, while memory is plentiful: it does the look-up and a one-page
 *     compaction on buffers inside the array, after which no call on that
 *     device causes any allocation, the driver's included.  From then on the
 *     GPU holds the array until the process exits.
 *   - All calls on a device share the one array.  Each call clears what it needs
 *     on `stream` before the kernel, so calls issued one after another are
 *     independent; two calls whose GPU work overlaps (different streams or
 *     threads) race on it and give wrong results -- serialise them.
 *   - Blocks until the work on `stream` is done (it reports *n_stored, read back
 *     through a small device-to-host copy).  The caller's current device is
 *     left as it was.  A device ordinal the process does not have is GRIT_EINVAL. */
#define GRIT_STATIC_MAX_PAGES (1u << 22)
int grit_compact_pages_static(int device, const void *d_image, uint64_t n_pages,
                              uint32_t *d_ref, void *d_packed,
                              uint64_t capacity_pages, uint64_t slot_base,
                              uint64_t *n_stored, void *stream, char *err,
                              size_t errlen);
/* Start-up companion of grit_compact_pages_static (no reference counterpart):
 * makes the scratch array and the compaction kernel resident on `device` by
 * compacting one page held in the array itself.  Uses the default stream and
 * blocks until done.  Idempotent; one call per device that will be used. */
int grit_compact_pages_static_warmup(int device, char *err, size_t errlen);

/* K1+K2+K3 on one device-resident image; the unit bench.py times as `value`.
 * For a full image (no parent, no peers) K1 first samples 4096 pages: if none of
 * them is a zero page the image is taken to be dense (weights, optimizer state:
 * compaction is then the identity) and K1 stores every page to its own slot of
 * d_packed while analysing it, up to the first zero page it meets; K3 leaves
 * pages that are already in place where they are.  The result is the same bytes
 * either way; stats->wrote_through says which way it went. */
int grit_pack_buffer(grit_ctx *ctx, const void *d_image, uint64_t n_pages,
                     const grit_table *parent, const grit_table *peers,
                     uint32_t n_peers, void *d_packed, uint64_t capacity_pages,
                     uint32_t *d_ref, uint32_t *d_crc, uint64_t *d_hash,
                     grit_pack_stats *stats, void *stream, char *err,
                     size_t errlen);

/* ---- cross-rank dedup of a multi-GPU job ----------------------------------
 * One process per GPU, each packs its own rank's image; the ranks exchange only
 * their per-page (hash64, CRC-32C) tables, 12 bytes per 4 KiB page, with NCCL
 * over NVLink inside the library.  The reference has no
 * counterpart: it runs one agent Job per pod and has no cross-pod step
 * (docs/proposals/20250221-gpu-workload-checkpointing-and-restoration.md:63-65).
 * NCCL (libnccl.so.2) is loaded at the first of these calls; a process that
 * never makes one does not need it.  A job has at most GRIT_MAX_PEERS ranks. */
#define GRIT_COMM_ID_BYTES 128 /* sizeof(ncclUniqueId) */
typedef struct grit_comm grit_comm;

/* Rank 0 makes the rendezvous id (ncclGetUniqueId) and hands its 128 bytes to the
 * other ranks by any channel it has (a file on the shared PV, a Job annotation,
 * an environment variable, MPI, torch.distributed). */
int grit_comm_new_id(void *id, char *err, size_t errlen);
/* Collective over all ranks: joins rank `rank` of `world` on the context's GPU. */
int grit_comm_create(grit_ctx *ctx, int rank, int world, const void *id,
                     grit_comm **out, char *err, size_t errlen);
void grit_comm_destroy(grit_comm *comm);
int grit_comm_rank(const grit_comm *comm);
int grit_comm_world(const grit_comm *comm);

/* grit_pack_buffer for one rank of a job (collective: every rank calls it, each
 * with its own image).  K1 runs in slices; each slice's hashes are all-gathered
 * on a second stream while K1 works on the next one, straight into the device
 * buffers K2 probes; pages a lower rank already holds become PEER references
 * (lowest rank owns a page, the parent before any rank).  Ranks may hold
 * different page counts: max_pages is the largest of them if the caller knows
 * it (the same value on every rank), or 0 to have the ranks agree on it first,
 * which costs one small exchange and a host synchronisation.
 * stats->stage_ms: K1 | exchange wait + K2 | K3. */
int grit_pack_buffer_ranked(grit_comm *comm, const void *d_image,
                            uint64_t n_pages, uint64_t max_pages,
                            const grit_table *parent, void *d_packed,
                            uint64_t capacity_pages, uint32_t *d_ref,
                            uint32_t *d_crc, uint64_t *d_hash,
                            grit_pack_stats *stats, void *stream, char *err,
                            size_t errlen);
/* ncclBroadcast of a device buffer from rank `root`, asynchronous on `stream`:
 * how the ranks hand each other restored images (or any other bytes) without a
 * second communication library, e.g. the lower ranks' images that
 * grit_restore_buffer needs in order to follow PEER references. */
int grit_comm_broadcast(grit_comm *comm, void *d_buf, uint64_t bytes, int root,
                        void *stream, char *err, size_t errlen);

/* K4: inverse scatter.  Every page of d_out is written: zeros, the packed slot,
 * the owner's slot, or the page of the parent / peer image (device pointers,
 * fully restored).  With verify != 0 the CRC-32C of every non-zero page is
 * recomputed from the bytes read and compared with d_crc; mismatches and
 * references that cannot be followed are counted in stats and make the call
 * return GRIT_ECRC (d_out is still fully written). */
int grit_restore_buffer(grit_ctx *ctx, const void *d_packed, uint64_t n_slots,
                        const uint32_t *d_ref, const uint32_t *d_crc,
                        uint64_t n_pages, const void *d_parent_image,
                        uint64_t parent_pages, const void *const *d_peer_images,
                        const uint64_t *peer_pages, uint32_t n_peers,
                        void *d_out, int verify, grit_restore_stats *stats,
                        void *stream, char *err, size_t errlen);

/* K5 (bench/tests only): seeded synthetic image, page classes zero / pool
 * duplicate / unique; same generator as the oracle's. */
int grit_synth_image(grit_ctx *ctx, void *d_image, uint64_t n_pages,
                     uint64_t page_base, uint64_t seed, uint32_t pct_zero,
                     uint32_t pct_dup, uint32_t dup_pool, void *stream,
                     char *err, size_t errlen);

/* ---- host buffers through the pinned, double-buffered pipeline --------
 * What copyFile (copy.go:66-90) is to one file: bytes in host memory in,
 * bytes in host memory out, H2D and D2H copies overlapped with the kernels.
 * parent holds HOST pointers here. */
int grit_pack_host(grit_ctx *ctx, const void *h_image, uint64_t n_pages,
                   const grit_table *parent, void *h_packed,
                   uint64_t capacity_pages, uint32_t *h_ref, uint32_t *h_crc,
                   uint64_t *h_hash, grit_pack_stats *stats, char *err,
                   size_t errlen);
int grit_restore_host(grit_ctx *ctx, const void *h_packed, uint64_t n_slots,
                      const uint32_t *h_ref, const uint32_t *h_crc,
                      uint64_t n_pages, const void *h_parent_image,
                      uint64_t parent_pages, void *h_out, int verify,
                      grit_restore_stats *stats, char *err, size_t errlen);
/* Restore straight into a device buffer, for images that fill the GPU (BASELINE
 * configs[4]: download + CRC verify + reassembly of a full-HBM image).  d_out,
 * n_pages * 4096 bytes, is the only large device allocation: the packed pages
 * are taken from host memory (h_packed: the whole packed image, pinned for full
 * PCIe rate) or from `read` (called from a helper thread, in slot order, to fill
 * a pinned staging buffer with slots [slot0, slot0 + n_slots); non-zero return
 * aborts) and pass through a ring of four chunks on the device (chunk_bytes
 * each, 0 = 64 MiB).  Slots must be numbered in page order, as every pack of
 * this library numbers them.  The restore of a duplicate copies its owner's page
 * from d_out.  h_ref / h_crc are host arrays; the parent and peer images are
 * DEVICE pointers as in grit_restore_buffer.  What copyFile's destination is to
 * the reference (copy.go:73-79) d_out is here: the consumer's memory itself,
 * with no file in between. */
typedef int (*grit_read_packed_fn)(void *user, uint64_t slot0, uint64_t n_slots,
                                   void *dst_pinned);
int grit_restore_to_device(grit_ctx *ctx, const void *h_packed,
                           grit_read_packed_fn read, void *user,
                           uint64_t n_slots, const uint32_t *h_ref,
                           const uint32_t *h_crc, uint64_t n_pages,
                           const void *d_parent_image, uint64_t parent_pages,
                           const void *const *d_peer_images,
                           const uint64_t *peer_pages, uint32_t n_peers,
                           void *d_out, int verify, uint64_t chunk_bytes,
                           grit_restore_stats *stats, char *err, size_t errlen);
/* cudaHostRegister / Unregister of a caller buffer, so that the two calls
 * above copy at full PCIe rate. */
int grit_pin_host(void *h_ptr, size_t bytes);
int grit_unpin_host(void *h_ptr);

/* ---- directories: the drop-in for TransferData ------------------------ */

/* Upload direction of TransferData (checkpoint.go:20): packs the tree under
 * src_dir into dst_dir on the PV as manifest + page map + unique pages.
 * Overwrites a previous (partial) pack of the same dst_dir, like os.Create
 * (copy.go:73); the manifest is written last. */
int grit_pack_dir(const char *src_dir, const char *dst_dir,
                  const grit_opts *opts, grit_stats *stats, char *err,
                  size_t errlen);

/* Download direction of TransferData (restore.go:16): recreates under dst_dir
 * the tree grit_pack_dir saw -- same relative paths, file bytes and mode bits,
 * directories 0777 & ~umask (copy.go:36,84-89).  A src_dir without a gritpack
 * manifest (a plain mirror uploaded by the stock agent, or a pack that did not
 * finish) is refused with GRIT_EFORMAT and nothing is written: the caller hands
 * plain mirrors to the stock TransferData (integration/copy_gritpack.go does).
 * Does not write the sentinel. */
int grit_restore_dir(const char *src_dir, const char *dst_dir,
                     const grit_opts *opts, grit_stats *stats, char *err,
                     size_t errlen);

/* Checks a packed directory and writes nothing: manifest and page-map checksums
 * on the host, then the whole restore pipeline (parents included) with every
 * stored page's CRC-32C recomputed on the GPU and the restored bytes dropped.
 * Returns GRIT_ECRC / GRIT_EFORMAT like grit_restore_dir would.  The reference
 * never verifies what it copied (copy.go:66-90); this is the "gritpack verify"
 * step of SURVEY 8(f). */
int grit_verify_dir(const char *src_dir, const grit_opts *opts,
                    grit_stats *stats, char *err, size_t errlen);

/* Which other packed directories a packed directory needs in order to be
 * restored: its parent (the previous checkpoint of an incremental chain) and
 * the lower ranks its page map refers to, resolved the way grit_restore_dir
 * resolves them; with recursive != 0 also what those need, each directory once.
 * Host work only (manifests and page maps are read, no GPU is touched).  The
 * list is written to out as NUL-terminated paths followed by an empty string;
 * *needed receives the bytes that takes, and GRIT_EINVAL is returned (nothing
 * written) when outlen is smaller.  This is what a garbage collector of old
 * checkpoints consults before deleting a directory from the PV;
 * the reference never deletes or links checkpoints, so there is no counterpart. */
int grit_image_deps(const char *dir, int recursive, char *out, size_t outlen,
                    size_t *needed, char *err, size_t errlen);

/* ---- a packed checkpoint served page by page -----------------
 * The reference restores the whole tree to local disk before the runtime may
 * start (restore.go:16-20; patched containerd waits for the sentinel,
 * contrib/containerd/grit-interceptor.diff:146-165) and CRIU then reads it all
 * again (cmd/containerd-shim-grit-v1/process/init_state.go:163).  A grit_image
 * keeps the packed checkpoint open instead: grit_image_open returns once the
 * metadata and the images it refers to (parent, lower ranks) are in place and
 * streams the stored pages into HBM in the background; a read waits only for
 * the slots it needs, restores and CRC-checks the pages on the GPU and hands
 * them over -- to host memory (a lazy-pages server answering CRIU's faults, a
 * FUSE view of the tree) or to device memory (device state that need not touch
 * the host again).  opts: device, verify, io_threads.  Reads are serialised. */
typedef struct grit_image grit_image;
int grit_image_open(const char *packed_dir, const grit_opts *opts,
                    grit_image **out, char *err, size_t errlen);
void grit_image_close(grit_image *img);
/* size, mode (with the file-type bits) and first stream page of a path of the tree */
int grit_image_stat(grit_image *img, const char *relpath, uint64_t *size,
                    uint32_t *mode, uint64_t *first_page, char *err, size_t errlen);
/* pread on a file of the tree: returns the bytes delivered (0 at or past the
 * end) or a negative GRIT_E* code. */
int64_t grit_image_read(grit_image *img, const char *relpath, uint64_t offset,
                        uint64_t len, void *h_dst, char *err, size_t errlen);
/* pages [page0, page0 + n_pages) of the page stream into host memory, or with
 * dst_on_device into memory of the image's GPU. */
int grit_image_read_pages(grit_image *img, uint64_t page0, uint64_t n_pages,
                          void *dst, int dst_on_device, char *err, size_t errlen);
/* how many stored pages have reached the GPU so far (n_slots when all have) */
uint64_t grit_image_resident_slots(grit_image *img);

/* ---- CRIU-aware ingest ---------------------------------------
 * The tree the agent uploads is a CRIU image directory written at
 * pkg/gritagent/checkpoint/runtime.go:122-134; the reference copies it as
 * opaque files.  grit_pack_dir parses every pagemap-<id>.img it finds (CRIU's
 * published image format; CRIU is not part of the reference tree) and records
 * beside the packed image, in criu-extents.txt, which pages of the page stream
 * belong to which virtual-address run of which pages-<id>.img.
 *
 * grit_criu_incremental turns an incremental pack into what CRIU itself calls
 * an incremental image: for every page map of the tree src_dir (which was
 * packed into packed_dir with a parent) it writes, under out_dir at the same
 * relative paths, a pagemap-<id>.img whose runs are marked in_parent wherever
 * the GPU found the page unchanged at the SAME virtual address in the parent
 * (packed reference PARENT to exactly that page), and a pages-<id>.img holding
 * only the remaining pages.  Together with a `parent` link to the previous
 * image directory that is a native CRIU parent chain -- the link the shim
 * plumbs but nobody sets (cmd/containerd-shim-grit-v1/runc/container.go:107,
 * process/init.go:192).  Host work only; the GPU's part is the page map
 * packed_dir already holds.  The other files of the image are not rewritten. */
typedef struct grit_criu_stats {
  uint64_t n_pagemaps, entries_in, entries_out;
  uint64_t pages_total, pages_in_parent; /* pages of present runs; those left to the parent */
  uint64_t bytes_written;
} grit_criu_stats;
/* What a pagemap-<id>.img says: the pages file it belongs to, its runs, how
 * many pages are present in that file and how many are left to a parent.
 * GRIT_EFORMAT when the file is not a CRIU page map.  Host work only. */
int grit_criu_pagemap_info(const char *path, uint32_t *pages_id,
                           uint64_t *n_entries, uint64_t *pages_present,
                           uint64_t *pages_in_parent, char *err, size_t errlen);
int grit_criu_incremental(const char *src_dir, const char *packed_dir,
                          const char *out_dir, grit_criu_stats *stats,
                          char *err, size_t errlen);

/* CreateSentinelFile (copy.go:92-102): dir/name holding "Transfer Completed\n". */
int grit_write_sentinel(const char *dir, const char *name);

#ifdef __cplusplus
}
#endif
#endif /* GRITPACK_H */
