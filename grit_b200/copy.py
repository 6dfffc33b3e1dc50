"""Host-side mirror of the reference package pkg/gritagent/copy (copy.go).

Same names and argument meaning as the Go functions; the work is done by
libgritpack.so.  The reference's TransferData is direction-blind (it mirrors
src into dst whichever way the agent Job swapped them,
pkg/gritmanager/agentmanager/manager.go:124-134); the packed format is not, so
the direction is explicit here: `UPLOAD` packs a host tree into the PV
(checkpoint.go:20), `DOWNLOAD` rebuilds the tree from the PV (restore.go:16).
"""
import ctypes as C
import os

from . import _lib
from ._lib import GritError, Opts, Stats

UPLOAD, DOWNLOAD = "upload", "download"


def _opts(device=0, verify=True, io_threads=0, chunk_bytes=0, parent_dir=None, peers=(), comm=None, peer_dirs=()):
    """peers: [(hash64 array, crc32c array, packed_dir)] of the ranks below this one, in rank order -- numpy arrays
    (uint64 / uint32), e.g. from grit_b200.dist.all_gather_tables or read_tables() of the lower ranks' directories."""
    import numpy as np
    o = Opts()
    o.struct_size = C.sizeof(Opts)
    o.device = device
    o.verify = 1 if verify else 0
    o.io_threads = io_threads
    o.chunk_bytes = chunk_bytes
    o.parent_dir = os.fsencode(parent_dir) if parent_dir else None
    if peers:
        keep = []
        tabs = (_lib.Table * len(peers))()
        dirs = (C.c_char_p * len(peers))()
        for i, (h, crc, d) in enumerate(peers):
            h = np.ascontiguousarray(h, dtype=np.uint64)
            crc = np.ascontiguousarray(crc, dtype=np.uint32)
            keep += [h, crc]
            tabs[i] = _lib.Table(h.ctypes.data, crc.ctypes.data, h.size)
            dirs[i] = os.fsencode(d)
        o.n_peers, o.peer_tables, o.peer_dirs = len(peers), tabs, dirs
        o._keep = (keep, tabs, dirs)
    if comm is not None:
        # a ranked upload (grit_b200.dist.Comm): the tables travel over NCCL inside the library; peer_dirs names the
        # packed directories of the ranks below this one, in rank order
        dirs = (C.c_char_p * max(len(peer_dirs), 1))(*[os.fsencode(d) for d in peer_dirs])
        o.comm, o.n_peers, o.peer_dirs = comm.handle, len(peer_dirs), dirs
        o._keep = (dirs, comm)
    return o


def read_tables(packed_dir):
    """(hash64, crc32c) per page of a packed directory, from its hashes.bin / pagemap.bin."""
    import numpy as np
    h = np.fromfile(os.path.join(packed_dir, "hashes.bin"), np.uint64)
    pm = np.fromfile(os.path.join(packed_dir, "pagemap.bin"), np.uint32)
    return h, pm[h.size:]


def TransferData(src_dir, dst_dir, direction, **kw):
    """copy.go:17 TransferData(ctx, srcDir, dstDir) error -- raises GritError instead of returning error.

    Returns the grit_stats of the call as a dict.
    """
    lib = _lib.load()
    if direction not in (UPLOAD, DOWNLOAD):
        raise ValueError(f"unknown direction {direction!r}")
    fn = lib.grit_pack_dir if direction == UPLOAD else lib.grit_restore_dir
    o, st, err = _opts(**kw), Stats(), _lib.errbuf()
    _lib.check(fn(os.fsencode(src_dir), os.fsencode(dst_dir), C.byref(o), C.byref(st), err, len(err)), err)
    return st.as_dict()


def CreateSentinelFile(dir, file_name):
    """copy.go:92 CreateSentinelFile(dir, fileName) error."""
    rc = _lib.load().grit_write_sentinel(os.fsencode(dir), os.fsencode(file_name))
    if rc != 0:
        raise GritError(rc, f"cannot create {os.path.join(dir, file_name)}")


def ImageDependencies(packed_dir, recursive=True):
    """The packed directories `packed_dir` needs in order to be restored (grit_image_deps): its parent and the lower
    ranks its page map refers to, and with `recursive` what those need.  Host work only.  What a garbage collector
    of old checkpoints on the PV consults before deleting one."""
    lib = _lib.load()
    need, err = C.c_size_t(0), _lib.errbuf()
    rc = lib.grit_image_deps(os.fsencode(packed_dir), int(bool(recursive)), None, 0, C.byref(need), err, len(err))
    if rc != _lib.EINVAL or need.value == 0:
        _lib.check(rc, err)
    buf = C.create_string_buffer(need.value)
    _lib.check(lib.grit_image_deps(os.fsencode(packed_dir), int(bool(recursive)), buf, need.value, C.byref(need), err, len(err)), err)
    return [os.fsdecode(x) for x in buf.raw[: need.value - 1].split(b"\0") if x]


def VerifyData(packed_dir, **kw):
    """Checks a packed checkpoint on the PV without restoring it (grit_verify_dir): checksums of the metadata,
    CRC-32C of every stored page recomputed on the GPU.  Raises GritError (ECRC / EFORMAT) on damage."""
    lib = _lib.load()
    o, st, err = _opts(**kw), Stats(), _lib.errbuf()
    _lib.check(lib.grit_verify_dir(os.fsencode(packed_dir), C.byref(o), C.byref(st), err, len(err)), err)
    return st.as_dict()


def CriuIncremental(src_dir, packed_dir, out_dir):
    """grit_criu_incremental: rewrites the pagemap-*.img / pages-*.img pairs of the CRIU image directory `src_dir`
    (packed into `packed_dir` against a parent) as a native CRIU incremental image under `out_dir`: runs the GPU found
    unchanged at the same virtual address in the parent become in_parent, only the other pages are kept."""
    lib = _lib.load()
    st, err = _lib.CriuStats(), _lib.errbuf()
    _lib.check(lib.grit_criu_incremental(os.fsencode(src_dir), os.fsencode(packed_dir), os.fsencode(out_dir), C.byref(st),
                                         err, len(err)), err)
    return st.as_dict()


class PackedImage:
    """grit_image: a packed checkpoint kept open and served on demand (what a lazy-pages server or a FUSE view of
    the checkpoint would sit on).  read() is pread on a file of the tree; read_pages() fills a CUDA tensor."""

    def __init__(self, packed_dir, **kw):
        self.lib = _lib.load()
        o, h, err = _opts(**kw), C.c_void_p(), _lib.errbuf()
        _lib.check(self.lib.grit_image_open(os.fsencode(packed_dir), C.byref(o), C.byref(h), err, len(err)), err)
        self.handle = h

    def close(self):
        if getattr(self, "handle", None):
            self.lib.grit_image_close(self.handle)
            self.handle = None

    __del__ = close

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def stat(self, relpath):
        size, mode, first, err = C.c_uint64(), C.c_uint32(), C.c_uint64(), _lib.errbuf()
        _lib.check(self.lib.grit_image_stat(self.handle, os.fsencode(relpath), C.byref(size), C.byref(mode), C.byref(first),
                                            err, len(err)), err)
        return dict(size=size.value, mode=mode.value, first_page=first.value)

    def read(self, relpath, offset, length):
        buf, err = C.create_string_buffer(max(length, 1)), _lib.errbuf()
        n = self.lib.grit_image_read(self.handle, os.fsencode(relpath), offset, length, buf, err, len(err))
        if n < 0:
            _lib.check(int(n), err)
        return buf.raw[:n]

    def read_pages(self, page0, n_pages, out):
        """out: a CUDA uint8 tensor on the image's device, or a CPU uint8 tensor."""
        err = _lib.errbuf()
        _lib.check(self.lib.grit_image_read_pages(self.handle, page0, n_pages, C.c_void_p(out.data_ptr()),
                                                  1 if out.is_cuda else 0, err, len(err)), err)
        return out

    @property
    def resident_slots(self):
        return int(self.lib.grit_image_resident_slots(self.handle))
