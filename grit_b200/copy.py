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


def _opts(device=0, verify=True, io_threads=0, chunk_bytes=0, parent_dir=None):
    o = Opts()
    o.struct_size = C.sizeof(Opts)
    o.device = device
    o.verify = 1 if verify else 0
    o.io_threads = io_threads
    o.chunk_bytes = chunk_bytes
    o.parent_dir = os.fsencode(parent_dir) if parent_dir else None
    return o


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


def VerifyData(packed_dir, **kw):
    """Checks a packed checkpoint on the PV without restoring it (grit_verify_dir): checksums of the metadata,
    CRC-32C of every stored page recomputed on the GPU.  Raises GritError (ECRC / EFORMAT) on damage."""
    lib = _lib.load()
    o, st, err = _opts(**kw), Stats(), _lib.errbuf()
    _lib.check(lib.grit_verify_dir(os.fsencode(packed_dir), C.byref(o), C.byref(st), err, len(err)), err)
    return st.as_dict()
