"""ctypes binding of oracle/liboracle.so -- the CPU checker (test infrastructure).

Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs import this.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
ORACLE_DIR = os.path.join(os.path.dirname(_HERE), "oracle")
PAGE = 4096
KIND_ZERO, KIND_DATA, KIND_DUP, KIND_PARENT, KIND_PEER = range(5)


class Table(C.Structure):
    _fields_ = [("hash", C.c_void_p), ("crc", C.c_void_p), ("n_pages", C.c_uint64)]


class PackStats(C.Structure):
    _fields_ = [(n, C.c_uint64) for n in ("n_pages", "n_zero", "n_data", "n_dup", "n_parent", "n_peer")]

    def as_dict(self):
        return {n: int(getattr(self, n)) for n, _ in self._fields_}


def build():
    subprocess.run(["make", "-s", "-C", ORACLE_DIR], check=True)


def load():
    path = os.path.join(ORACLE_DIR, "liboracle.so")
    if not os.path.exists(path):
        build()
    lib = C.CDLL(path)
    lib.orc_crc32c.restype = C.c_uint32
    lib.orc_crc32c.argtypes = [C.c_uint32, C.c_void_p, C.c_size_t]
    lib.orc_crc32c_hw.restype = C.c_uint32
    lib.orc_crc32c_hw.argtypes = [C.c_uint32, C.c_void_p, C.c_size_t]
    lib.orc_hash64_page.restype = C.c_uint64
    lib.orc_hash64_page.argtypes = [C.c_void_p]
    lib.orc_analyze.restype = None
    lib.orc_analyze.argtypes = [C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p, C.c_void_p]
    lib.orc_pack.restype = C.c_int
    lib.orc_pack.argtypes = [C.c_void_p, C.c_uint64, C.POINTER(Table), C.POINTER(Table), C.c_uint32,
                             C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(PackStats)]
    lib.orc_restore.restype = C.c_uint64
    lib.orc_restore.argtypes = [C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p,
                                C.c_uint64, C.POINTER(C.c_void_p), C.POINTER(C.c_uint64), C.c_uint32,
                                C.c_void_p, C.c_int, C.POINTER(C.c_uint64)]
    lib.orc_pack_mt.restype = C.c_int
    lib.orc_pack_mt.argtypes = [C.c_void_p, C.c_uint64, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p,
                                C.c_void_p, C.POINTER(PackStats)]
    lib.orc_restore_mt.restype = C.c_uint64
    lib.orc_restore_mt.argtypes = [C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p, C.c_uint64, C.c_int,
                                   C.c_void_p, C.c_int]
    lib.orc_synth_image.restype = None
    lib.orc_synth_image.argtypes = [C.c_void_p, C.c_uint64, C.c_uint64, C.c_uint64, C.c_uint32,
                                    C.c_uint32, C.c_uint32]
    lib.orc_transfer_data.restype = C.c_int
    lib.orc_transfer_data.argtypes = [C.c_char_p, C.c_char_p, C.c_char_p, C.c_size_t]
    lib.orc_create_sentinel_file.restype = C.c_int
    lib.orc_create_sentinel_file.argtypes = [C.c_char_p, C.c_char_p]
    return lib


_lib = None


def lib():
    global _lib
    if _lib is None:
        _lib = load()
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


def crc32c(data, crc=0, hw=False):
    b = bytes(data)
    f = lib().orc_crc32c_hw if hw else lib().orc_crc32c
    return int(f(crc, b, len(b)))


def hash64_page(page):
    b = bytes(page)
    assert len(b) == PAGE
    return int(lib().orc_hash64_page(b))


def synth_image(n_pages, seed, pct_zero=10, pct_dup=15, dup_pool=1024, page_base=0):
    img = np.empty(n_pages * PAGE, dtype=np.uint8)
    lib().orc_synth_image(_p(img), n_pages, page_base, seed, pct_zero, pct_dup, dup_pool)
    return img


def analyze(image):
    image = np.ascontiguousarray(image, dtype=np.uint8)
    n = image.size // PAGE
    crc = np.empty(n, np.uint32)
    h = np.empty(n, np.uint64)
    z = np.empty(n, np.uint8)
    lib().orc_analyze(_p(image), n, _p(crc), _p(h), _p(z))
    return crc, h, z


def _table(t):
    if t is None:
        return None
    h, c = t
    h = np.ascontiguousarray(h, np.uint64)
    c = np.ascontiguousarray(c, np.uint32)
    tab = Table(h.ctypes.data, c.ctypes.data, h.size)
    tab._keep = (h, c)
    return tab


def pack(image, parent=None, peers=(), threads=0):
    """-> dict(ref, crc, hash, packed, stats).  parent/peers entries are (hash, crc) arrays."""
    image = np.ascontiguousarray(image, dtype=np.uint8)
    n = image.size // PAGE
    ref = np.empty(n, np.uint32)
    crc = np.empty(n, np.uint32)
    h = np.empty(n, np.uint64)
    packed = np.empty(max(n, 1) * PAGE, np.uint8)
    st = PackStats()
    if threads:
        assert parent is None and not peers
        rc = lib().orc_pack_mt(_p(image), n, threads, _p(ref), _p(crc), _p(h), _p(packed), C.byref(st))
    else:
        ptab = _table(parent)
        peer_tabs = [_table(t) for t in peers]
        arr = (Table * max(len(peer_tabs), 1))(*peer_tabs)
        rc = lib().orc_pack(_p(image), n, C.byref(ptab) if ptab else None, arr, len(peer_tabs),
                            _p(ref), _p(crc), _p(h), _p(packed), C.byref(st))
    assert rc == 0
    return dict(ref=ref, crc=crc, hash=h, packed=packed[: st.n_data * PAGE], stats=st.as_dict())


def restore(packed, ref, crc, parent_image=None, peer_images=(), verify=True):
    """-> (image, n_bad, first_bad)"""
    packed = np.ascontiguousarray(packed, np.uint8)
    ref = np.ascontiguousarray(ref, np.uint32)
    crc = np.ascontiguousarray(crc, np.uint32)
    n = ref.size
    out = np.empty(n * PAGE, np.uint8)
    fb = C.c_uint64(0)
    npeer = len(peer_images)
    imgs = (C.c_void_p * max(npeer, 1))(*[(_p(i) if i is not None else None) for i in peer_images])
    cnts = (C.c_uint64 * max(npeer, 1))(*[(i.size // PAGE if i is not None else 0) for i in peer_images])
    bad = lib().orc_restore(_p(packed), packed.size // PAGE, _p(ref), _p(crc), n,
                            _p(parent_image), (parent_image.size // PAGE) if parent_image is not None else 0,
                            imgs, cnts, npeer, _p(out), int(verify), C.byref(fb))
    return out, int(bad), int(fb.value)


def transfer_data(src, dst):
    err = C.create_string_buffer(4096)
    rc = lib().orc_transfer_data(os.fsencode(src), os.fsencode(dst), err, len(err))
    return rc, err.value.decode()


def create_sentinel_file(d, name):
    return lib().orc_create_sentinel_file(os.fsencode(d), os.fsencode(name))
