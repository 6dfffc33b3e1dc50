"""ctypes binding of grit_b200/csrc/libgritpack.so (the C ABI in include/gritpack.h).

There is no fallback: if the shared library is missing or the device is not a
B200-class GPU, the calls raise.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("GRITPACK_LIB") or os.path.join(_HERE, "csrc", "libgritpack.so")  # override: kernel A/B builds

PAGE = 4096
MAX_PAGES = 1 << 26
KIND_ZERO, KIND_DATA, KIND_DUP, KIND_PARENT, KIND_PEER = range(5)
OK, EINVAL, ECUDA, EIO, ECRC, EFORMAT, ENOMEM = 0, -1, -2, -3, -4, -5, -6


class GritError(RuntimeError):
    def __init__(self, code, message):
        super().__init__(f"gritpack error {code}: {message}")
        self.code = code
        self.message = message


class Table(C.Structure):
    _fields_ = [("hash", C.c_void_p), ("crc", C.c_void_p), ("n_pages", C.c_uint64)]


class PackStats(C.Structure):
    _fields_ = [(n, C.c_uint64) for n in ("n_pages", "n_zero", "n_data", "n_dup", "n_parent", "n_peer")] + [
        ("gpu_ms", C.c_double), ("stage_ms", C.c_double * 3), ("wrote_through", C.c_uint64)]

    def as_dict(self):
        d = {n: getattr(self, n) for n, _ in self._fields_ if n != "stage_ms"}
        d["stage_ms"] = [float(x) for x in self.stage_ms]
        return d


class RestoreStats(C.Structure):
    _fields_ = [("n_pages", C.c_uint64), ("n_bad", C.c_uint64), ("first_bad", C.c_uint64), ("gpu_ms", C.c_double)]

    def as_dict(self):
        return {n: getattr(self, n) for n, _ in self._fields_}


class Opts(C.Structure):
    _fields_ = [("struct_size", C.c_uint32), ("device", C.c_int32), ("verify", C.c_uint32),
                ("io_threads", C.c_uint32), ("chunk_bytes", C.c_uint64), ("parent_dir", C.c_char_p),
                ("n_peers", C.c_uint32), ("peer_tables", C.POINTER(Table)), ("peer_dirs", C.POINTER(C.c_char_p)),
                ("comm", C.c_void_p)]


class Stats(C.Structure):
    _fields_ = [(n, C.c_uint64) for n in ("bytes_logical", "bytes_stored", "n_files", "n_dirs", "n_pages", "n_zero",
                                          "n_data", "n_dup", "n_parent", "n_peer", "n_bad")] + [
        (n, C.c_double) for n in ("seconds", "gpu_ms", "h2d_bytes", "d2h_bytes")]

    def as_dict(self):
        return {n: getattr(self, n) for n, _ in self._fields_}


class CriuStats(C.Structure):
    _fields_ = [(n, C.c_uint64) for n in ("n_pagemaps", "entries_in", "entries_out", "pages_total", "pages_in_parent",
                                          "bytes_written")]

    def as_dict(self):
        return {n: getattr(self, n) for n, _ in self._fields_}


# every symbol include/gritpack.h declares: name -> (restype, argtypes)
_ERR = [C.c_char_p, C.c_size_t]
_P, _U64, _U32, _I = C.c_void_p, C.c_uint64, C.c_uint32, C.c_int
SYMBOLS = {
    "grit_version": (_I, []),
    "grit_crc32c": (_U32, [_U32, _P, C.c_size_t]),
    "grit_ctx_create": (_I, [_I, C.POINTER(_P)] + _ERR),
    "grit_ctx_destroy": (None, [_P]),
    "grit_ctx_launches": (_U64, [_P]),
    "grit_warmup": (_I, [_I]),
    "grit_ctx_set_option": (_I, [_P, C.c_char_p, _U64]),
    "grit_ctx_dup_mismatches": (_U64, [_P]),
    "grit_analyze_pages": (_I, [_P, _P, _U64, _P, _P, _P, _P] + _ERR),
    "grit_analyze_pages_through": (_I, [_P, _P, _U64, _P, _P, _P, _P, _U64, _P] + _ERR),
    "grit_dedup_pages": (_I, [_P, _U64, _P, _P, _P, C.POINTER(Table), C.POINTER(Table), _U32, _P] + _ERR),
    "grit_verify_dup_pages": (_I, [_P, _P, _U64, _P, C.POINTER(_U64), _P] + _ERR),
    "grit_compact_pages": (_I, [_P, _P, _U64, _P, _P, _U64, _U64, C.POINTER(_U64), _P] + _ERR),
    "grit_compact_pages_static": (_I, [_I, _P, _U64, _P, _P, _U64, _U64, C.POINTER(_U64), _P] + _ERR),
    "grit_compact_pages_static_warmup": (_I, [_I] + _ERR),
    "grit_pack_buffer": (_I, [_P, _P, _U64, C.POINTER(Table), C.POINTER(Table), _U32, _P, _U64, _P, _P, _P,
                              C.POINTER(PackStats), _P] + _ERR),
    "grit_comm_new_id": (_I, [_P] + _ERR),
    "grit_comm_create": (_I, [_P, _I, _I, _P, C.POINTER(_P)] + _ERR),
    "grit_comm_destroy": (None, [_P]),
    "grit_comm_rank": (_I, [_P]),
    "grit_comm_world": (_I, [_P]),
    "grit_pack_buffer_ranked": (_I, [_P, _P, _U64, _U64, C.POINTER(Table), _P, _U64, _P, _P, _P, C.POINTER(PackStats),
                                     _P] + _ERR),
    "grit_comm_broadcast": (_I, [_P, _P, _U64, _I, _P] + _ERR),
    "grit_restore_buffer": (_I, [_P, _P, _U64, _P, _P, _U64, _P, _U64, C.POINTER(_P), C.POINTER(_U64), _U32, _P, _I,
                                 C.POINTER(RestoreStats), _P] + _ERR),
    "grit_synth_image": (_I, [_P, _P, _U64, _U64, _U64, _U32, _U32, _U32, _P] + _ERR),
    "grit_pack_host": (_I, [_P, _P, _U64, C.POINTER(Table), _P, _U64, _P, _P, _P, C.POINTER(PackStats)] + _ERR),
    "grit_restore_host": (_I, [_P, _P, _U64, _P, _P, _U64, _P, _U64, _P, _I, C.POINTER(RestoreStats)] + _ERR),
    "grit_restore_to_device": (_I, [_P, _P, _P, _P, _U64, _P, _P, _U64, _P, _U64, C.POINTER(_P), C.POINTER(_U64), _U32, _P,
                                    _I, _U64, C.POINTER(RestoreStats)] + _ERR),
    "grit_pin_host": (_I, [_P, C.c_size_t]),
    "grit_unpin_host": (_I, [_P]),
    "grit_pack_dir": (_I, [C.c_char_p, C.c_char_p, C.POINTER(Opts), C.POINTER(Stats)] + _ERR),
    "grit_restore_dir": (_I, [C.c_char_p, C.c_char_p, C.POINTER(Opts), C.POINTER(Stats)] + _ERR),
    "grit_verify_dir": (_I, [C.c_char_p, C.POINTER(Opts), C.POINTER(Stats)] + _ERR),
    "grit_image_deps": (_I, [C.c_char_p, C.c_int, C.c_char_p, C.c_size_t, C.POINTER(C.c_size_t)] + _ERR),
    "grit_image_open": (_I, [C.c_char_p, C.POINTER(Opts), C.POINTER(_P)] + _ERR),
    "grit_image_close": (None, [_P]),
    "grit_image_stat": (_I, [_P, C.c_char_p, C.POINTER(_U64), C.POINTER(_U32), C.POINTER(_U64)] + _ERR),
    "grit_image_read": (C.c_int64, [_P, C.c_char_p, _U64, _U64, _P] + _ERR),
    "grit_image_read_pages": (_I, [_P, _U64, _U64, _P, _I] + _ERR),
    "grit_image_resident_slots": (_U64, [_P]),
    "grit_criu_pagemap_info": (_I, [C.c_char_p, C.POINTER(_U32), C.POINTER(_U64), C.POINTER(_U64), C.POINTER(_U64)] + _ERR),
    "grit_criu_incremental": (_I, [C.c_char_p, C.c_char_p, C.c_char_p, C.POINTER(CriuStats)] + _ERR),
    "grit_write_sentinel": (_I, [C.c_char_p, C.c_char_p]),
}

READ_PACKED_FN = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_uint64, C.c_uint64, C.c_void_p)

_lib = None


def load():
    """Loads libgritpack.so and types every entry point; raises if it is not built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise GritError(ECUDA, f"{LIB_PATH} is not built; run `make -C grit_b200/csrc` "
                               "(or __graft_entry__.build()); there is no fallback path")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SYMBOLS.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(rc, err):
    if rc != 0:
        raise GritError(rc, err.value.decode(errors="replace"))


def errbuf():
    return C.create_string_buffer(1024)
