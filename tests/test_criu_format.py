"""CPU: the library's CRIU page-map parser against hand-built images (tests/criu_img.py writes them independently)."""
import ctypes as C

import pytest

import criu_img
from criu_img import PE_LAZY, PE_PARENT, PE_PRESENT
from grit_b200 import _lib


def info(path):
    lib = _lib.load()
    pid, n, present, inpar, err = C.c_uint32(), C.c_uint64(), C.c_uint64(), C.c_uint64(), _lib.errbuf()
    rc = lib.grit_criu_pagemap_info(str(path).encode(), C.byref(pid), C.byref(n), C.byref(present), C.byref(inpar), err, len(err))
    return rc, (pid.value, n.value, present.value, inpar.value), err.value.decode()


@pytest.mark.parametrize("unknown", [False, True])
def test_pagemap_images_are_read_like_criu_writes_them(tmp_path, unknown):
    entries = [(0x7F0000000000, 900, False, PE_PRESENT), (0x7F0010000000, 300, True, PE_PARENT),
               (0xFFFFFFFFFF600000, 1, False, None),                  # a 64-bit address, no flags field (older CRIU)
               (0x560000001000, 41, False, PE_PRESENT | PE_LAZY), (0x10000, 70000, False, PE_PARENT)]
    p = tmp_path / "pagemap-77.img"
    p.write_bytes(criu_img.encode_pagemap(1 << 20, entries, unknown_field=unknown))
    rc, got, msg = info(p)
    assert rc == 0, msg
    assert got == (1 << 20, 5, 900 + 1 + 41, 300 + 70000)
    assert criu_img.decode_pagemap(p.read_bytes()) == (1 << 20, entries)


def test_what_is_not_a_pagemap_is_refused(tmp_path):
    good = criu_img.encode_pagemap(3, [(0x1000, 5, False, PE_PRESENT)])
    cases = {"empty": b"", "other-magic": b"\x19\x43\x56\x54\x16\x31\x31\x58" + good[8:], "truncated": good[:-2],
             "no-head": good[:8], "garbage": bytes(range(200)), "size-overrun": good[:8] + b"\xff\xff\x00\x00" + good[12:]}
    for name, data in cases.items():
        p = tmp_path / name
        p.write_bytes(data)
        rc, _, msg = info(p)
        assert rc == _lib.EFORMAT, name
    assert info(tmp_path / "missing")[0] == _lib.EIO
