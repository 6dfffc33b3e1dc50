"""CRIU page-map images, written and read independently of grit_b200/csrc (test infrastructure).

Format as CRIU publishes it (criu.org/Images): u32 0x54564319, u32 0x56084025, then records {u32 size, protobuf}:
pagemap_head {pages_id = 1}, then pagemap_entry {vaddr = 1, nr_pages = 2, in_parent = 3, flags = 4};
pages-<pages_id>.img holds the raw pages of the entries that are not in_parent, in entry order.
"""
import struct

PAGE = 4096
PE_PARENT, PE_LAZY, PE_PRESENT = 1, 2, 4


def _varint(v):
    out = bytearray()
    while v >= 0x80:
        out.append((v & 0x7F) | 0x80)
        v >>= 7
    out.append(v)
    return bytes(out)


def _read_varint(b, i):
    v = shift = 0
    while True:
        c = b[i]
        i += 1
        v |= (c & 0x7F) << shift
        shift += 7
        if not c & 0x80:
            return v, i


def encode_pagemap(pages_id, entries, unknown_field=False):
    """entries: [(vaddr, nr_pages, in_parent, flags or None)]"""
    out = struct.pack("<II", 0x54564319, 0x56084025)
    msg = b"\x08" + _varint(pages_id)
    out += struct.pack("<I", len(msg)) + msg
    for vaddr, nr, in_parent, flags in entries:
        msg = b"\x08" + _varint(vaddr) + b"\x10" + _varint(nr)
        if in_parent:
            msg += b"\x18\x01"
        if flags is not None:
            msg += b"\x20" + _varint(flags)
        if unknown_field:                      # a field a later CRIU might add: length-delimited, number 9
            msg += b"\x4a\x03abc"
        out += struct.pack("<I", len(msg)) + msg
    return out


def decode_pagemap(b):
    assert struct.unpack_from("<II", b) == (0x54564319, 0x56084025)
    i, recs = 8, []
    while i < len(b):
        (size,) = struct.unpack_from("<I", b, i)
        i += 4
        end, f = i + size, {}
        while i < end:
            key, i = _read_varint(b, i)
            if key & 7 == 0:
                f[key >> 3], i = _read_varint(b, i)
            elif key & 7 == 2:
                n, i = _read_varint(b, i)
                i += n
            else:
                raise ValueError("unexpected wire type")
        recs.append(f)
    pages_id = recs[0][1]
    return pages_id, [(f[1], f[2], bool(f.get(3, 0)), f.get(4)) for f in recs[1:]]


def pages_by_vaddr(pagemap_bytes, pages_bytes, parent=None):
    """{vaddr: page bytes} of an image; in_parent runs are taken from `parent` (a dict of the same kind)."""
    _, entries = decode_pagemap(pagemap_bytes)
    out, off = {}, 0
    for vaddr, nr, in_parent, flags in entries:
        for k in range(nr):
            va = vaddr + k * PAGE
            if in_parent or (flags is not None and flags & PE_PARENT):
                assert parent is not None and va in parent, hex(va)
                out[va] = parent[va]
            else:
                out[va] = pages_bytes[off:off + PAGE]
                off += PAGE
    assert off == len(pages_bytes)
    return out
