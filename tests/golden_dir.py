"""The directory-level golden fixture (tests/golden/packed_base, tests/golden/packed_child): the tree it was packed
from, and an independent reader of the on-PV format as DESIGN.md section 2 specifies it.

The fixture is written by the CUDA path on a B200 (tests/make_golden_dir.py); the reader below is plain Python and
shares no code with grit_b200, so the CPU tests that unpack the fixture with it check the format against its
specification, and the GPU test that packs the same tree again checks that the format has not drifted."""
import hashlib
import os
import struct

PAGE = 4096
MAGIC = b"GRITPAK1"
VERSION = 3
KIND_ZERO, KIND_DATA, KIND_DUP, KIND_PARENT, KIND_PEER = range(5)


def _stream(tag, n):
    """n deterministic bytes (SHA-256 in counter mode): no dependence on numpy's generators."""
    out, i = bytearray(), 0
    while len(out) < n:
        out += hashlib.sha256(f"{tag}:{i}".encode()).digest()
        i += 1
    return bytes(out[:n])


def golden_files(dirty=False):
    """{relative path: (bytes, mode)} -- the cases SURVEY 8(c) asks a format fixture to hold: an all-zero file, two
    identical files, zero and duplicate pages inside a file, ragged and empty files, and (dirty) one page changed
    against the parent."""
    rnd = [_stream(f"page{i}", PAGE) for i in range(16)]
    pages1 = rnd[0:8] + [bytes(PAGE)] * 4 + rnd[0:4] + rnd[8:16]
    if dirty:
        pages1[5] = _stream("dirty", PAGE)
    return {
        "main/checkpoint/pages-1.img": (b"".join(pages1), 0o600),
        "main/checkpoint/pages-2.img": (b"".join(rnd[0:8]), 0o600),
        "main/checkpoint/zeros.img": (bytes(3 * PAGE), 0o644),
        "main/checkpoint/inventory.img": (_stream("inventory", 77), 0o600),
        "main/checkpoint/empty.img": (b"", 0o444),
        "main/container.log": (b"2025-01-01T00:00:00Z stdout F step\n" * 143, 0o640),
        "main/rootfs-diff.tar": (_stream("tar", 2 * PAGE + 100), 0o644),
    }


GOLDEN_DIRS = ["main", "main/checkpoint", "main/work"]


def build_golden_tree(root, dirty=False):
    for d in GOLDEN_DIRS:
        os.makedirs(os.path.join(root, d), exist_ok=True)
        os.chmod(os.path.join(root, d), 0o755)
    for rel, (data, mode) in golden_files(dirty).items():
        p = os.path.join(root, rel)
        with open(p, "wb") as f:
            f.write(data)
        os.chmod(p, mode)


class PackedDir:
    """gritpack.manifest + pagemap.bin + hashes.bin + pages-NNN.bin, parsed as DESIGN.md section 2 lays them out."""

    def __init__(self, d, crc32c):
        self.dir = d
        raw = open(os.path.join(d, "gritpack.manifest"), "rb").read()
        assert raw[:8] == MAGIC
        assert struct.unpack_from("<I", raw, len(raw) - 4)[0] == crc32c(raw[:-4]), "manifest checksum"
        off = 8
        (version, n_entries, self.n_pages, self.n_slots, self.bytes_logical, self.pagemap_crc, self.hashes_crc,
         self.n_shards, self.stripe) = struct.unpack_from("<IIQQQIIII", raw, off)
        off += struct.calcsize("<IIQQQIIII")
        assert version == VERSION

        def string():
            nonlocal off
            (n,) = struct.unpack_from("<I", raw, off)
            s = raw[off + 4: off + 4 + n]
            off += 4 + n
            return s.decode()

        self.parent = string()
        (n_peers,) = struct.unpack_from("<I", raw, off)
        off += 4
        self.peers = [string() for _ in range(n_peers)]
        self.entries = []
        for _ in range(n_entries):
            typ, mode, size, first_page = struct.unpack_from("<BIQQ", raw, off)
            off += struct.calcsize("<BIQQ")
            self.entries.append((typ, mode, size, first_page, string()))
        assert off + 4 == len(raw)
        pm = open(os.path.join(d, "pagemap.bin"), "rb").read()
        assert len(pm) == 8 * self.n_pages and crc32c(pm) == self.pagemap_crc
        self.ref = list(struct.unpack_from(f"<{self.n_pages}I", pm, 0))
        self.crc = list(struct.unpack_from(f"<{self.n_pages}I", pm, 4 * self.n_pages))
        hs = open(os.path.join(d, "hashes.bin"), "rb").read()
        assert len(hs) == 8 * self.n_pages and crc32c(hs) == self.hashes_crc
        self.hash = list(struct.unpack(f"<{self.n_pages}Q", hs))
        self.shards = [open(os.path.join(d, f"pages-{i:03d}.bin"), "rb").read() for i in range(self.n_shards)]
        assert sum(len(s) for s in self.shards) == self.n_slots * PAGE

    def slot(self, s):
        """Slot s is in stripe t = s // stripe, which is stripe t // k of file t % k."""
        t, k = s // self.stripe, self.n_shards
        o = ((t // k) * self.stripe + s % self.stripe) * PAGE
        return self.shards[t % k][o: o + PAGE]

    def image(self, parent_image=None):
        """The page stream, every reference followed."""
        pages = []
        for p, r in enumerate(self.ref):
            kind, val = r >> 29, r & ((1 << 29) - 1)
            if kind == KIND_ZERO:
                pages.append(bytes(PAGE))
            elif kind == KIND_DATA:
                pages.append(self.slot(val))
            elif kind == KIND_DUP:
                assert val < p
                pages.append(pages[val])
            elif kind == KIND_PARENT:
                pages.append(parent_image[val * PAGE: (val + 1) * PAGE])
            else:
                raise AssertionError(f"page {p}: unexpected reference kind {kind}")
        return b"".join(pages)

    def files(self, parent_image=None):
        """{path: (bytes, mode)} and the list of directories, as a restore would recreate them."""
        img = self.image(parent_image)
        files, dirs = {}, []
        for typ, mode, size, first_page, path in self.entries:
            if typ == 0:
                dirs.append(path)
            else:
                files[path] = (img[first_page * PAGE: first_page * PAGE + size], mode & 0o7777)
        return files, dirs
