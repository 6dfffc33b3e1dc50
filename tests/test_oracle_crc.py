"""The oracle's CRC-32C against the golden vectors of RFC 3720 / Go hash/crc32 and the CPU's CRC32 instruction."""
import json
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
KAT = json.load(open(os.path.join(HERE, "golden", "crc32c_kat.json")))
PATTERNS = {"32x00": bytes(32), "32xff": b"\xff" * 32, "00..1f": bytes(range(32)), "1f..00": bytes(range(31, -1, -1)),
            "4096x00": bytes(4096), "4096xff": b"\xff" * 4096}


def test_ascii_vectors(oracle):
    for s, want in KAT["ascii"].items():
        assert oracle.crc32c(s.encode()) == int(want, 16), s
        assert oracle.crc32c(s.encode(), hw=True) == int(want, 16), s


def test_pattern_vectors(oracle):
    for name, want in KAT["patterns"].items():
        assert oracle.crc32c(PATTERNS[name]) == int(want, 16), name
        assert oracle.crc32c(PATTERNS[name], hw=True) == int(want, 16), name


def test_update_is_incremental(oracle):
    """Go's crc32.Update(crc, tab, p) chains: Update(Update(0, a), b) == Checksum(a+b)."""
    data = bytes(range(256)) * 5
    for cut in (0, 1, 7, 255, 1279, 1280):
        assert oracle.crc32c(data[cut:], crc=oracle.crc32c(data[:cut])) == oracle.crc32c(data)


def test_table_agrees_with_cpu_instruction_on_random_pages(oracle):
    rng = np.random.default_rng(1)
    for n in (1, 3, 15, 16, 17, 4095, 4096, 4097, 65536 + 5):
        b = rng.integers(0, 256, n, dtype=np.uint8).tobytes()
        assert oracle.crc32c(b) == oracle.crc32c(b, hw=True)
