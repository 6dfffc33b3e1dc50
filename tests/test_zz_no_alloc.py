"""-m gpu, collected last: grit_compact_pages_static, the compaction entry point that allocates nothing.

Some callers of this library may not allocate device memory at all: a packer that runs while the frozen job's
allocator owns almost all of HBM, code that cannot take the device-wide synchronisation of cudaMalloc/cudaFree,
runtimes that hand out a fixed set of buffers.  For them the scan's working memory is a fixed array inside the
library, shared by every call on the device.  This file checks, through the C ABI, that the entry point gives
exactly what grit_compact_pages gives -- page references, stored count and packed bytes -- on two calls in a row
(the second on a smaller, different image, so stale state from the first call would show), and that an image
beyond the fixed limit is refused with a clear message.  It makes its own context and inputs, so it passes when
run alone in a fresh process as well as after the other tests.
"""
import ctypes as C

import pytest
import torch

import oracle_lib as O
from grit_b200 import _lib
from grit_b200.buffers import Context

pytestmark = pytest.mark.gpu
PAGE = 4096


def compact_static(image, ref, packed, slot_base=0):
    lib = _lib.load()
    stored, err = C.c_uint64(0), _lib.errbuf()
    rc = lib.grit_compact_pages_static(0, C.c_void_p(image.data_ptr()), image.numel() // PAGE,
                                       C.c_void_p(ref.data_ptr()), C.c_void_p(packed.data_ptr()),
                                       packed.numel() // PAGE, slot_base, C.byref(stored),
                                       C.c_void_p(torch.cuda.current_stream().cuda_stream), err, len(err))
    return rc, int(stored.value), err.value.decode()


def warmup(device=0):
    lib = _lib.load()
    err = _lib.errbuf()
    return lib.grit_compact_pages_static_warmup(device, err, len(err)), err.value.decode()


def test_static_scratch_compaction_matches_the_allocating_entry_point():
    assert torch.cuda.is_available()
    rc, msg = warmup()
    assert rc == 0, msg
    ctx = Context(0)
    # two calls in a row: a larger image first, then a smaller one with another page mix
    for n_pages, seed, mix in [(70_000, 11, dict(pct_zero=10, pct_dup=15, dup_pool=64)),
                               (9_001, 12, dict(pct_zero=40, pct_dup=30, dup_pool=5))]:
        image = torch.from_numpy(O.synth_image(n_pages, seed, **mix)).cuda()
        crc, h, ref = ctx.analyze(image)
        ctx.dedup(crc, h, ref)                       # ref now marks the pages that must be stored
        ref_a, ref_b = ref.clone(), ref.clone()
        packed_a = torch.zeros(n_pages * PAGE, dtype=torch.uint8, device="cuda")
        packed_b = torch.zeros(n_pages * PAGE, dtype=torch.uint8, device="cuda")
        n_a = ctx.compact(image, ref_a, packed_a)    # existing entry point: scratch from cudaMalloc
        rc, n_b, msg = compact_static(image, ref_b, packed_b)
        assert rc == 0, msg
        assert n_b == n_a and n_a > 0
        assert torch.equal(ref_b, ref_a)             # bit-exact, as in every other parity test here
        assert torch.equal(packed_b, packed_a)
        # and first hand against the oracle's pack of the same image
        o = O.pack(image.cpu().numpy())
        assert n_b == o["stats"]["n_data"]
        assert (ref_b.cpu().numpy().view("uint32") == o["ref"]).all()
        assert (packed_b[: n_b * PAGE].cpu().numpy() == o["packed"]).all()
    ctx.close()


def test_static_scratch_calls_allocate_nothing_after_the_warm_up():
    """The property the entry point exists for: once the warm-up has made array and kernel resident, a call leaves
    the device's free memory exactly where it was.  Also the corners: slots that continue a previous chunk
    (slot_base > 0), an empty image, a packed buffer that is too small, a device the process does not have."""
    rc, msg = warmup()
    assert rc == 0, msg
    ctx = Context(0)
    n_pages = 20_000
    image = torch.from_numpy(O.synth_image(n_pages, 21, pct_zero=20, pct_dup=10, dup_pool=16)).cuda()
    crc, h, ref = ctx.analyze(image)
    ctx.dedup(crc, h, ref)
    packed_a = torch.zeros((n_pages + 7) * PAGE, dtype=torch.uint8, device="cuda")
    packed_b = torch.zeros_like(packed_a)
    ref_a, ref_b, ref_c, ref_d = (ref.clone() for _ in range(4))
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    rc, n_b, msg = compact_static(image, ref_b, packed_b, slot_base=7)
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info()
    assert rc == 0, msg
    assert free1 == free0, f"the call changed free device memory by {free0 - free1} bytes"
    n_a = ctx.compact(image, ref_a, packed_a, slot_base=7)
    assert n_b == n_a and torch.equal(ref_b, ref_a) and torch.equal(packed_b, packed_a)
    assert int(ref_b.cpu().numpy().view("uint32")[ref_b.cpu().numpy().view("uint32") >> 29 == 1].min()) == (1 << 29) + 7
    # an empty image stores nothing and is not an error
    rc, n0, msg = compact_static(image[:0], ref_c[:0], packed_b)
    assert rc == 0 and n0 == 0, msg
    # a packed buffer that cannot take the stored pages is refused, like grit_compact_pages refuses it
    rc, _, msg = compact_static(image, ref_d, packed_b[: (n_a // 2) * PAGE])
    assert rc == _lib.EINVAL and "too small" in msg
    # a device ordinal this process does not have is a bad argument, not a CUDA failure
    lib = _lib.load()
    stored, err = C.c_uint64(0), _lib.errbuf()
    rc = lib.grit_compact_pages_static(torch.cuda.device_count() + 3, C.c_void_p(image.data_ptr()), 1,
                                       C.c_void_p(ref_c.data_ptr()), C.c_void_p(packed_b.data_ptr()), 1, 0,
                                       C.byref(stored), None, err, len(err))
    assert rc == _lib.EINVAL and b"out of range" in err.value
    assert torch.cuda.current_device() == 0
    ctx.close()


def test_static_scratch_rejects_images_beyond_its_fixed_limit():
    tiny = torch.zeros(PAGE, dtype=torch.uint8, device="cuda")
    ref = torch.zeros(1, dtype=torch.int32, device="cuda")
    lib = _lib.load()
    stored, err = C.c_uint64(0), _lib.errbuf()
    too_many = (1 << 22) + 1
    rc = lib.grit_compact_pages_static(0, C.c_void_p(tiny.data_ptr()), too_many, C.c_void_p(ref.data_ptr()),
                                       C.c_void_p(tiny.data_ptr()), 1, 0, C.byref(stored), None, err, len(err))
    assert rc == _lib.EINVAL
    assert b"GRIT_STATIC_MAX_PAGES" in err.value
