"""Page kernels on device / host buffers, through the C ABI.

torch is used for what it is here for -- device memory, streams, pinned host
memory -- and nothing else: every byte of work is done by libgritpack.so.
"""
import ctypes as C

import torch

from . import _lib
from ._lib import PAGE, GritError, PackStats, RestoreStats, Table


def _ptr(t):
    return C.c_void_p(t.data_ptr()) if t is not None else None


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


class Context:
    """grit_ctx: per-device state (streams, GF(2) tables, scratch)."""

    def __init__(self, device=0):
        self.lib = _lib.load()
        self.device = int(device)
        h = C.c_void_p()
        err = _lib.errbuf()
        _lib.check(self.lib.grit_ctx_create(self.device, C.byref(h), err, len(err)), err)
        self.handle = h

    def close(self):
        if getattr(self, "handle", None):
            self.lib.grit_ctx_destroy(self.handle)
            self.handle = None

    __del__ = close

    def set_option(self, name, value):
        """grit_ctx_set_option, e.g. ("verify_dup_bytes", 1)."""
        if self.lib.grit_ctx_set_option(self.handle, name.encode(), int(value)) != 0:
            raise GritError(_lib.EINVAL, f"unknown context option {name!r}")

    @property
    def dup_mismatches(self):
        return int(self.lib.grit_ctx_dup_mismatches(self.handle))

    @property
    def launches(self):
        return int(self.lib.grit_ctx_launches(self.handle))

    # -- helpers
    def _dev(self):
        return torch.device("cuda", self.device)

    @staticmethod
    def _pages(image):
        assert image.dtype == torch.uint8 and image.is_contiguous() and image.numel() % PAGE == 0
        return image.numel() // PAGE

    def _table(self, t):
        if t is None:
            return None, None
        h, c = t
        assert h.dtype == torch.int64 and c.dtype == torch.int32 and h.numel() == c.numel()
        tab = Table(h.data_ptr(), c.data_ptr(), h.numel())
        return tab, (h, c)

    def _tables(self, peers):
        keep = [self._table(p) for p in peers]
        arr = (Table * max(len(keep), 1))(*[k[0] for k in keep])
        return arr, keep

    # -- K5
    def synth_image(self, n_pages, seed, pct_zero=10, pct_dup=15, dup_pool=1024, page_base=0):
        img = torch.empty(n_pages * PAGE, dtype=torch.uint8, device=self._dev())
        err = _lib.errbuf()
        _lib.check(self.lib.grit_synth_image(self.handle, _ptr(img), n_pages, page_base, seed, pct_zero, pct_dup,
                                             dup_pool, _stream(), err, len(err)), err)
        return img

    # -- K1
    def analyze(self, image, through=None):
        """K1.  `through`: the packed buffer the following compact() will fill -- dense images are then written
        through to it while they are analysed (grit_analyze_pages_through)."""
        n = self._pages(image)
        crc = torch.empty(n, dtype=torch.int32, device=image.device)
        h = torch.empty(n, dtype=torch.int64, device=image.device)
        ref = torch.empty(n, dtype=torch.int32, device=image.device)
        err = _lib.errbuf()
        if through is None:
            rc = self.lib.grit_analyze_pages(self.handle, _ptr(image), n, _ptr(crc), _ptr(h), _ptr(ref), _stream(), err,
                                             len(err))
        else:
            rc = self.lib.grit_analyze_pages_through(self.handle, _ptr(image), n, _ptr(crc), _ptr(h), _ptr(ref),
                                                     _ptr(through), through.numel() // PAGE, _stream(), err, len(err))
        _lib.check(rc, err)
        return crc, h, ref

    # -- K2
    def dedup(self, crc, h, ref, parent=None, peers=()):
        ptab, _k1 = self._table(parent)
        arr, _k2 = self._tables(peers)
        err = _lib.errbuf()
        _lib.check(self.lib.grit_dedup_pages(self.handle, crc.numel(), _ptr(crc), _ptr(h), _ptr(ref),
                                             C.byref(ptab) if ptab else None, arr, len(peers), _stream(), err,
                                             len(err)), err)
        return ref

    def verify_dups(self, image, ref):
        """grit_verify_dup_pages: DUP references whose bytes differ from their owner's go back to "store"; -> how many."""
        n, err = C.c_uint64(0), _lib.errbuf()
        _lib.check(self.lib.grit_verify_dup_pages(self.handle, _ptr(image), self._pages(image), _ptr(ref), C.byref(n), _stream(),
                                                  err, len(err)), err)
        return int(n.value)

    # -- K3
    def compact(self, image, ref, packed, slot_base=0):
        n = self._pages(image)
        stored = C.c_uint64(0)
        err = _lib.errbuf()
        _lib.check(self.lib.grit_compact_pages(self.handle, _ptr(image), n, _ptr(ref), _ptr(packed),
                                               packed.numel() // PAGE, slot_base, C.byref(stored), _stream(), err,
                                               len(err)), err)
        return int(stored.value)

    # -- K1+K2+K3
    def pack(self, image, parent=None, peers=(), packed=None, out=None):
        """-> dict(ref, crc, hash, packed (view of the stored pages), stats)."""
        n = self._pages(image)
        if packed is None:
            packed = torch.empty(max(n, 1) * PAGE, dtype=torch.uint8, device=image.device)
        if out is None:
            out = (torch.empty(n, dtype=torch.int32, device=image.device),
                   torch.empty(n, dtype=torch.int32, device=image.device),
                   torch.empty(n, dtype=torch.int64, device=image.device))
        ref, crc, h = out
        ptab, _k1 = self._table(parent)
        arr, _k2 = self._tables(peers)
        st = PackStats()
        err = _lib.errbuf()
        _lib.check(self.lib.grit_pack_buffer(self.handle, _ptr(image), n, C.byref(ptab) if ptab else None, arr,
                                             len(peers), _ptr(packed), packed.numel() // PAGE, _ptr(ref), _ptr(crc),
                                             _ptr(h), C.byref(st), _stream(), err, len(err)), err)
        return dict(ref=ref, crc=crc, hash=h, packed=packed[: st.n_data * PAGE], stats=st.as_dict())

    # -- K4
    def restore(self, packed, ref, crc, parent_image=None, peer_images=(), verify=True, out=None, check=True):
        n = ref.numel()
        if out is None:
            out = torch.empty(n * PAGE, dtype=torch.uint8, device=ref.device)
        npeer = len(peer_images)
        imgs = (C.c_void_p * max(npeer, 1))(*[(i.data_ptr() if i is not None else None) for i in peer_images])
        cnts = (C.c_uint64 * max(npeer, 1))(*[(i.numel() // PAGE if i is not None else 0) for i in peer_images])
        st = RestoreStats()
        err = _lib.errbuf()
        rc = self.lib.grit_restore_buffer(self.handle, _ptr(packed), packed.numel() // PAGE, _ptr(ref), _ptr(crc), n,
                                          _ptr(parent_image),
                                          parent_image.numel() // PAGE if parent_image is not None else 0, imgs, cnts,
                                          npeer, _ptr(out), int(verify), C.byref(st), _stream(), err, len(err))
        if check:
            _lib.check(rc, err)
        return out, st.as_dict()

    # -- host buffers through the pinned pipeline
    def pack_host(self, image, parent=None, packed=None):
        """image / packed: CPU uint8 tensors (pinned for full PCIe rate)."""
        assert image.device.type == "cpu"
        n = self._pages(image)
        if packed is None:
            packed = torch.empty(max(n, 1) * PAGE, dtype=torch.uint8)
        ref = torch.empty(n, dtype=torch.int32)
        crc = torch.empty(n, dtype=torch.int32)
        h = torch.empty(n, dtype=torch.int64)
        ptab, _k = self._table(parent)
        st = PackStats()
        err = _lib.errbuf()
        _lib.check(self.lib.grit_pack_host(self.handle, _ptr(image), n, C.byref(ptab) if ptab else None, _ptr(packed),
                                           packed.numel() // PAGE, _ptr(ref), _ptr(crc), _ptr(h), C.byref(st), err,
                                           len(err)), err)
        return dict(ref=ref, crc=crc, hash=h, packed=packed[: st.n_data * PAGE], stats=st.as_dict())

    def restore_host(self, packed, ref, crc, parent_image=None, verify=True, out=None, check=True):
        n = ref.numel()
        if out is None:
            out = torch.empty(n * PAGE, dtype=torch.uint8)
        st = RestoreStats()
        err = _lib.errbuf()
        rc = self.lib.grit_restore_host(self.handle, _ptr(packed), packed.numel() // PAGE, _ptr(ref), _ptr(crc), n,
                                        _ptr(parent_image),
                                        parent_image.numel() // PAGE if parent_image is not None else 0, _ptr(out),
                                        int(verify), C.byref(st), err, len(err))
        if check:
            _lib.check(rc, err)
        return out, st.as_dict()


    def restore_to_device(self, packed, ref, crc, out, parent_image=None, peer_images=(), verify=True, chunk_bytes=0,
                          read=None, n_slots=None, check=True):
        """grit_restore_to_device: packed (CPU tensor, or None with `read(slot0, n, dst_ptr) -> int`) streams through a
        device ring into `out` (CUDA tensor of the whole image); ref / crc are CPU tensors."""
        n = ref.numel()
        npeer = len(peer_images)
        imgs = (C.c_void_p * max(npeer, 1))(*[(i.data_ptr() if i is not None else None) for i in peer_images])
        cnts = (C.c_uint64 * max(npeer, 1))(*[(i.numel() // PAGE if i is not None else 0) for i in peer_images])
        cb = _lib.READ_PACKED_FN(lambda user, s0, k, dst: int(read(s0, k, dst))) if read is not None else None
        slots = n_slots if n_slots is not None else packed.numel() // PAGE
        st, err = RestoreStats(), _lib.errbuf()
        rc = self.lib.grit_restore_to_device(
            self.handle, _ptr(packed), C.cast(cb, C.c_void_p) if cb else None, None, slots, _ptr(ref), _ptr(crc), n,
            _ptr(parent_image), parent_image.numel() // PAGE if parent_image is not None else 0, imgs, cnts, npeer,
            _ptr(out), int(verify), chunk_bytes, C.byref(st), err, len(err))
        if check:
            _lib.check(rc, err)
        return out, st.as_dict()


def crc32c(data, crc=0):
    """Host CRC-32C (Go: crc32.Update with the Castagnoli table)."""
    b = bytes(data)
    return int(_lib.load().grit_crc32c(crc, b, len(b)))


__all__ = ["Context", "GritError", "crc32c", "PAGE"]
