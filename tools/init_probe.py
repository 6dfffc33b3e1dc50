"""What the first call of a process pays: CUDA runtime initialisation against grit_ctx_create against a small
pack_dir, in a fresh process without torch.  Development aid (profiles/README.md quotes it)."""
import ctypes as C
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
t0 = time.perf_counter()
from grit_b200 import _lib  # noqa: E402

lib = _lib.load()
t1 = time.perf_counter()
rt = C.CDLL("libcudart.so.12")
rt.cudaFree.argtypes = [C.c_void_p]
rc = rt.cudaFree(None)                       # creates the primary context
t2 = time.perf_counter()
ctx, err = C.c_void_p(), _lib.errbuf()
_lib.check(lib.grit_ctx_create(0, C.byref(ctx), err, len(err)), err)
t3 = time.perf_counter()
lib.grit_ctx_destroy(ctx)
t4 = time.perf_counter()
_lib.check(lib.grit_ctx_create(0, C.byref(ctx), err, len(err)), err)
t5 = time.perf_counter()
lib.grit_ctx_destroy(ctx)
print({"dlopen_s": round(t1 - t0, 3), "cudaFree0_s": round(t2 - t1, 3), "rc": rc, "first_ctx_create_s": round(t3 - t2, 3),
       "ctx_destroy_s": round(t4 - t3, 3), "second_ctx_create_s": round(t5 - t4, 3)})
