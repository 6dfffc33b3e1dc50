"""gritpack: B200-native checkpoint-image pack/restore behind GRIT's TransferData boundary.

  grit_b200.copy     mirror of pkg/gritagent/copy (TransferData, CreateSentinelFile)
  grit_b200.agent    mirror of RunCheckpoint / RunRestore
  grit_b200.buffers  page kernels on device / host buffers (needs torch + a B200)
  grit_b200.csrc     CUDA kernels and the C ABI (libgritpack.so, include/gritpack.h)
"""
from ._lib import GritError, PAGE  # noqa: F401
