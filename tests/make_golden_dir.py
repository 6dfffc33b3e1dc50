"""Writes tests/golden/packed_base and tests/golden/packed_child with the CUDA path.  Needs a B200:

    gpurun -- 'python tests/make_golden_dir.py gpurun_out/golden && ...'   then copy gpurun_out/golden/* to tests/golden/

The child is packed against the base through a RELATIVE parent path (the current directory is the output directory),
so the manifest does not hold a path of the box it was made on; restore finds the parent as the sibling directory."""
import os
import shutil
import sys
import tempfile

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))

from golden_dir import build_golden_tree  # noqa: E402

from grit_b200 import copy  # noqa: E402


def main(out):
    os.makedirs(out, exist_ok=True)
    os.chdir(out)
    for d in ("packed_base", "packed_child"):
        shutil.rmtree(d, ignore_errors=True)
    with tempfile.TemporaryDirectory() as tmp:
        build_golden_tree(os.path.join(tmp, "base"))
        build_golden_tree(os.path.join(tmp, "child"), dirty=True)
        s0 = copy.TransferData(os.path.join(tmp, "base"), "packed_base", copy.UPLOAD)
        s1 = copy.TransferData(os.path.join(tmp, "child"), "packed_child", copy.UPLOAD, parent_dir="packed_base")
    print({k: v for k, v in s0.items() if k.startswith("n_")})
    print({k: v for k, v in s1.items() if k.startswith("n_")})


if __name__ == "__main__":
    main(sys.argv[1] if len(sys.argv) > 1 else os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden"))
