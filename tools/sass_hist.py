"""Opcode histogram of every kernel in grit_b200/csrc/kernels.o (cuobjdump -sass), written to profiles/: the
data-movement path of each kernel can then be audited without rebuilding (LDG / LDGSTS / STG / LDS, and for the
single-pass kernel the tensor-memory instructions STTM / LDTM, SASS for tcgen05.st / tcgen05.ld)."""
import collections
import os
import re
import subprocess
import sys

root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
obj = os.path.join(root, "grit_b200", "csrc", "kernels.o")
out = subprocess.run(["cuobjdump", "-sass", obj], capture_output=True, text=True, check=True).stdout
tag = sys.argv[1] if len(sys.argv) > 1 else "r2"
name, hist = None, {}
for ln in out.splitlines():
    m = re.search(r"Function : (\S+)", ln)
    if m:
        k = re.search(r"\d+(k_[a-z_]+)", m.group(1))
        name = k.group(1) if k else m.group(1)
        hist[name] = collections.Counter()
        continue
    m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_.]*)", ln)
    if m and name:
        hist[name][m.group(1)] += 1
with open(os.path.join(root, "profiles", f"{tag}_sass_opcodes.txt"), "w") as f:
    f.write("# cuobjdump -sass grit_b200/csrc/kernels.o (sm_100a), opcode counts per kernel, memory and tensor-memory ops first\n")
    for k, h in sorted(hist.items()):
        mem = {o: c for o, c in h.items() if re.match(r"(LD|ST|ATOM|RED|LDGSTS|LDTM|STTM|UTC|UBLKCP|UTMA|BAR|MEMBAR|CCTL|SYNCS)", o)}
        f.write(f"\n{k}: {sum(h.values())} instructions\n")
        f.write("  memory: " + ", ".join(f"{o} {c}" for o, c in sorted(mem.items(), key=lambda x: -x[1])) + "\n")
        f.write("  top:    " + ", ".join(f"{o} {c}" for o, c in h.most_common(12)) + "\n")
print("written", f"profiles/{tag}_sass_opcodes.txt")
