"""SASS opcode histogram of every kernel in the built library (cuobjdump -sass), so that claims such as
"FMNMX3 in the FfDL row loop" or "REDUX in the AFS-L pass" can be checked without the .so.
Usage: python profiles/sass_histogram.py > profiles/r2_sass_opcodes.txt"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
so = os.path.join(ROOT, "vodascheduler_b200", "libvoda_accel.so")
out = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True).stdout
git = subprocess.run(["git", "rev-parse", "--short", "HEAD"], capture_output=True, text=True, cwd=ROOT).stdout.strip()
print(f"# cuobjdump -sass vodascheduler_b200/libvoda_accel.so (sm_100a), source tree at git {git} (+ working changes)")
name, hist = None, None


def flush():
    if name and hist:
        total = sum(hist.values())
        demangled = subprocess.run(["c++filt", name], capture_output=True, text=True).stdout.strip()[:150]
        print(f"\n== {demangled}\n   {total} instructions")
        line = "   "
        for op, c in sorted(hist.items(), key=lambda kv: -kv[1]):
            item = f"{op}:{c}  "
            if len(line) + len(item) > 118:
                print(line)
                line = "   "
            line += item
        print(line)


for ln in out.splitlines():
    m = re.search(r"Function : (\S+)", ln)
    if m:
        flush()
        name, hist = m.group(1), collections.Counter()
        continue
    m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_]*(?:\.[A-Z0-9_]+)*)", ln)
    if m and hist is not None:
        op = m.group(1).split(".")
        key = op[0] + ("." + op[1] if len(op) > 1 and op[0] in ("REDUX", "CREDUX", "LDG", "STG", "LDS", "STS", "SHFL", "LDGSTS", "FMNMX3", "ATOMS", "BAR", "VOTE", "MUFU") else "")
        hist[key] += 1
flush()
