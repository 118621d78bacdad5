#!/usr/bin/env python3
"""Compile one benchmark kernel (no GPU needed) and print registers / smem / SASS statistics.
usage: tools/kinfo.py kepler|pendulum parity|fast [reg_budget] [block_threads] [flags]"""
import os, re, subprocess, sys, tempfile, collections
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import taylorintegration_b200 as tb
from taylorintegration_b200 import _lib as L, problems as P
name, mode = sys.argv[1], sys.argv[2]
rb = int(sys.argv[3]) if len(sys.argv) > 3 else 0
bt = int(sys.argv[4]) if len(sys.argv) > 4 else 0
fl = int(sys.argv[5]) if len(sys.argv) > 5 else 0
tape, order = {"kepler": (P.kepler_tape(), 28), "pendulum": (P.pendulum_tape(), 25)}[name]
k = tb.Kernel(tape, order, L.KIND_PLAIN, L.MODE_PARITY if mode == "parity" else L.MODE_FAST, reg_budget=rb, block_threads=bt, flags=fl | 1)
print(k.info())
with tempfile.NamedTemporaryFile(suffix=".cubin", delete=False) as f:
    f.write(k.cubin()); path = f.name
sass = subprocess.run(["cuobjdump", "-sass", path], capture_output=True, text=True).stdout
os.unlink(path)
fn = None; cnt = collections.defaultdict(collections.Counter)
for line in sass.splitlines():
    m = re.match(r"\s*Function : (\S+)", line)
    if m: fn = m.group(1); continue
    m = re.match(r"\s*/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
    if m and fn: cnt[fn][m.group(1).split(".")[0]] += 1
for f, c in cnt.items():
    tot = sum(c.values())
    print(f"{f}: {tot} instr ({tot*16/1024:.0f} KB)", {k: v for k, v in c.most_common(12)})
if os.environ.get("DUMP"): open(os.environ["DUMP"], "w").write(k.source)
