"""Count the SASS instructions that prove which pipes the library's kernels use (DMMA = fp64 tensor,
UBLKCP = 1-D TMA bulk copy, SYNCS = mbarrier) per kernel; `cuobjdump` runs without a GPU."""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
so = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "ate_replication_causalml_b200", "libate_b200.so")
out = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True, check=True).stdout
fn, counts, total = None, collections.OrderedDict(), collections.Counter()
pat = {"DMMA": re.compile(r"\bDMMA\."), "UBLKCP": re.compile(r"\bUBLKCP"), "SYNCS": re.compile(r"\bSYNCS\."),
       "BRX": re.compile(r"\bBRX\b"), "DFMA": re.compile(r"\bDFMA\b"), "LDS": re.compile(r"\bLDS\."), "inst": re.compile(r"/\*[0-9a-f]{4,}\*/")}
for line in out.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        fn = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
        fn = re.sub(r"\(anonymous namespace\)::|<unnamed>::", "", fn)
        counts[fn] = collections.Counter()
        continue
    if fn is None:
        continue
    for k, rx in pat.items():
        if rx.search(line):
            counts[fn][k] += 1
            total[k] += 1
print(f"# {os.path.relpath(so, ROOT)}: SASS instruction counts per kernel (sm_100a)")
print(f"{'kernel':100s} {'inst':>7s} {'DMMA':>5s} {'UBLKCP':>6s} {'SYNCS':>5s} {'DFMA':>5s} {'LDS':>5s} {'BRX':>4s}")
for fn, c in counts.items():
    print(f"{fn[:100]:100s} {c['inst']:7d} {c['DMMA']:5d} {c['UBLKCP']:6d} {c['SYNCS']:5d} {c['DFMA']:5d} {c['LDS']:5d} {c['BRX']:4d}")
print(f"{'TOTAL':100s} {total['inst']:7d} {total['DMMA']:5d} {total['UBLKCP']:6d} {total['SYNCS']:5d} {total['DFMA']:5d} {total['LDS']:5d} {total['BRX']:4d}")
