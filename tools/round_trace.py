"""Print the slot timeline of the step kernel from a trace CSV (library built with -DATE_ROUND_TRACE,
tools/build_trace_lib.sh; written by ate_round_kernel_bench when ATE_ROUND_TRACE_OUT is set)."""
import csv
import sys

rows = list(csv.DictReader(open(sys.argv[1])))
cols = [c for c in rows[0] if c != "slot"]
lo, hi = (int(sys.argv[2]), int(sys.argv[3])) if len(sys.argv) > 3 else (40, 64)
print("slot | " + " ".join(f"{c:>9s}" for c in cols) + " | period")
prev = None
for r in rows[lo:hi]:
    v = [int(r[c]) for c in cols]
    pf = v[0]
    print(f"{int(r['slot']):4d} | " + " ".join(f"{(x - pf) if x >= 0 else -1:9d}" for x in v) + f" | {pf - prev if prev is not None else 0}")
    prev = pf
