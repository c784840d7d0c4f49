"""Summarise a slot trace of the fused IRLS step (irls_step.cu built with -DATE_STEP_TRACE).

    ATE_EXTRA_NVCC_FLAGS=-DATE_STEP_TRACE python -c "from ate_replication_causalml_b200 import build; build.build(force=True)"
    ATE_STEP_TRACE_OUT=gpurun_out/trace.csv python bench.py --steps 1 --warmup 3      # on the GPU box
    python tools/step_trace.py gpurun_out/trace.csv

CTA (0,0) of the last traced launch stamps clock64() per slot: producer sees the slot free, eval
start/done, helper start/done, MMA warps 0-2 start/done.  Prints the steady-state slot period and the
length of each phase in cycles.  The stamps perturb the traced CTA a little (a global store sits in
front of the release barrier), so read phases, not gaps of a few hundred cycles.
"""
import csv
import statistics as st
import sys

rows = [r for r in csv.reader(open(sys.argv[1]))]
print(" ".join(rows[0]))
data = [[int(x) for x in r] for r in rows[2:] if r and r[0].isdigit()]
live = [r for r in data if r[1] > 0 and r[3] > 0]
lo, hi = len(live) // 4, len(live) - 4
R = live[lo:hi]
names = ["slot", "prod_free", "eval_start", "eval_done", "help_start", "help_done", "m0_start", "m0_done", "m1_start",
         "m1_done", "m2_start", "m2_done"]


def mean_delta(a, b):
    d = [r[b] - r[a] for r in R if r[a] > 0 and r[b] > 0]
    return round(st.mean(d)) if d else None


print("slots used", len(R), " period", round(st.mean(R[i + 1][1] - R[i][1] for i in range(len(R) - 1))), "cycles/slot")
for a, b, label in [(1, 2, "fill (slot free -> rows landed)"), (2, 3, "eval"), (3, 4, "eval done -> helper start"),
                    (4, 5, "helper"), (3, 6, "eval done -> MMA0 start"), (6, 7, "MMA warp 0"), (8, 9, "MMA warp 1"),
                    (10, 11, "MMA warp 2")]:
    print(f"  {label:34s} {mean_delta(a, b)}")
