#!/usr/bin/env python
"""How the CPU restatement's time grows with n (same p, K, folds, lambdas), on ONE machine with a
fixed thread count -- the evidence behind bench.py's extrapolation of its bounded CPU sample to the
full C1 workload.  Writes one JSON line per n to stdout (and to --out).

  python tools/cpu_linearity.py --threads 4 --ns 10000,100000,1000000 --out profiles/r02_cpu_linearity.jsonl
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--threads", type=int, default=4)
    ap.add_argument("--ns", default="10000,100000")
    ap.add_argument("--p0", type=int, default=16)
    ap.add_argument("--B", type=int, default=20)
    ap.add_argument("--out", default=None)
    args = ap.parse_args()
    import bench
    from oracle import oracle as O
    thr = O.set_threads(args.threads)
    out = open(args.out, "a") if args.out else None
    for n in [int(t) for t in args.ns.split(",")]:
        d, fo, ips, iout = bench.make_inputs(n, args.p0)
        xs, _, _ = O.scale(d["x_raw"], cols=np.arange(d["n_continuous"]))
        xd = O.expand_deg2(xs, False)
        t0 = time.perf_counter()
        r = O.crossfit_aipw(xd, d["w"], d["y"], fo, ips, iout)
        t_fit = time.perf_counter() - t0
        rng = np.random.Generator(np.random.Philox(key=9))
        bidx = rng.integers(1, n + 1, size=(args.B, n)).astype(np.int32)
        t0 = time.perf_counter()
        O.bootstrap_idx(d["w"], d["y"], r["phat"], r["m0"], r["m1"], bidx)
        t_boot = time.perf_counter() - t0
        line = dict(n=n, p=args.p0 ** 2, threads=thr, fit_s=t_fit, fit_s_per_row=t_fit / n, boot_s=t_boot,
                    boot_s_per_draw=t_boot / (args.B * n), B=args.B, tau=r["tau"], se=r["se"],
                    host=os.uname().nodename, cpus=os.cpu_count())
        print(json.dumps(line), flush=True)
        if out:
            out.write(json.dumps(line) + "\n")
            out.flush()


if __name__ == "__main__":
    main()
