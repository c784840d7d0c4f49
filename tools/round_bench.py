"""The IRLS step kernel alone on synthetic rows: ms per launch, cycles per row and fit, and what each
role costs (skip masks).  python tools/round_bench.py [--n 800000] [--ni 256]"""
import argparse
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ate_replication_causalml_b200.engine import Engine  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--n", type=int, default=800_000)
ap.add_argument("--ni", type=int, default=256)
ap.add_argument("--iters", type=int, default=5)
ap.add_argument("--cases", default="1x1,5x1,1x10,5x10")
ap.add_argument("--lists", default="0,32,64,96")
ap.add_argument("--skips", default="0,1,2,4,8,3,15")
ap.add_argument("--list-mode", action="store_true")
args = ap.parse_args()
eng = Engine(0)
lib, h = eng._lib, eng._h
print("nprob x fits | list | skip | ms/launch | cycles per row-fit (148 SMs, 1.965 GHz) | TFLOP/s alg | GB/s alg | CTAs")
for case in args.cases.split(","):
    P, F = (int(t) for t in case.split("x"))
    for L in (int(t) for t in args.lists.split(",")):
        for sk in (int(t) for t in args.skips.split(",")):
            ms, fl, by, ct = C.c_double(), C.c_double(), C.c_double(), C.c_int()
            rc = lib.ate_round_kernel_bench(h, args.n, args.ni, P, F, L, 0 if args.list_mode else 1, args.iters, sk,
                                            C.byref(ms), C.byref(fl), C.byref(by), C.byref(ct))
            if rc != 0:
                print(case, L, sk, "FAILED", lib.ate_last_error(h))
                continue
            rf = args.n * P * F
            print(f"{P:2d} x {F:2d} | {L:3d} | {sk:2d} | {ms.value:8.3f} | {ms.value * 1e-3 * 148 * 1.965e9 / rf:7.1f} | "
                  f"{fl.value / ms.value / 1e9:6.2f} | {by.value / ms.value / 1e6:7.0f} | {ct.value}", flush=True)
eng.close()
