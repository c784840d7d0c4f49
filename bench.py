#!/usr/bin/env python
"""bench.py -- cross-fitted AIPW ATE + bootstrap-SE fits/sec (BASELINE.json's metric).

Default workload C1 (BASELINE.json configs[1]); a "step" is one complete fit of SURVEY.md section
8's composite: n = 1,000,000 rows, p0 = 16 base covariates -> the 256 ordered degree-2 products
(ate_functions.R:290-296), K = 5 cross-fitting folds, cv.glmnet with 10 inner folds and a
100-lambda path for both nuisances (L1-logistic propensity, lasso outcome), AIPW score, sandwich
SE and B = 2000 Philox bootstrap replicates.

  value        fits/s, inputs already resident in HBM (device pointers into the C ABI), timed with
               CUDA events on the library's stream, profiling hooks OFF
  e2e          fits/s through the same C-ABI call with HOST (pinned) buffers: the H2D copies of
               the inputs and fold ids and the D2H of the nuisances / replicates are inside
  roofline     the dominant kernel, from a separate short profiled pass (CUDA events around
               every launch of it on the library's stream)
  cpu_baseline the CPU oracle (a port: R is not installed) on a bounded sample, all host threads,
               plus the same sample on ONE thread (what R itself would use)

Other workloads (--workload C2|C3|C4) are the remaining BASELINE.json configs; their lines carry
their own metric names and are committed under profiles/.

N > 1 (torchrun): one process per GPU.  The rows are sharded: rank r keeps rows
[r n / N, (r + 1) n / N) of the design and of every fit's private row copy; each IRLS round's packed
partial sums (strong-set Grams, X^T r, X^T v, deviance) meet in ONE ncclAllReduce on the library's
stream; the bootstrap replicates are sharded by replicate id.  One fit is timed => "strong".
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

METRIC = "cross-fitted AIPW ATE+bootstrap-SE fits/sec at n=1M,p=256"
UNIT = "fits/s"

WORKLOADS = {
    "C1": dict(n=1_000_000, p0=16, expand=1, B=2000,
               text="C1: cross-fit AIPW, L1-logistic propensity + lasso outcome, degree-2 products, 5 folds, "
                    "10 inner cv folds, 100-lambda path, Philox bootstrap"),
    "C2": dict(n=1_000_000, p0=16, expand=1, B=10_000,
               text="C2: bootstrap ATE standard error, B = 10,000 replicates over the n = 1M AIPW scores"),
    "C3": dict(n=4_000_000, p0=0, p=1024, B=0,
               text="C3: 10-fold cv.glmnet lambda selection (gaussian lasso), n = 4M, p = 1024"),
    "C4": dict(n=16_000_000, p0=22, expand=2, B=2000,
               text="C4: cross-fit AIPW, n = 16M, p0 = 22 -> 22 mains + 484 ordered products = 506 columns "
                    "(BASELINE's 'p = 512'), 5 folds, 10 inner cv folds, 100-lambda path, B = 2000"),
}


def design_width(p0, expand):
    return p0 if expand == 0 else (p0 * p0 if expand == 1 else p0 + p0 * p0)


def workload(args):
    p = design_width(args.p0, args.expand)
    return dict(workload=WORKLOADS[args.workload]["text"], n=args.n, p0=args.p0, p=p, K=5, nfolds_inner=10,
                nlambda=args.nlambda, B=args.B,
                l2="inputs_exceed_l2 (design matrix %.2f GB)" % (args.n * (p + 32) * 8 / 1e9))


def make_inputs(n, p0):
    from ate_replication_causalml_b200 import synth
    d = synth.make_dataset(n, p0)
    K = 5
    fo = synth.fold_ids(n, K, 11)
    ips = np.zeros((K, n), np.int32)
    iout = np.zeros((K, n), np.int32)
    for k in range(K):
        nt = int((fo != k + 1).sum())
        ips[k, :nt] = synth.fold_ids(nt, 10, 100 + k)
        iout[k, :nt] = synth.fold_ids(nt, 10, 200 + k)
    return d, fo, ips, iout


# ------------------------------------------------------------------------------------------
class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.gpu = gpu_index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "200", "-i", str(self.gpu)], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except Exception:
            self.proc = None

    def _read(self):
        for ln in self.proc.stdout:
            self.lines.append(ln.strip())

    def stop(self):
        if not self.proc:
            return dict(sm_mhz=None, sm_max_mhz=None, reasons=["nvidia-smi unavailable"])
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [t.strip() for t in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1]))
                mx.append(float(f[2]))
            except ValueError:
                continue
            for nm, val in zip(names, f[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(nm)
        return dict(sm_mhz=float(np.median(sm)) if sm else None, sm_max_mhz=max(mx) if mx else None,
                    reasons=sorted(reasons), samples=len(sm))


def measure_dgemm_peak(torch, dev):
    """fp64 tensor ceiling: cuBLAS DGEMM 8192^3, best of 5 (MEASURED_PEAKS.json has no fp64 row)."""
    N = 8192
    a = torch.randn(N, N, dtype=torch.float64, device=dev)
    b = torch.randn(N, N, dtype=torch.float64, device=dev)
    for _ in range(2):
        a @ b
    torch.cuda.synchronize()
    best = 1e30
    for _ in range(5):
        s = torch.cuda.Event(enable_timing=True)
        e = torch.cuda.Event(enable_timing=True)
        s.record()
        a @ b
        e.record()
        torch.cuda.synchronize()
        best = min(best, s.elapsed_time(e))
    del a, b
    torch.cuda.empty_cache()
    return 2.0 * N ** 3 / best / 1e9  # TFLOP/s


def hbm_peak():
    try:
        mp = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        return float(mp["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs (burst copy): of measured"
    except Exception:
        return 6650.0, "fallback 6.65 TB/s (B200_PROFILING.md): of fallback"


# ------------------------------------------------------------------------------------------
def host_threads():
    """Host threads this process may use (torchrun exports OMP_NUM_THREADS=1: not a property of the box)."""
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def per_row_scaling(n_sample, n_full):
    """seconds-per-row of the CPU port's nuisance fits at n_full relative to n_sample, from the one-machine
    measurements in profiles/r02_cpu_linearity.jsonl (tools/cpu_linearity.py: n = 10k .. 1M at a fixed thread
    count).  glmnet needs fewer passes per row as n grows, so a small sample scaled linearly would OVERSTATE
    the CPU time of the full workload (by 2.0x from n = 10k); this factor takes that out.  1.0 when the
    file is missing or the sizes fall outside what was measured."""
    try:
        pts = sorted((json.loads(l)["n"], json.loads(l)["fit_s_per_row"])
                     for l in open(os.path.join(ROOT, "profiles", "r02_cpu_linearity.jsonl")) if l.strip())
    except Exception:
        return 1.0, "no linearity file: scaled linearly in n"
    if len(pts) < 2:
        return 1.0, "no linearity file: scaled linearly in n"
    xs = np.log([p[0] for p in pts])
    ys = np.log([p[1] for p in pts])
    at = lambda n: float(np.exp(np.interp(np.log(n), xs, ys)))
    f = at(n_full) / at(n_sample)
    return f, (f"seconds per row at n={n_full} are {f:.3f} x those at n={n_sample} on one machine "
               "(profiles/r02_cpu_linearity.jsonl)")


def cpu_sample(args, n_sample, threads):
    """One fit of the CPU oracle on a bounded sample of the same workload (same p, K, inner folds and
    lambdas; n_sample rows; Bs replicates), on exactly `threads` OpenMP threads.  The two legs scale
    differently and are extrapolated separately: the nuisance fits are linear in n (glmnet's sweeps
    are O(n) each; checked on one machine in profiles/r02_cpu_linearity.jsonl), the bootstrap is
    linear in B * n."""
    from oracle import oracle as O
    used = O.set_threads(threads)
    d, fo, ips, iout = make_inputs(n_sample, args.p0)
    xs, _, _ = O.scale(d["x_raw"], cols=np.arange(d["n_continuous"]))
    xd = O.expand_deg2(xs, include_main=(args.expand == 2)) if args.expand else xs
    Bs = max(20, int(round(args.B * n_sample / args.n))) if args.B > 0 else 0
    t = time.perf_counter()
    r = O.crossfit_aipw(xd, d["w"], d["y"], fo, ips, iout, nlambda=args.nlambda)
    t_fit = time.perf_counter() - t
    t_boot = 0.0
    if Bs > 0:
        rng = np.random.Generator(np.random.Philox(key=9))
        bidx = rng.integers(1, n_sample + 1, size=(Bs, n_sample)).astype(np.int32)
        t = time.perf_counter()
        O.bootstrap_idx(d["w"], d["y"], r["phat"], r["m0"], r["m1"], bidx)
        t_boot = time.perf_counter() - t
    corr, corr_text = per_row_scaling(n_sample, args.n)
    full_s = t_fit * (args.n / n_sample) * corr + (t_boot * (args.B * args.n) / (Bs * n_sample) if Bs > 0 else 0.0)
    return dict(n=n_sample, B=Bs, threads=used, fit_s=t_fit, boot_s=t_boot, sample_s=t_fit + t_boot,
                full_workload_s=full_s, fits_per_s=1.0 / full_s, per_row_scaling=corr, per_row_scaling_text=corr_text)


def sample_text(args, s):
    return (f"one full fit of the CPU port at n={s['n']} (same p={design_width(args.p0, args.expand)}, K=5, 10 inner "
            f"folds, {args.nlambda} lambdas) took {s['fit_s']:.2f} s and B={s['B']} index-mode replicates {s['boot_s']:.2f} s "
            f"on {s['threads']} threads; extrapolated to n={args.n}, B={args.B}: nuisance fits x{args.n / s['n']:.0f} x "
            f"{s['per_row_scaling']:.3f} ({s['per_row_scaling_text']}), bootstrap x{(args.B * args.n) / max(1, s['B'] * s['n']):.0f} "
            f"(linear in B*n) = {s['full_workload_s']:.0f} s per fit.  R and glmnet are not installed on this box: the timed code is oracle/ate_oracle.c, a C "
            "restatement of glmnet's coordinate descent, cv.glmnet and the reference's bootstrap loop")


def run_reference(args, rank):
    """--impl reference: the reference's CPU path (the oracle port) on rank 0, all host threads."""
    if rank != 0:
        return
    t0 = time.perf_counter()
    threads = args.cpu_threads or host_threads()
    steps = max(1, args.steps)
    warm = 1 if args.warmup >= 1 else 0  # one untimed sample pages the oracle library and its threads in
    for _ in range(warm):
        cpu_sample(args, args.cpu_sample_n, threads)
    runs = [cpu_sample(args, args.cpu_sample_n, threads) for _ in range(steps)]
    s = dict(runs[0])
    for k in ("fit_s", "boot_s", "sample_s", "full_workload_s"):
        s[k] = float(np.mean([r[k] for r in runs]))
    fps = 1.0 / s["full_workload_s"]
    one = None
    if not args.no_single_thread:
        o = cpu_sample(args, args.cpu_sample_n, 1)
        one = dict(value=o["fits_per_s"], unit=UNIT, cores=1, kind="port", sample=sample_text(args, o))
    lin = None
    if args.linearity:
        b = cpu_sample(args, 2 * args.cpu_sample_n, threads)
        lin = dict(n_a=s["n"], fit_s_a=s["fit_s"], n_b=b["n"], fit_s_b=b["fit_s"],
                   per_row_ratio=(b["fit_s"] / b["n"]) / (s["fit_s"] / s["n"]),
                   note="seconds per row of the nuisance fits at 2x the sample (1.0 = linear in n)")
    cfg = workload(args)
    cfg["sample"] = dict(n=s["n"], B=s["B"], threads=s["threads"],
                         note="what each timed step actually ran; `value` is that time extrapolated to the n and B "
                              "above (cpu_baseline.sample states the arithmetic), `ms_per_step` is the sample's own time")
    line = dict(metric=METRIC, value=fps, unit=UNIT, n_gpus=args.gpus, steps=steps, warmup=warm,
                ms_per_step=1e3 * s["sample_s"], ms_per_step_extrapolated=1e3 * s["full_workload_s"],
                higher_is_better=True, scaling="strong", vs_baseline=None, dtype="f64",
                data="synthetic", impl="reference", config=cfg,
                cpu_baseline=dict(value=fps, unit=UNIT, cores=s["threads"], kind="port", sample=sample_text(args, s)),
                cpu_baseline_1thread=one, linearity=lin,
                e2e=dict(value=fps, unit=UNIT, h2d_bytes_per_step=0, d2h_bytes_per_step=0),
                gpu_launches=0, wall_s=time.perf_counter() - t0)
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=2)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="C1", choices=sorted(WORKLOADS))
    ap.add_argument("--n", type=int, default=None)
    ap.add_argument("--p0", type=int, default=None)
    ap.add_argument("--B", type=int, default=None)
    ap.add_argument("--nlambda", type=int, default=100)
    ap.add_argument("--cpu-sample-n", type=int, default=10_000)
    ap.add_argument("--cpu-threads", type=int, default=0, help="threads of the CPU legs (0 = every host thread)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-single-thread", action="store_true")
    ap.add_argument("--linearity", action="store_true", help="reference arm: time a second sample at 2x the rows")
    ap.add_argument("--no-e2e", action="store_true")
    args = ap.parse_args()
    wl = WORKLOADS[args.workload]
    if args.n is None:
        args.n = wl["n"]
    if args.p0 is None:
        args.p0 = wl["p0"]
    if args.B is None:
        args.B = wl["B"]
    args.expand = wl.get("expand", 1)

    import faulthandler
    faulthandler.dump_traceback_later(int(os.environ.get("ATE_BENCH_WATCHDOG", "1500")), exit=True)  # a stuck run leaves a traceback
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))

    if args.impl == "reference":
        if args.workload != "C1":
            raise SystemExit("bench.py: the reference arm times the headline workload (C1) only")
        run_reference(args, rank)
        return

    import torch
    import torch.distributed as dist

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (the engine has no CPU fallback)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    # stdout carries exactly ONE JSON line (rank 0).  NCCL prints its version banner on fd 1 when the
    # box exports NCCL_DEBUG=VERSION, so everything up to the final print goes to stderr instead.
    sys.stdout.flush()
    json_fd = os.dup(1)
    os.dup2(2, 1)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    import bench_workloads as W
    ctx = W.BenchCtx(args=args, rank=rank, world=world, local_rank=local_rank, dev=dev, torch=torch, dist=dist)
    if args.workload in ("C1", "C4"):
        line = W.run_crossfit(ctx)
    elif args.workload == "C2":
        line = W.run_bootstrap(ctx)
    else:
        line = W.run_cv_gaussian(ctx)
    if rank == 0:
        sys.stdout.flush()
        os.write(json_fd, (json.dumps(line) + "\n").encode())
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
