#!/usr/bin/env python
"""bench.py -- cross-fitted AIPW ATE + bootstrap-SE fits/sec (BASELINE.json's metric).

A "step" is one complete fit of SURVEY.md section 8's composite on the C1 workload:
n = 1,000,000 rows, p0 = 16 base covariates -> the 256 ordered degree-2 products
(ate_functions.R:290-296), K = 5 cross-fitting folds, cv.glmnet with 10 inner folds and a
100-lambda path for both nuisances (L1-logistic propensity, lasso outcome), AIPW score,
sandwich SE and B = 2000 Philox bootstrap replicates.

  value : fits/s with the inputs already resident in HBM (device pointers into the C ABI)
  e2e   : fits/s through the same C-ABI call with HOST (pinned) buffers, H2D/D2H inside
  roofline : the dominant kernel (fp64 DMMA weighted Gram), algorithmic flop / CUDA-event time
             of every launch in the timed region, against cuBLAS DGEMM measured in this run
  cpu_baseline : the CPU oracle (a port: R is not installed) on a bounded sample

N > 1 (torchrun): one process per GPU, data replicated, the 50 CV-fold paths of each nuisance
and the bootstrap replicates are sharded over ranks; held-out CV deviances cross ranks through
one small all-reduce, replicate estimates through one all-gather.  One fit is timed => "strong".
"""
from __future__ import annotations

import argparse
import ctypes
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


def workload(args):
    return dict(workload="C1: cross-fit AIPW, L1-logistic propensity + lasso outcome, degree-2 "
                         "products, 5 folds, 10 inner cv folds, 100-lambda path, Philox bootstrap",
                n=args.n, p0=args.p0, p=args.p0 * args.p0, K=5, nfolds_inner=10, nlambda=args.nlambda,
                B=args.B, l2="inputs_exceed_l2 (design matrix %.2f GB)" % (args.n * (args.p0 ** 2 + 32) * 8 / 1e9))


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


# ------------------------------------------------------------------------------------------
def cpu_sample(args, n_sample, threads=None):
    """The CPU oracle on a bounded sample of the same workload.  Returns (seconds, fits/s
    scaled linearly in n to the full workload, threads used, description)."""
    from oracle import oracle as O
    if threads:
        os.environ["OMP_NUM_THREADS"] = str(threads)
    ncores = threads or os.cpu_count() or 1
    d, fo, ips, iout = make_inputs(n_sample, args.p0)
    xs, _, _ = O.scale(d["x_raw"], cols=np.arange(d["n_continuous"]))
    xd = O.expand_deg2(xs, False)
    Bs = max(20, int(args.B * n_sample / args.n))
    t = time.perf_counter()
    r = O.crossfit_aipw(xd, d["w"], d["y"], fo, ips, iout, nlambda=args.nlambda)
    rng = np.random.Generator(np.random.Philox(key=9))
    bidx = rng.integers(1, n_sample + 1, size=(Bs, n_sample)).astype(np.int32)
    O.bootstrap_idx(d["w"], d["y"], r["phat"], r["m0"], r["m1"], bidx)
    dt = time.perf_counter() - t
    # glmnet's per-sweep cost and the bootstrap are linear in n: scale the sample to n
    fits_per_s = (1.0 / dt) * (n_sample / args.n)
    desc = (f"one full fit at n={n_sample} (same p={args.p0 ** 2}, K=5, 10 inner folds, {args.nlambda} lambdas, "
            f"B={Bs}) took {dt:.1f} s; scaled linearly in n to n={args.n}")
    return dt, fits_per_s, min(ncores, 10), desc


def run_reference(args, rank):
    if rank != 0:
        return
    t0 = time.perf_counter()
    vals = []
    steps = max(1, args.steps)
    desc = ""
    cores = 1
    warm = 1 if args.warmup >= 1 else 0  # one untimed sample pages the oracle library and its threads in
    for _ in range(warm):
        cpu_sample(args, args.cpu_sample_n)
    for _ in range(steps):
        dt, fps, cores, desc = cpu_sample(args, args.cpu_sample_n)
        vals.append((dt, fps))
    fps = float(np.mean([v[1] for v in vals]))
    ms = 1e3 / fps
    line = dict(metric=METRIC, value=fps, unit=UNIT, n_gpus=args.gpus, steps=steps, warmup=warm,
                ms_per_step=ms, higher_is_better=True, scaling="strong", vs_baseline=None, dtype="f64",
                data="synthetic", impl="reference", config=workload(args),
                cpu_baseline=dict(value=fps, unit=UNIT, cores=cores, kind="port",
                                  sample=desc + " (R and glmnet are not installed on this box: the timed code is "
                                                "oracle/ate_oracle.c, a C restatement of glmnet's coordinate descent)"),
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
    ap.add_argument("--n", type=int, default=1_000_000)
    ap.add_argument("--p0", type=int, default=16)
    ap.add_argument("--B", type=int, default=2000)
    ap.add_argument("--nlambda", type=int, default=100)
    ap.add_argument("--cpu-sample-n", type=int, default=10_000)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    args = ap.parse_args()

    import faulthandler
    faulthandler.dump_traceback_later(int(os.environ.get("ATE_BENCH_WATCHDOG", "1500")), exit=True)  # a stuck run leaves a traceback
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))

    if args.impl == "reference":
        run_reference(args, rank)
        return

    import torch
    import torch.distributed as dist
    from ate_replication_causalml_b200.engine import Engine
    from ate_replication_causalml_b200 import parallel

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

    d, fo, ips, iout = make_inputs(args.n, args.p0)
    n, p0 = args.n, args.p0
    # pinned host buffers (e2e leg) and HBM-resident copies (value leg); column-major base matrix
    xb_host = torch.from_numpy(np.asfortranarray(d["x_raw"]).T.copy()).contiguous().pin_memory()  # (p0, n) == col-major n x p0
    w_host = torch.from_numpy(d["w"]).pin_memory()
    y_host = torch.from_numpy(d["y"]).pin_memory()
    xb_dev, w_dev, y_dev = xb_host.to(dev), w_host.to(dev), y_host.to(dev)

    eng = Engine(local_rank)
    stream = torch.cuda.ExternalStream(eng.stream_ptr, device=dev)
    comm = parallel.TorchComm(dist, dev) if world > 1 else None

    def one_fit(on_device: bool):
        if on_device:
            return parallel.crossfit_sharded(eng, xb_dev.data_ptr(), w_dev.data_ptr(), y_dev.data_ptr(), n, p0,
                                             fo, ips, iout, comm=comm, on_device=True, n_continuous=d["n_continuous"],
                                             expand=1, B=args.B, nlambda=args.nlambda)
        return parallel.crossfit_sharded(eng, xb_host.numpy().T, w_host.numpy(), y_host.numpy(), n, p0, fo, ips,
                                         iout, comm=comm, on_device=False, n_continuous=d["n_continuous"],
                                         expand=1, B=args.B, nlambda=args.nlambda)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(nsteps, on_device):
        barrier()
        ev0 = torch.cuda.Event(enable_timing=True)
        ev1 = torch.cuda.Event(enable_timing=True)
        ev0.record(stream)
        last = None
        for _ in range(nsteps):
            last = one_fit(on_device)
        ev1.record(stream)
        barrier()
        ms = ev0.elapsed_time(ev1)
        if world > 1:
            t = torch.tensor([ms], dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms, last

    for _ in range(args.warmup):
        one_fit(True)
    sampler = ClockSampler(local_rank)
    launches0 = eng.launches
    eng.profile_enable(True)
    sampler.start()
    ms_total, last = timed(args.steps, True)
    clocks = sampler.stop()
    prof = eng.profile_read()
    eng.profile_enable(False)
    launches = eng.launches - launches0
    ms_step = ms_total / args.steps
    value = 1e3 / ms_step

    e2e = None
    if not args.no_e2e:
        one_fit(False)  # warm the host path once
        ms_e2e, last_h = timed(args.steps, False)
        h2d = (n * p0 + 2 * n) * 8
        d2h = 3 * n * 8 + args.B * 8 + 256
        e2e = dict(value=1e3 / (ms_e2e / args.steps), unit=UNIT, h2d_bytes_per_step=h2d, d2h_bytes_per_step=d2h,
                   ms_per_step=ms_e2e / args.steps)

    # ---- roofline of the dominant kernel ----
    # The dominant kernel is the fused IRLS step (irls_step_kernel): it streams the design rows
    # once per step (HBM-bound once the Gram is restricted to glmnet's strong set) and runs the
    # strong-set Gram on the fp64 DMMA pipe from the same staged rows.
    roof = None
    if rank == 0:
        peak_tensor = measure_dgemm_peak(torch, dev)
        gl = max(1, prof["gram_launches"])
        avg_ms = prof["gram_ms"] / gl
        hbm_peak, peak_src = 6650.0, "fallback 6.65 TB/s (B200_PROFILING.md): of fallback"
        try:
            mp = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
            hbm_peak, peak_src = float(mp["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs (burst copy): of measured"
        except Exception:
            pass
        achieved = prof["gram_bytes"] / gl / (avg_ms * 1e-3) / 1e9 if avg_ms > 0 else 0.0
        # DRAM bytes of one launch from the committed ncu --set full capture (a smaller launch than
        # this run's: the ratio to its algorithmic bytes is what carries over)
        traffic, traffic_capture = None, None
        try:
            traffic_capture = json.load(open(os.path.join(ROOT, "profiles", "r01b_irls_step_traffic.json")))
            traffic = traffic_capture["dram_bytes_per_launch"]
        except Exception:
            pass
        tf = prof["gram_flops"] / gl / (avg_ms * 1e-3) / 1e12 if avg_ms > 0 else 0.0
        roof = dict(bound="hbm", kernel="irls_step_kernel (fused IRLS step: eta/q/v/r, X^T r, X^T v and the "
                                        "strong-set Gram X_S^T V X_S on DMMA from one pass over the rows)",
                    achieved=achieved, peak=hbm_peak, unit="GB/s", frac=achieved / hbm_peak, traffic=traffic,
                    traffic_capture=traffic_capture,
                    launches=prof["gram_launches"], avg_launch_ms=avg_ms, share_of_step=prof["gram_ms"] / ms_total,
                    algorithmic="nrows*(p+2)*8 bytes per launch (each staged row read once, v written once)",
                    peak_source=peak_src,
                    tensor=dict(achieved_tflops=tf, peak_tflops=peak_tensor, frac=tf / peak_tensor if peak_tensor else None,
                                note="fp64 DMMA work of the same launches: nrows*|L|*(|L|+1) flop over the column list L; "
                                     "peak = cuBLAS DGEMM 8192^3 measured in this run (MEASURED_PEAKS.json has no fp64 "
                                     "row). The full-width Gram (gram_dmma_kernel, used by the gaussian fits) reaches "
                                     "0.81 of that peak: profiles/"))

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        dt, fps, cores, desc = cpu_sample(args, args.cpu_sample_n)
        cpu = dict(value=fps, unit=UNIT, cores=cores, kind="port",
                   sample=desc + " (R/glmnet are not installed here: oracle/ate_oracle.c restates them)")

    if rank == 0:
        line = dict(metric=METRIC, value=value, unit=UNIT, n_gpus=world, steps=args.steps, warmup=args.warmup,
                    ms_per_step=ms_step, higher_is_better=True, scaling="strong", vs_baseline=None, dtype="f64",
                    data="synthetic", config=dict(workload(args), parallelism=("1 GPU" if world == 1 else
                                                  f"cv-fold paths + bootstrap replicates sharded over {world} GPUs")),
                    clocks=clocks, e2e=e2e, gpu_launches=int(launches), roofline=roof, cpu_baseline=cpu,
                    result=dict(tau=last["tau"], se_sandwich=last["se_sandwich"], se_bootstrap=last["se_bootstrap"],
                                true_ate=d["true_ate"], rounds_irls=last.get("rounds_irls"),
                                irls_solves=last.get("irls_solves"), strong_mean=last.get("strong_mean"),
                                strong_frac2=last.get("strong_frac2"),
                                stage_ms={k: last.get(k) for k in ("ms_design", "ms_outcome", "ms_propensity",
                                                                    "ms_score", "ms_bootstrap")}))
        sys.stdout.flush()
        os.write(json_fd, (json.dumps(line) + "\n").encode())
    eng.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
