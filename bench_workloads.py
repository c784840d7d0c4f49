"""GPU arms of bench.py, one function per BASELINE.json config (see bench.py for the contract)."""
from __future__ import annotations

import dataclasses
import json
import os
import time

import numpy as np

import bench as B

ROOT = os.path.dirname(os.path.abspath(__file__))


@dataclasses.dataclass
class BenchCtx:
    args: object
    rank: int
    world: int
    local_rank: int
    dev: object
    torch: object
    dist: object


def _barrier(c):
    if c.world > 1:
        c.dist.barrier()
    c.torch.cuda.synchronize()


def _timed(c, stream, nsteps, fn):
    """nsteps calls of fn between two CUDA events on the library's stream; max over ranks (ms)."""
    torch = c.torch
    _barrier(c)
    ev0 = torch.cuda.Event(enable_timing=True)
    ev1 = torch.cuda.Event(enable_timing=True)
    ev0.record(stream)
    last = None
    for _ in range(nsteps):
        last = fn()
    ev1.record(stream)
    _barrier(c)
    ms = ev0.elapsed_time(ev1)
    if c.world > 1:
        t = torch.tensor([ms], dtype=torch.float64, device=c.dev)
        c.dist.all_reduce(t, op=c.dist.ReduceOp.MAX)
        ms = float(t.item())
    return ms, last


def _traffic(name):
    try:
        t = json.load(open(os.path.join(ROOT, "profiles", name)))
        return t.get("dram_bytes_per_launch"), t
    except Exception:
        return None, None


# ------------------------------------------------------------------------------------------
def run_crossfit(c: BenchCtx):
    """C1 / C4: the whole composite, one fit per step."""
    torch, args = c.torch, c.args
    from ate_replication_causalml_b200 import parallel
    from ate_replication_causalml_b200.engine import Engine
    n, p0, K = args.n, args.p0, 5
    d, fo, ips, iout = B.make_inputs(n, p0)
    # pinned host buffers (e2e leg) and HBM-resident copies (value leg); column-major base matrix
    xb_host = torch.from_numpy(np.asfortranarray(d["x_raw"]).T.copy()).contiguous().pin_memory()  # (p0, n) == col-major n x p0
    w_host = torch.from_numpy(d["w"]).pin_memory()
    y_host = torch.from_numpy(d["y"]).pin_memory()
    fo_host = torch.from_numpy(fo).pin_memory()
    ips_host = torch.from_numpy(ips).pin_memory()
    iout_host = torch.from_numpy(iout).pin_memory()
    xb_dev, w_dev, y_dev = xb_host.to(c.dev), w_host.to(c.dev), y_host.to(c.dev)
    fo_dev, ips_dev, iout_dev = fo_host.to(c.dev), ips_host.to(c.dev), iout_host.to(c.dev)

    eng = Engine(c.local_rank)
    stream = torch.cuda.ExternalStream(eng.stream_ptr, device=c.dev)
    comm = None
    if c.world > 1:
        comm = parallel.TorchComm(c.dist, c.dev)
        parallel.init_nccl(eng, comm)
    common = dict(n_continuous=d["n_continuous"], expand=args.expand, B=args.B, nlambda=args.nlambda)

    def fit_device():
        return parallel.crossfit_sharded(eng, xb_dev.data_ptr(), w_dev.data_ptr(), y_dev.data_ptr(), n, p0,
                                         fo_dev.data_ptr(), ips_dev.data_ptr(), iout_dev.data_ptr(), comm=comm,
                                         on_device=True, **common)

    def fit_host():
        return parallel.crossfit_sharded(eng, xb_host.numpy().T, w_host.numpy(), y_host.numpy(), n, p0, fo_host.numpy(),
                                         ips_host.numpy(), iout_host.numpy(), comm=comm, on_device=False,
                                         want_nuisances=True, **common)

    for _ in range(args.warmup):
        fit_device()
    sampler = B.ClockSampler(c.local_rank)
    launches0 = eng.launches
    sampler.start()
    ms_total, last = _timed(c, stream, args.steps, fit_device)   # headline: profiling hooks off
    clocks = sampler.stop()
    launches = (eng.launches - launches0) / args.steps
    ms_step = ms_total / args.steps
    value = 1e3 / ms_step

    # ---- a separate profiled fit for the roofline of the dominant kernel ----
    eng.profile_enable(1)
    ms_prof, last_p = _timed(c, stream, 1, fit_device)
    prof = eng.profile_read()
    eng.profile_enable(0)

    e2e = None
    if not args.no_e2e:
        fit_host()  # warm the host path once
        ms_e2e, last_h = _timed(c, stream, args.steps, fit_host)
        h2d = (n * p0 + 2 * n) * 8 + (n + 2 * K * n) * 4          # base covariates, w, y; outer + inner fold ids
        d2h = 3 * n * 8 + args.B * 8 + 2 * K * 8 + 208             # phat, m0, m1; replicates; lambdas; result struct
        e2e = dict(value=1e3 / (ms_e2e / args.steps), unit=B.UNIT, h2d_bytes_per_step=h2d, d2h_bytes_per_step=d2h,
                   ms_per_step=ms_e2e / args.steps,
                   stage_ms={k: last_h.get(k) for k in ("ms_design", "ms_folds", "ms_outcome", "ms_propensity", "ms_score",
                                                        "ms_bootstrap")})

    roof = None
    if c.rank == 0:
        peak_tensor = B.measure_dgemm_peak(torch, c.dev)
        hbm, peak_src = B.hbm_peak()
        gl = max(1, prof["gram_launches"])
        avg_ms = prof["gram_ms"] / gl
        gbs = prof["gram_bytes"] / gl / (avg_ms * 1e-3) / 1e9 if avg_ms > 0 else 0.0
        tf = prof["gram_flops"] / gl / (avg_ms * 1e-3) / 1e12 if avg_ms > 0 else 0.0
        traffic, cap = _traffic("r02_irls_round_traffic.json")
        staged_gb = last_p.get("rows_staged", 0.0) * (B.design_width(p0, args.expand) + 8) * 8 / 1e9
        hbm_obj = dict(achieved=gbs, peak=hbm, unit="GB/s", frac=gbs / hbm, peak_source=peak_src,
                       algorithmic="sum over the launch's live fits of rows * (p * 8 + 24) bytes (SURVEY 8d: every fit "
                                   "reads each of its rows once per IRLS step); the kernel stages a row once for the "
                                   "fits that share a CTA and L2 serves the CTAs walking the same rows, so the DRAM "
                                   "traffic (`traffic`) is a fraction of this",
                       staged_gb_per_fit=staged_gb)
        ten_obj = dict(achieved=tf, peak=peak_tensor, unit="TFLOP/s", frac=tf / peak_tensor if peak_tensor else None,
                       algorithmic="sum over the launch's live fits of rows * (|L| (|L| + 1) + 6 p) flop: the strong-set "
                                   "Gram's lower triangle plus eta, X^T r, X^T v; all of it issues as DMMA.8x8x4 "
                                   "(CTAs serving a single fit run X^T r / X^T v on the FP64 FMA pipe). The padded tiles "
                                   "the kernel really issues are about 2x this count",
                       peak_source="cuBLAS DGEMM 8192^3 measured in this run (MEASURED_PEAKS.json has no fp64 row)")
        # The step is bound by the fp64 tensor (DMMA) pipe: rows are staged once per CTA for up to four fits and
        # L2 serves most repeats, so the ALGORITHMIC bytes (every fit reads its rows) are not DRAM traffic and
        # their rate may exceed the HBM peak; the DMMA work is real work.  `hbm` is kept beside it.
        primary = "tensor"
        p_obj = ten_obj
        roof = dict(bound=primary, kernel="irls_round_kernel (one launch = one IRLS step of every live path fit: eta/q/v/r, "
                                          "X^T r, X^T v and the strong-set Gram X_L^T V X_L on DMMA from staged rows)",
                    achieved=p_obj["achieved"], peak=p_obj["peak"], unit=p_obj["unit"], frac=p_obj["frac"],
                    traffic=traffic, traffic_capture=cap, launches=prof["gram_launches"], avg_launch_ms=avg_ms,
                    share_of_step=prof["gram_ms"] / ms_prof, profiled_fit_ms=ms_prof, hbm=hbm_obj, tensor=ten_obj,
                    dram_gbs=(traffic / (avg_ms * 1e-3) / 1e9 if traffic and avg_ms > 0 else None),
                    dmma_pipe_pct=(cap or {}).get("dmma_pipe_pct"),
                    note="both ceilings are reported: the step is a stream of rows (HBM -> L2 -> shared memory by 1-D TMA) "
                         "feeding fp64 DMMA work; `traffic` / `dram_gbs` are the measured DRAM bytes of one launch "
                         "(ncu) and their rate over the average launch; `dmma_pipe_pct` is ncu's "
                         "sm__pipe_tensor_subpipe_dmma_cycles_active of the same capture (issued DMMAs incl. padded "
                         "tiles and the helpers' X^T r / X^T v tiles, which the algorithmic count leaves out)")

    cpu = None
    one = None
    if c.rank == 0 and c.world == 1 and not args.no_cpu_baseline and args.workload == "C1":
        s = B.cpu_sample(args, args.cpu_sample_n, args.cpu_threads or B.host_threads())
        cpu = dict(value=s["fits_per_s"], unit=B.UNIT, cores=s["threads"], kind="port", sample=B.sample_text(args, s))
        if not args.no_single_thread:
            o = B.cpu_sample(args, args.cpu_sample_n, 1)
            one = dict(value=o["fits_per_s"], unit=B.UNIT, cores=1, kind="port", sample=B.sample_text(args, o))

    metric = B.METRIC if args.workload == "C1" else "cross-fitted AIPW ATE+bootstrap-SE fits/sec at n=16M,p=506 (BASELINE config C4)"
    line = None
    if c.rank == 0:
        keys = ("ms_design", "ms_folds", "ms_outcome", "ms_propensity", "ms_score", "ms_bootstrap")
        line = dict(metric=metric, value=value, unit=B.UNIT, n_gpus=c.world, steps=args.steps, warmup=args.warmup,
                    ms_per_step=ms_step, higher_is_better=True, scaling="strong", vs_baseline=None, dtype="f64",
                    data="synthetic",
                    config=dict(B.workload(args), parallelism=("1 GPU" if c.world == 1 else
                                f"rows sharded over {c.world} GPUs, one NCCL all-reduce of the packed partial sums per IRLS "
                                "round, bootstrap replicates sharded by id")),
                    clocks=clocks, e2e=e2e, gpu_launches=int(launches), roofline=roof, cpu_baseline=cpu,
                    cpu_baseline_1thread=one,
                    result=dict(tau=last["tau"], se_sandwich=last["se_sandwich"], se_bootstrap=last["se_bootstrap"],
                                true_ate=d["true_ate"], rounds_irls=last.get("rounds_irls"),
                                irls_solves=last.get("irls_solves"), strong_mean=last.get("strong_mean"),
                                strong_frac2=last.get("strong_frac2"), step_launches=last.get("step_launches"),
                                list_misses=last.get("list_misses"), row_fits=last.get("row_fits"),
                                rows_staged=last.get("rows_staged"),
                                stage_ms={k: last.get(k) for k in keys},
                                stage_ms_sum=float(sum(last.get(k) or 0.0 for k in keys))))
    eng.close()
    return line


# ------------------------------------------------------------------------------------------
def run_bootstrap(c: BenchCtx):
    """C2: B = 10,000 bootstrap replicates over the n = 1M AIPW scores, replicates sharded by id over the
    ranks.  `value` = replicates/s in the production (Philox) mode; the validation (index-matrix) mode is
    reported as HBM GB/s of the int32 index stream in `roofline`."""
    torch, args = c.torch, c.args
    from ate_replication_causalml_b200.engine import Engine
    from ate_replication_causalml_b200.parallel import replicate_shard, r_sd
    n, Bt = args.n, args.B
    g = torch.Generator(device=c.dev)
    g.manual_seed(1991)
    w = (torch.rand(n, device=c.dev, generator=g, dtype=torch.float64) < 0.4).double()
    p = torch.rand(n, device=c.dev, generator=g, dtype=torch.float64) * 0.8 + 0.1
    m0 = torch.rand(n, device=c.dev, generator=g, dtype=torch.float64) * 0.5 + 0.1
    m1 = m0 + 0.08
    y = (torch.rand(n, device=c.dev, generator=g, dtype=torch.float64) < (w * m1 + (1 - w) * m0)).double()
    eng = Engine(c.local_rank)
    stream = torch.cuda.ExternalStream(eng.stream_ptr, device=c.dev)
    lo, hi = replicate_shard(Bt, c.rank, c.world)
    ptrs = [t.data_ptr() for t in (w, y, p, m0, m1)]

    def philox():
        _, boot = eng.bootstrap(*ptrs, B=hi - lo, seed=1991, rep_offset=lo, on_device=True, n=n)
        if c.world > 1:
            parts = [torch.zeros(replicate_shard(Bt, r, c.world)[1] - replicate_shard(Bt, r, c.world)[0],
                                 dtype=torch.float64, device=c.dev) for r in range(c.world)]
            c.dist.all_gather(parts, torch.from_numpy(boot).to(c.dev))
            boot = torch.cat(parts).cpu().numpy()
        return r_sd(boot)

    for _ in range(args.warmup):
        philox()
    sampler = B.ClockSampler(c.local_rank)
    l0 = eng.launches
    sampler.start()
    ms, se = _timed(c, stream, args.steps, philox)
    clocks = sampler.stop()
    launches = (eng.launches - l0) / args.steps
    ms_step = ms / args.steps
    # validation mode: a device-resident index matrix of Bi replicates (8 GB at Bi = 2000)
    roof = None
    Bi = min(2000, hi - lo)
    idx = torch.randint(1, n + 1, (Bi, n), device=c.dev, dtype=torch.int32, generator=g)
    for _ in range(2):
        eng.bootstrap(*ptrs, B=Bi, idx=idx.data_ptr(), on_device=True, n=n)
    ms_i, _ = _timed(c, stream, 3, lambda: eng.bootstrap(*ptrs, B=Bi, idx=idx.data_ptr(), on_device=True, n=n))
    if c.rank == 0:
        hbm, src = B.hbm_peak()
        gbs = Bi * n * 4 / (ms_i / 3 * 1e-3) / 1e9
        traffic, cap = _traffic("r02_bootstrap_index_traffic.json")
        roof = dict(bound="hbm", kernel="bootstrap_kernel<index mode> (validation mode: R-generated index matrix)",
                    achieved=gbs, peak=hbm, unit="GB/s", frac=gbs / hbm, traffic=traffic, traffic_capture=cap,
                    algorithmic=f"B * n * 4 bytes of int32 indices ({Bi} x {n}); the 8 MB score table stays in L2",
                    peak_source=src, avg_launch_ms=ms_i / 3,
                    philox=dict(draws_per_s=Bt * n / (ms_step * 1e-3), equivalent_GBps=Bt * n * 4 / (ms_step * 1e-3) / 1e9,
                                note="production mode draws the indices on the fly: no index traffic; the GB/s figure is "
                                     "an equivalence (B n 4 bytes / time), not DRAM traffic"))
    line = None
    if c.rank == 0:
        line = dict(metric="bootstrap ATE standard error: replicates/sec, B=10,000 over n=1M AIPW scores (BASELINE config C2)",
                    value=Bt / (ms_step * 1e-3), unit="replicates/s", n_gpus=c.world, steps=args.steps, warmup=args.warmup,
                    ms_per_step=ms_step, higher_is_better=True, scaling="strong", vs_baseline=None, dtype="f64",
                    data="synthetic", config=dict(workload=B.WORKLOADS["C2"]["text"], n=n, B=Bt,
                                                  parallelism=f"replicates sharded by id over {c.world} GPU(s)",
                                                  l2="score table (8 MB) is L2-resident by design; index mode streams "
                                                     f"{Bi * n * 4 / 1e9:.1f} GB per step (exceeds L2)"),
                    clocks=clocks, e2e=None, gpu_launches=int(launches), roofline=roof, cpu_baseline=None,
                    result=dict(se_bootstrap=se))
    eng.close()
    return line


# ------------------------------------------------------------------------------------------
def run_cv_gaussian(c: BenchCtx):
    """C3: 10-fold cv.glmnet (gaussian lasso) on a device-resident n x p design."""
    torch, args = c.torch, c.args
    from ate_replication_causalml_b200 import parallel
    from ate_replication_causalml_b200.engine import Engine
    n, p = args.n, B.WORKLOADS["C3"]["p"]
    G, rank = c.world, c.rank
    chunk = ((n + G - 1) // G + 3) // 4 * 4
    lo, hi = min(n, rank * chunk), min(n, (rank + 1) * chunk)
    nloc = hi - lo
    ld = (p + 2 + 31) // 32 * 32
    g = torch.Generator(device=c.dev)
    g.manual_seed(1991 + rank)
    # this rank's rows of the row-major design [x | y | 1 | 0]: generated on the device, block by block
    X = torch.zeros((nloc, ld), device=c.dev, dtype=torch.float64)
    beta = torch.zeros(p, device=c.dev, dtype=torch.float64)
    gb = torch.Generator(device=c.dev)
    gb.manual_seed(7)
    nz = torch.randperm(p, device=c.dev, generator=gb)[:48]
    beta[nz] = torch.randn(48, device=c.dev, dtype=torch.float64, generator=gb) * 0.3
    step = 1 << 18
    for r0 in range(0, nloc, step):
        r1 = min(nloc, r0 + step)
        xb = torch.randn((r1 - r0, p), device=c.dev, dtype=torch.float64, generator=g)
        X[r0:r1, :p] = xb
        X[r0:r1, p] = xb @ beta + torch.randn(r1 - r0, device=c.dev, dtype=torch.float64, generator=g)
        X[r0:r1, p + 1] = 1.0
        del xb
    gf = torch.Generator(device=c.dev)
    gf.manual_seed(11)
    foldid = (torch.randperm(n, device=c.dev, generator=gf) % 10 + 1).to(torch.int32)   # the same on every rank
    eng = Engine(c.local_rank)
    stream = torch.cuda.ExternalStream(eng.stream_ptr, device=c.dev)
    if c.world > 1:
        parallel.init_nccl(eng, parallel.TorchComm(c.dist, c.dev))

    def fit():
        return eng.cv_glmnet_device(X.data_ptr(), ld, n, p, foldid.data_ptr(), nfolds=10, family="gaussian",
                                    nlambda=args.nlambda, shard_rank=rank, shard_count=G)

    for _ in range(args.warmup):
        fit()
    sampler = B.ClockSampler(c.local_rank)
    l0 = eng.launches
    sampler.start()
    ms, last = _timed(c, stream, args.steps, fit)
    clocks = sampler.stop()
    launches = (eng.launches - l0) / args.steps
    ms_step = ms / args.steps
    eng.profile_enable(2)
    ms_p, _ = _timed(c, stream, 1, fit)
    prof = eng.profile_read()
    eng.profile_enable(0)
    roof = None
    if c.rank == 0:
        peak_tensor = B.measure_dgemm_peak(torch, c.dev)
        gl = max(1, prof["gram_launches"])
        tf = prof["gram_flops"] / (prof["gram_ms"] * 1e-3) / 1e12 if prof["gram_ms"] > 0 else 0.0
        traffic, cap = _traffic("r02_gram_panel_traffic.json")
        roof = dict(bound="tensor", kernel="gram_dmma_kernel (moment matrix [X y 1]'[X y 1] of each inner fold's rows, "
                                           "96-column panels and panel pairs at this width)",
                    achieved=tf, peak=peak_tensor, unit="TFLOP/s", frac=tf / peak_tensor if peak_tensor else None,
                    traffic=traffic, traffic_capture=cap, launches=prof["gram_launches"],
                    avg_launch_ms=prof["gram_ms"] / gl, share_of_step=prof["gram_ms"] / ms_p,
                    algorithmic="rows * p * (p + 1) flop per Gram (lower triangle incl. diagonal)",
                    peak_source="cuBLAS DGEMM 8192^3 measured in this run")
    line = None
    if c.rank == 0:
        line = dict(metric="10-fold cv.glmnet lambda selections/sec at n=4M,p=1024 (BASELINE config C3)",
                    value=1e3 / ms_step, unit="cv.glmnet fits/s", n_gpus=c.world, steps=args.steps, warmup=args.warmup,
                    ms_per_step=ms_step, higher_is_better=True, scaling="strong", vs_baseline=None, dtype="f64",
                    data="synthetic (generated on the device: 48 of 1024 coefficients non-zero, unit noise)",
                    config=dict(workload=B.WORKLOADS["C3"]["text"], n=n, p=p, nfolds=10, nlambda=args.nlambda,
                                parallelism=("1 GPU" if G == 1 else f"rows sharded over {G} GPUs; the ten moment matrices "
                                             "meet in one NCCL all-reduce; the 11 path fits are solved by every rank"),
                                l2=f"inputs_exceed_l2 (design {nloc * ld * 8 / 1e9:.1f} GB per GPU)"),
                    clocks=clocks, e2e=None, gpu_launches=int(launches), roofline=roof, cpu_baseline=None,
                    result=dict(lmu=last["lmu"], idx_min=last["idx_min"], idx_1se=last["idx_1se"],
                                lambda_min=last["lambda_min"], lambda_1se=last["lambda_1se"],
                                nonzero_at_1se=int((last["beta"][:, last["idx_1se"]] != 0).sum())))
    eng.close()
    return line
