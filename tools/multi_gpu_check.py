"""torchrun check: a fit whose rows are shared by N ranks equals the single-GPU fit to rounding (run under
torchrun).  The sharded fit sums its partial Grams in another order, so agreement is ~1e-12 relative,
not bit for bit; the lambdas selected and the bootstrap replicate ids are the same."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch, torch.distributed as dist
from ate_replication_causalml_b200.engine import Engine
from ate_replication_causalml_b200 import parallel, synth
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
from conftest import make_folds

rank = int(os.environ["RANK"]); world = int(os.environ["WORLD_SIZE"]); lr = int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr)
dev = torch.device("cuda", lr)
dist.init_process_group("nccl", device_id=dev)
eng = Engine(lr)
comm = parallel.TorchComm(dist, dev)
rel = lambda a, b: abs(a - b) / abs(b)
all_ok = True
# (a) a twin-free design (no products): the shared fit follows the same iterates, so 1e-9;
# (b) the benchmark's ordered products (twin columns x_i x_j / x_j x_i): rounding in the summed moments may
#     admit the other twin, the fitted values then agree to the convergence threshold -- the 1e-6 bar.
for tag, n, p0, expand, tol in (("twin-free", int(os.environ.get("ATE_CHECK_N", "20000")), 12, 0, 1e-9),
                                ("products", int(os.environ.get("ATE_CHECK_N", "20000")), 8, 1, 1e-6)):
    d = synth.make_dataset(n, p0)
    fo, ips, iout = make_folds(n)
    kw = dict(n_continuous=d["n_continuous"], expand=expand, B=200, nlambda=100, want_nuisances=True)
    r = parallel.crossfit_sharded(eng, d["x_raw"], d["w"], d["y"], n, p0, fo, ips, iout, comm=comm, on_device=False, **kw)
    ref = parallel.crossfit_sharded(eng, d["x_raw"], d["w"], d["y"], n, p0, fo, ips, iout, comm=None, on_device=False, **kw)
    diffs = dict(tau=rel(r["tau"], ref["tau"]), se=rel(r["se_sandwich"], ref["se_sandwich"]),
                 se_boot=rel(r["se_bootstrap"], ref["se_bootstrap"]),
                 lam_ps=float(np.abs(r["lambda_ps"] / ref["lambda_ps"] - 1).max()),
                 lam_out=float(np.abs(r["lambda_out"] / ref["lambda_out"] - 1).max()),
                 phat=float(np.abs(r["phat"] - ref["phat"]).max()), m0=float(np.abs(r["m0"] - ref["m0"]).max()),
                 m1=float(np.abs(r["m1"] - ref["m1"]).max()))
    # pointwise fitted values on the twin design agree to the convergence-threshold level (sqrt(thresh) =
    # 3e-4, the bar DESIGN.md section 2 states for coefficients); the estimates themselves to 1e-6
    tol_of = lambda k: 3e-4 if (expand and k in ("phat", "m0", "m1")) else tol
    ok = all(v < tol_of(k) for k, v in diffs.items())
    all_ok &= ok
    print(f"rank {rank}/{world} {tag} n={n} p0={p0} expand={expand}: tau {r['tau']:.15f} vs {ref['tau']:.15f} | "
          + " ".join(f"{k}={v:.1e}" for k, v in diffs.items()) + f" | tol {tol:g} agree={ok}", flush=True)
dist.barrier()
eng.close()
dist.destroy_process_group()
sys.exit(0 if all_ok else 1)
