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
n, p0 = int(os.environ.get("ATE_CHECK_N", "20000")), 8
d = synth.make_dataset(n, p0)
fo, ips, iout = make_folds(n)
eng = Engine(lr)
comm = parallel.TorchComm(dist, dev)
kw = dict(n_continuous=d["n_continuous"], expand=1, B=200, nlambda=100, want_nuisances=True)
r = parallel.crossfit_sharded(eng, d["x_raw"], d["w"], d["y"], n, p0, fo, ips, iout, comm=comm, on_device=False, **kw)
ref = parallel.crossfit_sharded(eng, d["x_raw"], d["w"], d["y"], n, p0, fo, ips, iout, comm=None, on_device=False, **kw)
rel = lambda a, b: abs(a - b) / abs(b)
ok = (rel(r["tau"], ref["tau"]) < 1e-9 and rel(r["se_sandwich"], ref["se_sandwich"]) < 1e-9
      and rel(r["se_bootstrap"], ref["se_bootstrap"]) < 1e-9 and np.allclose(r["lambda_ps"], ref["lambda_ps"], rtol=1e-9)
      and np.allclose(r["lambda_out"], ref["lambda_out"], rtol=1e-9) and np.abs(r["phat"] - ref["phat"]).max() < 1e-9
      and np.abs(r["m0"] - ref["m0"]).max() < 1e-9)
print(f"rank {rank}/{world}: shared tau {r['tau']:.15f} se_b {r['se_bootstrap']:.15f} | single tau {ref['tau']:.15f} "
      f"se_b {ref['se_bootstrap']:.15f} | max |dphat| {np.abs(r['phat'] - ref['phat']).max():.2e} | agree={ok}", flush=True)
dist.barrier()
eng.close()
dist.destroy_process_group()
sys.exit(0 if ok else 1)
