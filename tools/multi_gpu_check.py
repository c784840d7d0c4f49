"""torchrun check: a fit shared by N ranks equals the single-GPU fit (run under torchrun)."""
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
n, p0 = 20000, 8
d = synth.make_dataset(n, p0)
fo, ips, iout = make_folds(n)
eng = Engine(lr)
comm = parallel.TorchComm(dist, dev)
kw = dict(n_continuous=d["n_continuous"], expand=1, B=200, nlambda=100)
r = parallel.crossfit_sharded(eng, d["x_raw"], d["w"], d["y"], n, p0, fo, ips, iout, comm=comm, on_device=False, **kw)
ref = parallel.crossfit_sharded(eng, d["x_raw"], d["w"], d["y"], n, p0, fo, ips, iout, comm=None, on_device=False, **kw)
ok = (r["tau"] == ref["tau"] and r["se_sandwich"] == ref["se_sandwich"] and np.array_equal(r["boot"], ref["boot"])
      and abs(r["se_bootstrap"] - ref["se_bootstrap"]) < 1e-15 and np.array_equal(r["lambda_ps"], ref["lambda_ps"]))
print(f"rank {rank}/{world}: sharded tau {r['tau']:.12f} se_b {r['se_bootstrap']:.12f} | single tau {ref['tau']:.12f} se_b {ref['se_bootstrap']:.12f} | identical={ok}", flush=True)
dist.barrier()
eng.close()
dist.destroy_process_group()
sys.exit(0 if ok else 1)
