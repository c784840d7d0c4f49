"""One fit shared by several GPUs.

One process per GPU (torchrun); every rank holds the full inputs (n = 1M x p0 = 16 doubles is
128 MB) and builds its own design.  What is sharded:
  * all 55 lambda paths of the propensity model -- the 5 full-data fits, then the 50 cv-fold
    paths, each set dealt round-robin over ranks inside the library; the full-data paths (1 MB)
    and the held-out deviances of all folds (K * nfolds * nlambda doubles) meet through the host
    all-reduce hook, after which every rank holds the same models and selects the same lambda;
  * the bootstrap replicates -- rank r draws replicates [r*B/N, (r+1)*B/N) with the Philox
    stream keyed by replicate id, so the union is identical to the 1-GPU run.
The K gaussian outcome fits are dealt over ranks the same way (their selected models meet through
the hook); design construction, held-out prediction and the score are replicated.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib as L


class TorchComm:
    """all-reduce / all-gather over torch.distributed (nccl on GPUs, gloo in CPU tests)."""

    def __init__(self, dist, device=None):
        self.dist = dist
        self.device = device
        self.rank = dist.get_rank()
        self.world = dist.get_world_size()

    def allreduce_sum_host(self, arr: np.ndarray):
        import torch
        t = torch.from_numpy(arr)
        if self.device is not None and self.dist.get_backend() == "nccl":
            g = t.to(self.device)
            self.dist.all_reduce(g, op=self.dist.ReduceOp.SUM)
            t.copy_(g.cpu())
        else:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.SUM)
        return arr

    def allgather_host(self, arr: np.ndarray, counts):
        """Concatenate per-rank 1-d arrays (lengths `counts`)."""
        out = np.zeros(int(sum(counts)))
        off = int(sum(counts[: self.rank]))
        out[off: off + counts[self.rank]] = arr
        return self.allreduce_sum_host(out)


def replicate_shard(B: int, rank: int, world: int):
    """Replicates [lo, hi) handled by `rank`; sizes differ by at most one."""
    base, rem = divmod(B, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def r_sd(x: np.ndarray) -> float:
    """sd() as R computes it (two passes, n - 1)."""
    x = np.asarray(x, dtype=np.float64)
    m = float(np.mean(x))
    m += float(np.mean(x - m))
    return float(np.sqrt(np.sum((x - m) ** 2) / (x.size - 1)))


def install_allreduce(eng, comm):
    """Bind the library's multi-GPU hook to `comm`.  Keeps the callback alive on the engine."""
    def _cb(user, buf, count):
        try:
            arr = np.ctypeslib.as_array(buf, shape=(int(count),))
            comm.allreduce_sum_host(arr)
            return 0
        except Exception:  # never let an exception cross the C boundary
            return 1
    cb = L.ALLREDUCE_FN(_cb)
    eng._allreduce_cb = cb
    rc = eng._lib.ate_ctx_set_allreduce(eng._h, C.cast(cb, C.c_void_p), None)
    if rc != 0:
        raise L.AteError("ate_ctx_set_allreduce failed")


def crossfit_sharded(eng, x_base, w, y, n, p0, foldid, inner_ps, inner_out, comm=None, on_device=False,
                     **opts):
    """One cross-fitted AIPW fit on `comm.world` GPUs (or one GPU when comm is None)."""
    B = int(opts.get("B", 1000))
    if comm is None or comm.world == 1:
        return eng.crossfit_aipw(x_base, w, y, foldid=foldid, inner_ps=inner_ps, inner_out=inner_out,
                                 n=n, p0=p0, on_device=int(on_device), boot_mode=1, want_nuisances=False, **opts)
    if getattr(eng, "_allreduce_cb", None) is None:
        install_allreduce(eng, comm)
    lo, hi = replicate_shard(B, comm.rank, comm.world)
    r = eng.crossfit_aipw(x_base, w, y, foldid=foldid, inner_ps=inner_ps, inner_out=inner_out, n=n, p0=p0,
                          on_device=int(on_device), boot_mode=1, want_nuisances=False,
                          shard_rank=comm.rank, shard_count=comm.world, boot_rep_offset=lo,
                          boot_reps_here=hi - lo, **opts)
    if B > 0:
        counts = [replicate_shard(B, q, comm.world) for q in range(comm.world)]
        counts = [b - a for a, b in counts]
        boot = comm.allgather_host(np.asarray(r["boot"], dtype=np.float64), counts)
        r["boot"] = boot
        r["se_bootstrap"] = r_sd(boot)
        r["lower_ci"] = r["tau"] - 1.96 * r["se_bootstrap"]
        r["upper_ci"] = r["tau"] + 1.96 * r["se_bootstrap"]
    return r
