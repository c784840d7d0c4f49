"""One fit shared by several GPUs.

One process per GPU (torchrun).  Every rank passes the same inputs; the library keeps rows
[r c, (r + 1) c) (c = ceil(n / ranks)) of the design and of every cv.glmnet problem's row copy on rank
r.  What crosses the ranks, all inside the library and on its own stream:
  * per IRLS round ONE all-reduce of the packed partial sums of every live path fit (strong-set
    Grams, X^T r, X^T v, deviances), after which every rank runs the same p x p bookkeeping and
    coordinate descent on the same numbers and takes the same decisions;
  * the moment matrices of the gaussian outcome fits and the held-out deviances (one all-reduce each);
  * the stitched nuisances (all-gather) -- the score is then evaluated on all rows by every rank;
  * the bootstrap replicates: rank r draws replicates [r B / N, (r + 1) B / N) with the Philox stream
    keyed by replicate id, so the union is identical to the 1-GPU run; one all-gather.
The collectives are NCCL calls issued by the library (ate_ctx_nccl_init); this module only carries
the 128-byte communicator id from rank 0 to the others through torch.distributed.  Without NCCL
(CPU tests with gloo, hosts that only have MPI) the library falls back to a host all-reduce hook.
Summation order differs from the 1-GPU run, so results agree to ~1e-12 relative, not bit for bit.
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

    def broadcast_bytes(self, data: bytes | None, nbytes: int) -> bytes:
        """Rank 0's `data` to everyone."""
        import torch
        if self.device is not None and self.dist.get_backend() == "nccl":
            t = torch.zeros(nbytes, dtype=torch.uint8, device=self.device)
            if self.rank == 0:
                t.copy_(torch.frombuffer(bytearray(data), dtype=torch.uint8))
            self.dist.broadcast(t, src=0)
            return bytes(t.cpu().numpy().tobytes())
        t = torch.zeros(nbytes, dtype=torch.uint8)
        if self.rank == 0:
            t.copy_(torch.frombuffer(bytearray(data), dtype=torch.uint8))
        self.dist.broadcast(t, src=0)
        return bytes(t.numpy().tobytes())


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
    """Bind the library's host all-reduce hook to `comm` (the fallback transport).  Keeps the callback
    alive on the engine."""
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


def init_nccl(eng, comm):
    """Give the engine's ctx its own NCCL communicator over the ranks of `comm` (collective call)."""
    ident = None
    if comm.rank == 0:
        buf = C.create_string_buffer(128)
        if eng._lib.ate_nccl_unique_id(buf) != 0:
            raise L.AteError("ate_nccl_unique_id failed (libnccl.so.2 not loadable?)")
        ident = buf.raw
    ident = comm.broadcast_bytes(ident, 128)
    rc = eng._lib.ate_ctx_nccl_init(eng._h, C.c_char_p(ident), comm.rank, comm.world)
    if rc != 0:
        msg = eng._lib.ate_last_error(eng._h)
        raise L.AteError(f"ate_ctx_nccl_init failed: {msg.decode() if msg else rc}")
    eng._nccl_ready = True


def crossfit_sharded(eng, x_base, w, y, n, p0, foldid, inner_ps, inner_out, comm=None, on_device=False,
                     want_nuisances=False, **opts):
    """One cross-fitted AIPW fit on `comm.world` GPUs (or one GPU when comm is None).  Every rank gets
    the complete result."""
    if comm is None or comm.world == 1:
        return eng.crossfit_aipw(x_base, w, y, foldid=foldid, inner_ps=inner_ps, inner_out=inner_out,
                                 n=n, p0=p0, on_device=int(on_device), boot_mode=1, want_nuisances=want_nuisances, **opts)
    if not getattr(eng, "_nccl_ready", False) and getattr(eng, "_allreduce_cb", None) is None:
        init_nccl(eng, comm)
    return eng.crossfit_aipw(x_base, w, y, foldid=foldid, inner_ps=inner_ps, inner_out=inner_out, n=n, p0=p0,
                             on_device=int(on_device), boot_mode=1, want_nuisances=want_nuisances,
                             shard_rank=comm.rank, shard_count=comm.world, **opts)
