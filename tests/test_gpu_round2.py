"""GPU tests (through the C ABI) of what round 2 added: rows shared by several ranks, validation of
bootstrap indices, the streamed index matrix, wide glm / lm designs, the device-resident cv.glmnet."""
import os
import subprocess
import sys

import numpy as np
import pytest

from ate_replication_causalml_b200 import synth
from conftest import make_folds

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

_SHARED_WORKER = r"""
import os, sys
sys.path.insert(0, sys.argv[1]); sys.path.insert(0, os.path.join(sys.argv[1], "tests"))
import numpy as np
import torch.distributed as dist
from ate_replication_causalml_b200 import parallel, synth
from ate_replication_causalml_b200.engine import Engine
from conftest import make_folds
rank, W = int(sys.argv[3]), int(sys.argv[4])
dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%s" % sys.argv[2], rank=rank, world_size=W)
comm = parallel.TorchComm(dist, None)
n, p0 = 6000, 7
d = synth.make_dataset(n, p0)
fo, ips, iout = make_folds(n)
eng = Engine(0)                                   # every rank on cuda:0: the row slices and the sums are what is tested
parallel.install_allreduce(eng, comm)             # host transport (gloo); on a multi-GPU box the library uses NCCL
kw = dict(n_continuous=d["n_continuous"], expand=0, B=40, want_nuisances=True)
r = parallel.crossfit_sharded(eng, d["x_raw"], d["w"], d["y"], n, p0, fo, ips, iout, comm=comm, **kw)
ref = parallel.crossfit_sharded(eng, d["x_raw"], d["w"], d["y"], n, p0, fo, ips, iout, comm=None, **kw)
rel = lambda a, b: abs(a - b) / abs(b)
assert rel(r["tau"], ref["tau"]) < 1e-10 and rel(r["se_sandwich"], ref["se_sandwich"]) < 1e-10, (r["tau"], ref["tau"])
assert rel(r["se_bootstrap"], ref["se_bootstrap"]) < 1e-10
assert np.allclose(r["lambda_ps"], ref["lambda_ps"], rtol=1e-10) and np.allclose(r["lambda_out"], ref["lambda_out"], rtol=1e-10)
assert np.abs(r["phat"] - ref["phat"]).max() < 1e-10 and np.abs(r["m0"] - ref["m0"]).max() < 1e-10
assert np.abs(r["boot"] - ref["boot"]).max() < 1e-10        # same replicate ids, whoever drew them
eng.close()
dist.destroy_process_group()
print("ok", rank)
"""


@pytest.mark.parametrize("world", [2, 3])
def test_rows_shared_by_several_ranks_reproduce_the_single_rank_fit(tmp_path, world):
    """shard_count > 1: every rank keeps a slice of the rows and the per-round partial sums are summed
    over the ranks.  Here the ranks are processes on ONE GPU talking through the host all-reduce hook
    (gloo); the fit must agree with the unshared one to summation-order rounding (1e-10; bar 1e-6)."""
    script = tmp_path / "worker.py"
    script.write_text(_SHARED_WORKER)
    port = str(31000 + (os.getpid() + 13 * world) % 2000)
    procs = [subprocess.Popen([sys.executable, str(script), ROOT, port, str(r), str(world)], stdout=subprocess.PIPE,
                              stderr=subprocess.STDOUT, text=True) for r in range(world)]
    outs = [p.communicate(timeout=600)[0] for p in procs]
    for p, o in zip(procs, outs):
        assert p.returncode == 0, o[-3000:]


def test_bootstrap_index_mode_rejects_out_of_range_indices(engine):
    from ate_replication_causalml_b200.engine import AteError
    rng = np.random.default_rng(0)
    n, B = 5000, 6
    w = (rng.random(n) < 0.5).astype(float); y = (rng.random(n) < 0.4).astype(float)
    p = rng.uniform(0.2, 0.8, n); m0 = rng.uniform(0.1, 0.5, n); m1 = m0 + 0.1
    idx = rng.integers(1, n + 1, (B, n)).astype(np.int32)
    engine.bootstrap(w, y, p, m0, m1, idx=idx)                       # fine
    for bad in (0, n + 1, -2147483648):                              # 0, one past the end, NA_integer_
        j = idx.copy()
        j[3, 1234] = bad
        with pytest.raises(AteError, match="outside 1"):
            engine.bootstrap(w, y, p, m0, m1, idx=j)
    engine.bootstrap(w, y, p, m0, m1, idx=idx)                       # the ctx is still usable afterwards


def test_bootstrap_index_matrix_streamed_in_blocks_matches_oracle(engine, oracle):
    """A host index matrix larger than the 256 MB staging blocks (n = 2^21 + 3: odd row length, so rows
    are not 16-byte aligned either) gives the oracle's replicates."""
    rng = np.random.default_rng(1)
    n, B = (1 << 21) + 3, 70
    w = (rng.random(n) < 0.5).astype(float); y = (rng.random(n) < 0.4).astype(float)
    p = rng.uniform(0.2, 0.8, n); m0 = rng.uniform(0.1, 0.5, n); m1 = m0 + 0.1
    idx = rng.integers(1, n + 1, (B, n)).astype(np.int32)
    se, boot = engine.bootstrap(w, y, p, m0, m1, idx=idx)
    sd, ob = oracle.bootstrap_idx(w, y, p, m0, m1, idx)
    assert np.abs(boot - ob).max() < 1e-12 and abs(se - sd) <= 1e-10 * sd


@pytest.mark.parametrize("p", [600, 1100])
def test_glm_and_lm_on_designs_wider_than_512_columns(engine, oracle, p):
    """columns beyond 512 used to be dropped silently from the row dot products."""
    rng = np.random.default_rng(p)
    n = 6000
    X = rng.standard_normal((n, p))
    b = np.zeros(p); b[-40:] = rng.standard_normal(40) * 0.2       # the signal sits in the LAST columns
    yb = (rng.random(n) < 1 / (1 + np.exp(-(X @ b - 0.2)))).astype(float)
    g, o = engine.glm_binomial(X, yb), oracle.glm_binomial(X, yb)
    assert g["iters"] == o["iters"] and np.abs(g["coef"] - o["coef"]).max() < 1e-7
    yl = X @ b + rng.standard_normal(n)
    gl, ol = engine.lm(X, yl), oracle.lm(X, yl)
    assert np.abs(gl["coef"] - ol["coef"]).max() < 1e-8 and np.abs(gl["std_err"] - ol["std_err"]).max() < 1e-8
    assert abs(gl["sigma"] - ol["sigma"]) < 1e-9


def test_cv_glmnet_on_a_device_resident_design_equals_the_host_call(engine):
    import torch
    n, p0 = 5000, 9
    d = synth.make_dataset(n, p0)
    fid = synth.fold_ids(n, 10, 7)
    dev = torch.device("cuda:0")
    ld = 32
    for fam, yy in (("gaussian", d["y"]), ("binomial", d["w"])):
        h = engine.cv_glmnet(d["x_raw"], yy, fam, foldid=fid)
        X = torch.zeros((n, ld), dtype=torch.float64, device=dev)
        X[:, :p0] = torch.from_numpy(d["x_raw"]).to(dev)
        X[:, p0] = torch.from_numpy(yy).to(dev)
        X[:, p0 + 1] = 1.0
        f = torch.from_numpy(fid).to(dev)
        g = engine.cv_glmnet_device(X.data_ptr(), ld, n, p0, f.data_ptr(), nfolds=10, family=fam)
        assert g["lmu"] == h["lmu"] and g["idx_min"] == h["idx_min"] and g["idx_1se"] == h["idx_1se"]
        assert np.abs(g["beta"] - h["beta"]).max() < 1e-12 and np.allclose(g["cvm"], h["cvm"], rtol=1e-12)


def test_crossfit_draws_identical_folds_on_identical_calls(engine):
    """with the ids drawn by the library (R's Mersenne-Twister from r_seed) two identical
    calls must select the same lambdas and give the same estimate."""
    n, p0 = 3000, 6
    d = synth.make_dataset(n, p0)
    kw = dict(n_continuous=d["n_continuous"], expand=0, B=0, r_seed=77)
    a = engine.crossfit_aipw(d["x_raw"], d["w"], d["y"], **kw)
    b = engine.crossfit_aipw(d["x_raw"], d["w"], d["y"], **kw)
    assert a["tau"] == b["tau"] and np.array_equal(a["lambda_ps"], b["lambda_ps"])
    fo = engine.r_foldid(10, 5)     # and the stream the user sees is untouched in kind
    assert fo.min() == 1 and fo.max() == 5


def test_fold_fits_joining_their_full_data_fit_change_nothing(engine):
    """The cv-fold fits of a binomial cv.glmnet start as soon as their full-data fit has fixed the lambda
    sequence (three lambdas recorded) instead of after it has finished (ATE_TWO_PHASE=1, the order R runs
    them in): same lambdas, same fold errors, same coefficients, bit for bit -- a fit's iterates depend on
    its own sums only."""
    from test_gpu_wide_binomial import unique_product_design
    n, p = 12_000, 120
    X, y = unique_product_design(n, p, seed=21)
    fid = synth.fold_ids(n, 10, 4)
    # the switch is read once per process: run the sequential order in a child process
    code = (
        "import sys, numpy as np; sys.path.insert(0, %r); sys.path.insert(0, %r)\n"
        "from ate_replication_causalml_b200.engine import Engine\n"
        "d = np.load(sys.argv[1]); eng = Engine(0)\n"
        "g = eng.cv_glmnet(np.asfortranarray(d['X']), d['y'], 'binomial', foldid=d['fid'])\n"
        "np.savez(sys.argv[2], cvm=g['cvm'], cvsd=g['cvsd'], beta=g['beta'], a0=g['a0'], lam=g['lambda_'], idx=np.array([g['idx_min'], g['idx_1se'], g['lmu']]))\n"
    ) % (ROOT, os.path.join(ROOT, "tests"))
    import tempfile
    with tempfile.TemporaryDirectory() as td:
        inp, out = os.path.join(td, "in.npz"), os.path.join(td, "out.npz")
        np.savez(inp, X=X, y=y, fid=fid)
        env = dict(os.environ, ATE_TWO_PHASE="1")
        r = subprocess.run([sys.executable, "-c", code, inp, out], env=env, capture_output=True, text=True, timeout=600)
        assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
        seq = np.load(out)
        g = engine.cv_glmnet(X, y, "binomial", foldid=fid)
        assert [g["idx_min"], g["idx_1se"], g["lmu"]] == list(seq["idx"])
        assert np.array_equal(g["lambda_"], seq["lam"])
        assert np.array_equal(g["cvm"], seq["cvm"]) and np.array_equal(g["cvsd"], seq["cvsd"])
        assert np.array_equal(g["beta"], seq["beta"]) and np.array_equal(g["a0"], seq["a0"])
