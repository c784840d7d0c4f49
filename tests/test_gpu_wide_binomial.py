"""GPU parity (through the C ABI) of the L1-logistic path at the widths the benchmark and the wide
BASELINE configs use -- every column-slice instantiation of the fused IRLS step, the narrow-ring
regime, lists wide enough to be split over several CTAs, and the row-list (no private copy) mode.

The designs are TWIN-FREE (unique products of continuous base columns), so the lasso solution is
unique and the Gram-space solver must reproduce glmnet's iterates: identical lambda counts and
pass counts, coefficients to 1e-8 (bar: glmnet's convergence threshold, ~3e-4)."""
import os

import numpy as np
import pytest

from ate_replication_causalml_b200 import synth

pytestmark = pytest.mark.gpu


def unique_product_design(n, p, seed, dense=0):
    """p unique pairwise products (i <= j) of N(0,1) base columns, standardised; binary response
    from a logistic model with `dense` (or a handful of) non-zero coefficients."""
    p0 = 2
    while p0 * (p0 + 1) // 2 < p:
        p0 += 1
    rng = np.random.Generator(np.random.Philox(key=seed))
    base = rng.standard_normal((n, p0))
    cols = [base[:, i] * base[:, j] for i in range(p0) for j in range(i, p0)][:p]
    X = np.column_stack(cols)
    X = (X - X.mean(0)) / X.std(0)
    nz = dense if dense else max(4, p // 16)
    b = np.zeros(p)
    idx = rng.choice(p, nz, replace=False)
    b[idx] = rng.choice([-1.0, 1.0], nz) * rng.uniform(0.15, 0.5, nz) / np.sqrt(max(1.0, nz / 16.0))
    eta = -0.3 + X @ b
    y = (rng.random(n) < 1.0 / (1.0 + np.exp(-eta))).astype(np.float64)
    return np.asfortranarray(X), y


def _cmp(g, o, tol=1e-8):
    assert g["jerr"] == o["jerr"] == 0
    assert g["lmu"] == o["lmu"]
    assert g["npasses"] == o["npasses"]
    assert np.allclose(g["lambda_"], o["lambda_"], rtol=1e-12)
    assert np.abs(g["beta"] - o["beta"]).max() < tol
    assert np.abs(g["a0"] - o["a0"]).max() < tol
    assert np.abs(g["dev_ratio"] - o["dev_ratio"]).max() < tol


# staged width = round_up(p + 1, 8): 104, 152, 208, 264 (the C1 width), 352 (two-slot ring with
# 32-row slots), 408; 520 is BASELINE config C4's width
@pytest.mark.parametrize("p,n", [(100, 4000), (150, 4000), (200, 4000), (256, 6000), (344, 5000), (400, 5000),
                                 (512, 6000)])
def test_binomial_path_every_staged_width(engine, oracle, p, n):
    X, y = unique_product_design(n, p, seed=300 + p)
    _cmp(engine.glmnet(X, y, "binomial", nlambda=40), oracle.glmnet(X, y, "binomial", nlambda=40))


def test_binomial_path_with_lists_beyond_one_cta(engine, oracle):
    """A dense truth: late in the path the strong set passes 128 columns, so the list's Gram tiles no
    longer fit one CTA's MMA warps and the work is split (ate_round_plan_describe: ngroups >= 2)."""
    X, y = unique_product_design(12_000, 256, seed=77, dense=170)
    g = engine.glmnet(X, y, "binomial")
    o = oracle.glmnet(X, y, "binomial")
    _cmp(g, o)
    assert (g["beta"][:, -1] != 0).sum() > 128


def test_cv_binomial_at_the_benchmark_width(engine, oracle):
    """cv.glmnet(family = "binomial") on 256 twin-free columns: the shape of one C1 propensity problem
    (ate_functions.R:139-141 on the design of :288-296)."""
    n, p = 20_000, 256
    X, y = unique_product_design(n, p, seed=5)
    fid = synth.fold_ids(n, 10, 7)
    g = engine.cv_glmnet(X, y, "binomial", foldid=fid)
    o = oracle.cv_glmnet(X, y, fid, "binomial")
    assert g["lmu"] == o["lmu"] and g["idx_min"] == o["idx_min"] and g["idx_1se"] == o["idx_1se"]
    assert np.allclose(g["lambda_"], o["lambda_"], rtol=1e-12)
    assert np.allclose(g["cvm"], o["cvm"], rtol=1e-9) and np.allclose(g["cvsd"], o["cvsd"], rtol=1e-7)
    assert np.abs(g["beta"] - o["beta"]).max() < 1e-8
    assert np.abs(g["a0"] - o["a0"]).max() < 1e-8


def test_cv_binomial_two_slot_ring(engine, oracle):
    n, p = 9000, 344
    X, y = unique_product_design(n, p, seed=6)
    fid = synth.fold_ids(n, 10, 8)
    g = engine.cv_glmnet(X, y, "binomial", foldid=fid, nlambda=50)
    o = oracle.cv_glmnet(X, y, fid, "binomial", nlambda=50)
    assert g["lmu"] == o["lmu"] and g["idx_min"] == o["idx_min"] and g["idx_1se"] == o["idx_1se"]
    assert np.allclose(g["cvm"], o["cvm"], rtol=1e-9)
    assert np.abs(g["beta"] - o["beta"]).max() < 1e-8


def test_cv_binomial_reading_the_shared_design_through_row_lists(engine, oracle):
    """ATE_NO_PRIVATE_COPY=1: no private contiguous row copies; the step stages rows one bulk copy each."""
    n, p = 8000, 256
    X, y = unique_product_design(n, p, seed=9)
    fid = synth.fold_ids(n, 10, 3)
    o = oracle.cv_glmnet(X, y, fid, "binomial", nlambda=40)
    old = os.environ.get("ATE_NO_PRIVATE_COPY")
    os.environ["ATE_NO_PRIVATE_COPY"] = "1"
    try:
        g = engine.cv_glmnet(X, y, "binomial", foldid=fid, nlambda=40)
    finally:
        if old is None:
            del os.environ["ATE_NO_PRIVATE_COPY"]
        else:
            os.environ["ATE_NO_PRIVATE_COPY"] = old
    assert g["lmu"] == o["lmu"] and g["idx_min"] == o["idx_min"] and g["idx_1se"] == o["idx_1se"]
    assert np.allclose(g["cvm"], o["cvm"], rtol=1e-9)
    assert np.abs(g["beta"] - o["beta"]).max() < 1e-8


def test_crossfit_on_the_benchmark_design(engine):
    """bench.py's C1 design itself (p0 = 16, expand = 1: the 256 ORDERED products, 120 twin pairs) at a
    convergence threshold tight enough that both sides sit on the unique optimum of the fitted values:
    the same lambdas are selected, and ATE / SE agree to the 1e-6 bar -- asserted unconditionally.
    The oracle side takes minutes at this threshold, so it is a committed fixture
    (tools/make_golden_c1.py -> tests/golden/c1_design_crossfit.npz)."""
    from conftest import GOLDEN, make_folds
    o = np.load(os.path.join(GOLDEN, "c1_design_crossfit.npz"))
    n, p0 = int(o["n"]), int(o["p0"])
    d = synth.make_dataset(n, p0)
    fo, ips, iout = make_folds(n)
    g = engine.crossfit_aipw(d["x_raw"], d["w"], d["y"], foldid=fo, inner_ps=ips, inner_out=iout,
                             n_continuous=d["n_continuous"], expand=1, B=0, thresh=float(o["thresh"]))
    assert np.allclose(g["lambda_ps"], o["lambda_ps"], rtol=1e-9)
    assert np.allclose(g["lambda_out"], o["lambda_out"], rtol=1e-9)
    assert np.abs(g["phat"] - o["phat"]).max() < 1e-6
    assert np.abs(g["m0"] - o["m0"]).max() < 1e-6 and np.abs(g["m1"] - o["m1"]).max() < 1e-6
    assert abs(g["tau"] - float(o["tau"])) <= 1e-6 * abs(float(o["tau"]))
    assert abs(g["se_sandwich"] - float(o["se"])) <= 1e-6 * abs(float(o["se"]))
