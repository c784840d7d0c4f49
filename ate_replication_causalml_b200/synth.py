"""Synthetic "social-pressure-shaped" inputs.

Shaped by the reference's data preparation: 15 continuous + 6 binary covariates
(/root/reference/ate_replication.Rmd:49-58), continuous ones standardised with scale()
(:72-74), binary treatment W and binary outcome Y (:90-93), and treatment assignment that
depends on the covariates (the selection-bias injection of :99-121).

The generator is counter-based (numpy Philox) so the CPU oracle, the tests and the GPU box
all see bit-identical inputs without shipping files.  Seed 1991 echoes Rmd:42.
"""
from __future__ import annotations

import numpy as np


def base_covariates(n: int, p0: int, seed: int = 1991):
    """n x p0 column-major float64; the first ceil(5/7 p0) columns are N(0,1) draws, the rest
    Bernoulli(0.5).  (p0 = 21 gives exactly 15 + 6.)  Returns (x, n_continuous)."""
    rng = np.random.Generator(np.random.Philox(key=seed))
    ncont = int(np.ceil(p0 * 5.0 / 7.0))
    x = np.empty((n, p0), order="F")
    for j in range(p0):
        if j < ncont:
            x[:, j] = rng.standard_normal(n)
        else:
            x[:, j] = (rng.random(n) < 0.5).astype(np.float64)
    return x, ncont


def treatment_outcome(x: np.ndarray, seed: int = 1991, effect: float = 0.4):
    """W ~ Bernoulli(plogis(-0.5 + x a)), Y ~ Bernoulli(plogis(-0.8 + x b + effect W)) with
    a, b sparse over the base covariates plus a few pairwise products, so that both a main
    design and a degree-2 design carry signal.  Returns (w, y, true_ate)."""
    n, p0 = x.shape
    rng = np.random.Generator(np.random.Philox(key=seed + 1))
    a = np.zeros(p0)
    b = np.zeros(p0)
    nz = max(2, p0 // 4)
    ia = rng.choice(p0, nz, replace=False)
    ib = rng.choice(p0, nz, replace=False)
    a[ia] = rng.choice([-0.35, 0.35], nz)
    b[ib] = rng.choice([-0.3, 0.3], nz)
    # a handful of interaction terms
    npair = max(1, p0 // 4)
    pa = rng.choice(p0, (npair, 2))
    pb = rng.choice(p0, (npair, 2))
    ca = rng.choice([-0.25, 0.25], npair)
    cb = rng.choice([-0.25, 0.25], npair)
    eta_w = -0.5 + x @ a
    eta_y = -0.8 + x @ b
    for t in range(npair):
        eta_w += ca[t] * x[:, pa[t, 0]] * x[:, pa[t, 1]]
        eta_y += cb[t] * x[:, pb[t, 0]] * x[:, pb[t, 1]]
    pw = 1.0 / (1.0 + np.exp(-eta_w))
    w = (rng.random(n) < pw).astype(np.float64)
    py1 = 1.0 / (1.0 + np.exp(-(eta_y + effect)))
    py0 = 1.0 / (1.0 + np.exp(-eta_y))
    u = rng.random(n)
    y = np.where(w > 0.5, u < py1, u < py0).astype(np.float64)
    return w, y, float(np.mean(py1 - py0))


def make_dataset(n: int, p0: int, seed: int = 1991):
    """Unscaled base covariates + (w, y).  Continuous columns get a location/scale so that
    standardisation has something to do."""
    x, ncont = base_covariates(n, p0, seed)
    w, y, ate = treatment_outcome(x, seed)
    rng = np.random.Generator(np.random.Philox(key=seed + 2))
    loc = rng.uniform(-3.0, 3.0, ncont)
    sc = rng.uniform(0.5, 4.0, ncont)
    raw = x.copy(order="F")
    raw[:, :ncont] = raw[:, :ncont] * sc + loc
    return dict(x_raw=raw, n_continuous=ncont, w=w, y=y, true_ate=ate)


def fold_ids(n: int, nfolds: int, seed: int):
    """Balanced random fold ids 1..nfolds (int32) in cv.glmnet's shape, i.e. a permutation of
    rep(seq(nfolds), length = n); drawn with Philox, not R's Mersenne-Twister."""
    rng = np.random.Generator(np.random.Philox(key=seed))
    return (rng.permutation(n) % nfolds + 1).astype(np.int32)
