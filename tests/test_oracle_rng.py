"""The oracle's restatement of R's Mersenne-Twister / sample() against known answers of real R
(the only golden vectors of this path that come from the reference's own runtime)."""
import numpy as np


def test_runif_known_answers(oracle, golden):
    oracle.set_seed(1)
    got = [oracle.unif_rand() for _ in range(3)]
    assert np.allclose(got, golden["rng"]["set.seed(1); runif(3)"], atol=5e-8)
    oracle.set_seed(42)
    assert abs(oracle.unif_rand() - golden["rng"]["set.seed(42); runif(1)"][0]) < 5e-7


def test_sample_permutations_both_sampler_kinds(oracle, golden):
    k = golden["rng"]
    oracle.set_seed(1)
    assert oracle.sample_perm(10, 1).tolist() == k["set.seed(1); sample(10) [R >= 3.6, Rejection]"]
    oracle.set_seed(1)
    assert oracle.sample_perm(10, 0).tolist() == k["set.seed(1); sample(10) [R <= 3.5, Rounding]"]
    oracle.set_seed(123)
    assert oracle.sample_perm(10, 1).tolist() == k["set.seed(123); sample(1:10) [R >= 3.6]"]
    oracle.set_seed(42)
    assert oracle.sample_perm(10, 1).tolist() == k["set.seed(42); sample(1:10) [R >= 3.6]"]


def test_sample_replace_range_and_determinism(oracle):
    oracle.set_seed(7)
    a = oracle.sample_replace(1000, 5000, 1)
    oracle.set_seed(7)
    b = oracle.sample_replace(1000, 5000, 1)
    assert (a == b).all() and a.min() >= 1 and a.max() <= 1000
    # roughly uniform
    assert abs(a.mean() - 500.5) < 20


def test_foldid_is_balanced_permutation(oracle):
    oracle.set_seed(3)
    f = oracle.foldid(1003, 10, 1)
    counts = np.bincount(f, minlength=11)[1:]
    assert counts.sum() == 1003 and counts.max() - counts.min() <= 1


def test_r_mean_sd(oracle):
    x = np.array([1.0, 2.0, 4.0, 8.0, 1e10, -1e10])
    assert abs(oracle.r_mean(x) - 2.5) < 1e-12
    assert abs(oracle.r_sd(np.array([2.0, 4.0, 4.0, 4.0, 5.0, 5.0, 7.0, 9.0])) - np.std([2, 4, 4, 4, 5, 5, 7, 9], ddof=1)) < 1e-14
