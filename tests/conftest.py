import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def oracle():
    from oracle import oracle as O
    O.build()
    return O


@pytest.fixture(scope="session")
def engine():
    """One Engine (C-ABI ctx) for the GPU tests.  Fails -- never skips -- when the CUDA library
    or device is missing: a GPU run that silently fell back would prove nothing."""
    from ate_replication_causalml_b200 import build as b
    b.build()
    from ate_replication_causalml_b200.engine import Engine
    eng = Engine(0)
    yield eng
    eng.close()


@pytest.fixture(scope="session")
def golden():
    import json
    import numpy as np
    return dict(
        rng=json.load(open(os.path.join(GOLDEN, "rng_known_answers.json"))),
        sk=np.load(os.path.join(GOLDEN, "sklearn_paths.npz")),
        reg=np.load(os.path.join(GOLDEN, "oracle_regression.npz")),
    )


def make_folds(n, K=5, F=10):
    """outer fold ids + inner cv.glmnet ids in the layout ate_crossfit_aipw expects."""
    import numpy as np
    from ate_replication_causalml_b200 import synth
    fo = synth.fold_ids(n, K, 11)
    ips = np.zeros((K, n), np.int32)
    iout = np.zeros((K, n), np.int32)
    for k in range(K):
        nt = int((fo != k + 1).sum())
        ips[k, :nt] = synth.fold_ids(nt, F, 100 + k)
        iout[k, :nt] = synth.fold_ids(nt, F, 200 + k)
    return fo, ips, iout
