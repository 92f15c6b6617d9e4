import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run with -m gpu on the GPU box)")


def pytest_collection_modifyitems(config, items):
    import torch
    if torch.cuda.is_available():
        return
    skip = pytest.mark.skip(reason="no CUDA device in this container (gpu tests run on the B200 box)")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


def load_golden(name):
    return np.load(os.path.join(GOLDEN, name))


@pytest.fixture(scope="session")
def wheat():
    g = load_golden("wheat.npz")
    n, p = int(g["n"]), int(g["p"])
    X = np.unpackbits(g["x_bits"], axis=1)[:, :p].astype(np.int8)
    assert X.shape == (n, p)
    return X, g["y"].astype(np.float64)


@pytest.fixture(scope="session")
def golden():
    return load_golden


@pytest.fixture(scope="session")
def wheat_csv(tmp_path_factory, wheat):
    """The WHEAT file rebuilt from the packed fixture (headerless: phenotype, then markers as 1.0 / 0.0)."""
    X, y = wheat
    path = tmp_path_factory.mktemp("csv") / "WHEAT_combined.csv"
    with open(path, "w") as fp:
        for yi, row in zip(y, X):
            fp.write(repr(float(yi)) + "," + ",".join("1.0" if v else "0.0" for v in row) + "\n")
    return str(path)
