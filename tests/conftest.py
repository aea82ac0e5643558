import pathlib
import sys

import pytest

ROOT = pathlib.Path(__file__).resolve().parents[1]
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")
    config.addinivalue_line("markers", "multigpu: needs at least 2 CUDA devices")
    config.addinivalue_line("markers", "slow: long-running")


def pytest_collection_modifyitems(config, items):
    try:
        import torch

        have = torch.cuda.is_available()
        ngpu = torch.cuda.device_count() if have else 0
    except Exception:  # pragma: no cover
        have, ngpu = False, 0
    skip_gpu = pytest.mark.skip(reason="no CUDA device")
    skip_multi = pytest.mark.skip(reason="needs >= 2 CUDA devices")
    for item in items:
        if "gpu" in item.keywords and not have:
            item.add_marker(skip_gpu)
        if "multigpu" in item.keywords and ngpu < 2:
            item.add_marker(skip_multi)


@pytest.fixture(scope="session")
def ref():
    """The unmodified reference's ``mpitree.tree`` module (oracle), or skip."""
    from baseline.oracle import load_reference

    mod = load_reference()
    if mod is None:
        pytest.skip("reference not available (baseline/_ref missing)")
    return mod


@pytest.fixture(scope="session")
def iris():
    from sklearn.datasets import load_iris

    return load_iris()
