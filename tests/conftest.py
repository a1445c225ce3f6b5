import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def _has_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


@pytest.fixture(scope="session")
def oracle():
    """CPU oracle (test infrastructure): the -ffp-contract=off build, the exact checker."""
    from oracle import loader
    return loader.load_oracle()


@pytest.fixture(scope="session")
def oracle_fma():
    from oracle import loader
    return loader.load_oracle(fma=True)


@pytest.fixture(scope="session")
def juceref():
    """The reference's own compiled JUCE objects, when oracle/_ref was built (container only)."""
    from oracle import loader
    lib = loader.load_juceref()
    if lib is None:
        pytest.skip("oracle/_ref/libzl_juceref.so not built (needs /root/reference)")
    return lib


@pytest.fixture(scope="session")
def golden():
    return np.load(os.path.join(ROOT, "tests", "golden", "juce_golden.npz"))


@pytest.fixture(scope="session")
def product():
    """The CUDA product library. GPU tests must run the native code: a missing .so is an error, not a skip."""
    if not _has_gpu():
        pytest.skip("no CUDA device")
    import libzl_b200
    return libzl_b200.load()


@pytest.fixture(scope="session")
def thumb_golden():
    """Vectors from the reference's own compiled juce::AudioThumbnail (tests/golden/make_thumb_golden.py)."""
    return np.load(os.path.join(ROOT, "tests", "golden", "juce_thumb_golden.npz"))


@pytest.fixture(scope="session")
def juceref_optional():
    """oracle/_ref/libzl_juceref.so when it was built (it travels to the GPU box), else None."""
    from oracle import loader
    return loader.load_juceref()
