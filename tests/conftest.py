import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200); run with -m gpu on the GPU box")


def _has_gpu():
    try:
        import orb_slam2_2021_b200 as pkg

        return pkg.lib().orbx_device_count() > 0
    except Exception:
        return False


@pytest.fixture(scope="session")
def has_gpu():
    return _has_gpu()


@pytest.fixture(scope="session", autouse=True)
def _built():
    """Both libraries are built in-tree before any test (nvcc cross-compiles without a GPU)."""
    import orb_slam2_2021_b200 as pkg
    from oracle import oracle

    if not os.path.exists(pkg.SO_PATH):
        pkg.build()
    oracle.build()


@pytest.fixture(scope="session", autouse=True)
def _reference_anchor_present():
    """On a GPU box the parity anchor (oracle/_ref: the reference's own translation units, built in the container that has
    /root/reference and shipped with the snapshot) must be there: its absence is a failure, not a silent skip of every
    reference-pin test."""
    if _has_gpu():
        from oracle import ref

        if not (ref.available() and ref.gpu_available()):
            pytest.fail("oracle/_ref is missing on a GPU box: run `make -C oracle _ref` where /root/reference exists", pytrace=False)


def pytest_collection_modifyitems(config, items):
    if _has_gpu():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for it in items:
        if "gpu" in it.keywords:
            it.add_marker(skip)
