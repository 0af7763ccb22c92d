import json
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "vllm-production-stack_b200")
for p in (ROOT, PKG):
    if p not in sys.path:
        sys.path.insert(0, p)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def pytest_sessionstart(session):
    """Build the libraries when they are missing or older than their sources (nvcc / g++
    cross-compile without a GPU); a failure here is reported by the tests that need them."""
    try:
        import build as b200_build
        if b200_build.needs_build():
            b200_build.build()
    except Exception as e:                      # pragma: no cover
        print("conftest: could not build the libraries: %r" % (e,))


@pytest.fixture(autouse=True, scope="session")
def _hash_mode_from_env():
    """B200R_TEST_HASH_MODE=0|1 runs the whole GPU suite with that chunk-hashing kernel."""
    mode = os.environ.get("B200R_TEST_HASH_MODE")
    if mode is not None:
        from b200_router import _lib
        _lib.check(_lib.lib().b200r_set_hash_mode(int(mode)))
    yield


def pytest_collection_modifyitems(config, items):
    # Without a device the gpu-marked tests cannot run; the driver selects them
    # with -m gpu on the B200 box and with -m "not gpu" here.
    try:
        import torch
        have = torch.cuda.is_available()
    except Exception:
        have = False
    if have:
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for it in items:
        if "gpu" in it.keywords:
            it.add_marker(skip)


def load_golden(name):
    with open(os.path.join(GOLDEN, name)) as f:
        return json.load(f)


@pytest.fixture(scope="session")
def golden_prefix():
    return load_golden("prefix_traces.json")


@pytest.fixture(scope="session")
def golden_xxh():
    return load_golden("xxh64_kat.json")


@pytest.fixture(scope="session")
def golden_md5():
    return load_golden("md5_ring_kat.json")


@pytest.fixture(scope="session")
def golden_go():
    return load_golden("go_picker_kat.json")
