import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")
    # build the test oracles (and, where /root/reference exists, the reference shim) if absent
    oracle_dir = os.path.join(ROOT, "oracle")
    if not os.path.exists(os.path.join(oracle_dir, "_ref", "libansoracle.so")):
        subprocess.check_call(["make", "-C", oracle_dir])


def _have_gpu():
    try:
        import torch

        return torch.cuda.is_available()
    except Exception:
        return False


def pytest_collection_modifyitems(config, items):
    if _have_gpu():
        return
    skip = pytest.mark.skip(reason="no CUDA device here")
    for it in items:
        if "gpu" in it.keywords:
            it.add_marker(skip)


@pytest.fixture(scope="session")
def orc():
    from _libs import oracle

    return oracle()


@pytest.fixture(scope="session")
def ref():
    from _libs import have_reference, reference

    if not have_reference():
        pytest.skip("oracle/_ref/libansref.so not built (needs /root/reference)")
    return reference()


@pytest.fixture(scope="session")
def host():
    import ans_large_alphabet_b200 as P

    return P.host_api()


@pytest.fixture(scope="session")
def _session_codec():
    import ans_large_alphabet_b200 as P

    c = P.Codec(device=0)
    yield c
    c.close()


@pytest.fixture
def codec(_session_codec):
    """one context per session; every test starts from the default kernel path and block size"""
    _session_codec.set_kernel_path(0)
    _session_codec.set_block_syms(8192)
    yield _session_codec
    _session_codec.set_kernel_path(0)
