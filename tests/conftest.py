import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "formula-vad_b200"))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
sys.path.insert(0, ROOT)

SEED = 0xF1AD5EED00000000


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def _have_gpu():
    try:
        import fvad
        return fvad.lib().fvad_device_count() > 0
    except Exception:
        return False


def pytest_collection_modifyitems(config, items):
    # `-m gpu` on a box without a GPU must fail loudly, not skip: there is no CPU fallback to test.
    pass


@pytest.fixture(scope="session")
def fv():
    import fvad
    fvad.lib()
    return fvad


@pytest.fixture(scope="session")
def port():
    from oracle import Oracle
    return Oracle("port")


@pytest.fixture(scope="session")
def ref():
    from oracle import Oracle, have_reference, build
    if not have_reference():
        try:
            build("reference")
        except Exception:
            pass
    if not have_reference():
        pytest.skip("oracle/_ref/liboracle_ref.so not built (needs /root/reference at build time)")
    return Oracle("reference")


@pytest.fixture(scope="session")
def checker(request):
    """The strongest oracle available: the reference build if present, else the port."""
    from oracle import Oracle, have_reference
    return Oracle("reference") if have_reference() else Oracle("port")


def synth(fvmod, stream_id, seconds, first=0, seed=SEED):
    return fvmod.synth_host(seed, stream_id, first, int(seconds * 48000))


def assert_segments_equal(got, want):
    assert len(got) == len(want), (len(got), len(want), got, want)
    assert np.array_equal(got["sample_from"], want["sample_from"])
    assert np.array_equal(got["sample_to"], want["sample_to"])


@pytest.fixture
def knobs(fv):
    """Set library knobs (fvad_debug_set) for one test; cleared afterwards."""
    def set_(**kw):
        for k, v in kw.items():
            fv.debug_set(k, v)
    yield set_
    fv.debug_set()
