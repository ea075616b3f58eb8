import os, sys
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
SETUP_JSON = os.path.join(ROOT, "tests", "golden", "trusted_setup_4096.json")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def oracle_setup():
    from oracle import oracle as O
    return O.Setup.load_json(SETUP_JSON)


# ---- the CUDA engine under test -----------------------------------------------------------------------------------------
# Every engine test runs in BOTH table forms that ship: plain c = 13 (28.6 GiB; what window_bits 0 picks when HBM is short) and the one
# bench.py measures and window_bits 0 picks on an idle B200 (16 -> GLV, 8 shared windows, 100 GiB table).  One 100 GiB table and a second big one do not
# fit in HBM together, so at most one engine is alive at a time: `engine(wb)` closes whatever else is loaded, and tests that load
# their own setups take the `gpu_free` fixture first.
_ENGINES = {}
ENGINE_CONFIGS = [13, 16]


def release_engines():
    for k in list(_ENGINES):
        _ENGINES.pop(k).close()


def engine(wb):
    import kateth_b200 as K
    if wb not in _ENGINES:
        release_engines()
        _ENGINES[wb] = K.Setup.load_json(SETUP_JSON, device=0, window_bits=wb)
    return _ENGINES[wb]


@pytest.fixture(scope="session", params=ENGINE_CONFIGS, ids=lambda wb: "c13" if wb == 13 else "glv%d" % wb)
def engine_setup(request):
    """The CUDA engine (fails loudly without the built .so or a GPU).  KB200_WINDOW_BITS pins one configuration."""
    wb = request.param
    if "KB200_WINDOW_BITS" in os.environ and int(os.environ["KB200_WINDOW_BITS"]) != wb:
        pytest.skip("KB200_WINDOW_BITS pins another configuration")
    yield _Proxy(wb)


class _Proxy:
    """Resolves to the live engine of one configuration on every use (another test may have released it in between)."""

    def __init__(self, wb):
        self._wb = wb

    def __getattr__(self, name):
        return getattr(engine(self._wb), name)


@pytest.fixture
def gpu_free():
    """For tests that load their own setups: nothing else may hold HBM."""
    release_engines()
    yield
    release_engines()


def pytest_sessionfinish(session, exitstatus):
    release_engines()


@pytest.fixture
def any_engine():
    """Whichever engine is alive (after the other tests: the last configuration they used), else the small table -- for tests that do not
    depend on the table form and must not be reordered by a parametrised session fixture (tests/test_zz_no_alloc.py runs last)."""
    if _ENGINES:
        return next(iter(_ENGINES.values()))
    return engine(int(os.environ.get("KB200_WINDOW_BITS", "13")))
