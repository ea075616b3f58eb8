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


@pytest.fixture(scope="session")
def engine_setup():
    """The CUDA engine with its default (shipping) configuration.  Fails loudly without the built .so or a GPU."""
    import kateth_b200 as K
    wb = int(os.environ.get("KB200_WINDOW_BITS", "0"))
    s = K.Setup.load_json(SETUP_JSON, device=0, window_bits=wb)
    yield s
    s.close()
