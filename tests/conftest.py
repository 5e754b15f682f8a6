import json, os, sys
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def golden():
    return json.load(open(os.path.join(ROOT, "tests", "golden", "golden.json")))


@pytest.fixture(scope="session")
def gz():
    import gzippier_b200
    gzippier_b200.lib()
    return gzippier_b200


@pytest.fixture(scope="session")
def orc():
    import oracle_lib
    oracle_lib.lib()
    return oracle_lib


@pytest.fixture(scope="session")
def ctx(gz):
    try:
        c = gz.Context(0)      # raises GzbError(E_NO_DEVICE) without a B200: there is no CPU fallback
    except gz.GzbError as e:
        if e.code == gz.E_NO_DEVICE:
            pytest.skip("no B200 here (the engine has no CPU fallback); run with -m gpu on the GPU box")
        raise
    yield c
    c.close()


# boundary sizes: reference tests/mixed:51-57 (0..64, 32831-32833, 131071-131073) plus the chunk/tile/window edges
EDGE_SIZES = [0, 1, 2, 3, 4, 5, 31, 32, 33, 63, 64, 255, 258, 259, 2047, 2048, 2049, 4095, 4096, 4097, 16383, 16384, 16385,
              32767, 32768, 32769, 32831, 32832, 32833, 65535, 65536, 65537, 131071, 131072, 131073, 163840, 262143, 262144, 262145, 393217]
