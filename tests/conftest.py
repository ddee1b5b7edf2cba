import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def ccjson_text():
    return open(os.path.join(ROOT, "tests", "golden", "CC.json")).read()


@pytest.fixture(scope="session")
def oracle_c1(ccjson_text):
    import oracle
    return oracle.Oracle.from_ccjson(ccjson_text)
