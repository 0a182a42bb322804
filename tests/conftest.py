import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200); run with -m gpu on the GPU box")
    config.addinivalue_line("markers", "slow: full-size BASELINE configs (minutes on one B200); deselect with -m 'gpu and not slow'")


@pytest.fixture(scope="session")
def golden():
    import json
    with open(os.path.join(ROOT, "tests", "golden", "addresses.json")) as f:
        return json.load(f)


@pytest.fixture(scope="session")
def regex_golden():
    import json
    with open(os.path.join(ROOT, "tests", "golden", "regex_cases.json")) as f:
        return json.load(f)


@pytest.fixture(scope="session")
def published():
    """Vectors copied from published specifications (tests/golden/published.json; sources cited inside)."""
    import json
    with open(os.path.join(ROOT, "tests", "golden", "published.json")) as f:
        return json.load(f)
