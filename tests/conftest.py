import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def load_hand_vectors():
    import json
    with open(os.path.join(ROOT, "tests", "golden", "hand_vectors.json")) as f:
        cases = json.load(f)["cases"]
    for c in cases:
        if "transactions_rle" in c:
            c["transactions"] = [list(tx) for n, tx in c["transactions_rle"] for _ in range(n)]
    return cases


@pytest.fixture(scope="session")
def hand_vectors():
    return load_hand_vectors()
