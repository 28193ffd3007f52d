import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA (B200) device")


@pytest.fixture(scope="session")
def golden_dir():
    return GOLDEN


_METRICS = os.path.join(ROOT, "gpurun_out", "test_metrics.jsonl")


@pytest.fixture(scope="session")
def record_metric():
    """record_metric(test, name, value): achieved parity errors of the GPU tests, appended to
    gpurun_out/test_metrics.jsonl when that directory can be written (so a tolerance is stated against what the
    kernels actually reach; profiles/ keeps the summary)."""
    import json

    def rec(test, name, value, tol=None):
        try:
            os.makedirs(os.path.dirname(_METRICS), exist_ok=True)
            with open(_METRICS, "a") as f:
                f.write(json.dumps({"test": test, "name": name, "value": float(value), "tol": tol}) + "\n")
        except OSError:
            pass

    return rec
