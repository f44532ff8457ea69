import os, sys
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, 'tests', 'golden')


def pytest_configure(config):
    config.addinivalue_line('markers', 'gpu: needs a CUDA device (run on the B200 box with -m gpu)')


@pytest.fixture(scope='session')
def golden():
    import numpy as np
    return {name[:-4]: np.load(os.path.join(GOLDEN, name)) for name in os.listdir(GOLDEN) if name.endswith('.npz')}


# ---- parity report: tests append measured deviations / sign-mismatch counts per row; written where gpurun brings it back
PARITY_REPORT = {}


def pytest_sessionfinish(session, exitstatus):
    out_dir = os.path.join(ROOT, 'gpurun_out')
    if PARITY_REPORT and os.path.isdir(out_dir):
        import json
        with open(os.path.join(out_dir, 'parity_report.json'), 'w') as f:
            json.dump(PARITY_REPORT, f, indent=1)
