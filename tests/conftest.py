import glob
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN_DIR = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu on the GPU box)")


def golden_names():
    return sorted(os.path.splitext(os.path.basename(p))[0] for p in glob.glob(os.path.join(GOLDEN_DIR, "*.npz")))


class Golden(object):
    """One reference-generated fixture (tests/golden/make_golden.py)."""

    def __init__(self, name):
        self.name = name
        self.z = np.load(os.path.join(GOLDEN_DIR, name + ".npz"))
        self.no = int(self.z["no"])
        self.nv = int(self.z["nv"])
        self.F = self.z["F"]
        self.blocks = {k[4:]: self.z[k] for k in self.z.files if k.startswith("blk_")}

    def __getitem__(self, k):
        return self.z[k]


@pytest.fixture(params=golden_names())
def golden(request):
    return Golden(request.param)
