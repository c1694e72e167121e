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
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def golden_names():
    return sorted(os.path.basename(p)[:-4] for p in glob.glob(os.path.join(GOLDEN_DIR, "*.npz")))


def load_golden(name):
    z = np.load(os.path.join(GOLDEN_DIR, name + ".npz"))
    g = {k: z[k] for k in z.files}
    for k in ("suf", "lcp", "llv", "bck"):
        g[k] = g[k].tobytes()
    g["alphabet"] = str(g["alphabet"])
    g["averagelcp"] = str(g["averagelcp"])
    for k in ("prefixlength", "longest", "maxbranchdepth", "largelcpvalues", "specialcharacters", "totallength"):
        g[k] = int(g[k])
    g["sigma"] = int(g["sigma"]) if "sigma" in g else (4 if g["alphabet"] == "dna" else 20)
    return g


@pytest.fixture(params=golden_names())
def golden(request):
    g = load_golden(request.param)
    g["name"] = request.param
    return g
