"""Shared fixtures.  GPU tests carry ``@pytest.mark.gpu``; everything else runs on CPU."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN_PATH = os.path.join(ROOT, "tests", "golden", "ntxent_golden.npz")
REDUCTIONS = {0: "none", 1: "mean", 2: "sum"}


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


class GoldenCase:
    def __init__(self, name, npz):
        self.name = name
        n, d, tau, red, bf16 = npz[f"{name}/meta"]
        self.n, self.d, self.tau = int(n), int(d), float(tau)
        self.reduction = REDUCTIONS[int(red)]
        self.bf16 = bool(bf16)
        for key in ("view0", "view1", "loss", "grad_out", "dview0", "dview1"):
            setattr(self, key, npz[f"{name}/{key}"])

    def __repr__(self):
        return f"GoldenCase({self.name})"


def load_golden():
    npz = np.load(GOLDEN_PATH)
    return [GoldenCase(str(name), npz) for name in npz["names"]]


GOLDEN = load_golden()


@pytest.fixture(params=GOLDEN, ids=[c.name for c in GOLDEN])
def golden_case(request):
    return request.param
