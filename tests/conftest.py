import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run with -m gpu on the B200 box)")


class Golden:
    """One tests/golden/*.npz fixture: parameter tables + inputs + reference outputs."""

    def __init__(self, name):
        from molecular_dynamics_b200 import _cabi
        from molecular_dynamics_b200.parameters import Parameters

        with np.load(os.path.join(GOLDEN, name + ".npz")) as z:
            self.z = {k: z[k] for k in z.files}
        self.par = Parameters.from_arrays({k[4:]: v for k, v in self.z.items() if k.startswith("par_")})
        self.ava_idx = self.z["ava_idx"]
        self.packed = _cabi.PackedTopology(self.par, self.ava_idx)

    def __getitem__(self, k):
        return self.z[k]

    def __contains__(self, k):
        return k in self.z


_cache = {}


def load_golden(name):
    if name not in _cache:
        _cache[name] = Golden(name)
    return _cache[name]


@pytest.fixture(scope="session")
def golden():
    return load_golden


MODES = {"plain": dict(), "rf": dict(cutoff=9, rfa=True, switch_dist=7.5)}
SYSTEMS = ("bb40", "bb40_improper", "ala10", "diala", "ala10_ff14sb")
