import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def golden():
    import numpy as np

    d = os.path.join(ROOT, "tests", "golden")
    # two weight sets, both outputs of the REFERENCE (oracle/pin_against_reference.py): init_scale 1.0 draws the
    # reference's init_scale=0 tensors at full fan-avg scale (worst case for error growth: every residual branch
    # is as loud as the skip path), 0.01 is the trained-like set (branches quiet, as in a trained score net)
    return {"ref": np.load(os.path.join(d, "reference_outputs.npz")),
            "c1": np.load(os.path.join(d, "sampler_c1.npz")),
            "ref_s0p01": np.load(os.path.join(d, "reference_outputs_s0p01.npz")),
            "c1_s0p01": np.load(os.path.join(d, "sampler_c1_s0p01.npz")),
            "variants": np.load(os.path.join(d, "variants.npz")),
            "sb": np.load(os.path.join(d, "sb.npz")),
            "loss": np.load(os.path.join(d, "loss.npz"))}
