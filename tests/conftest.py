import os
import sys

import numpy as np
import pytest

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(os.path.dirname(__file__), "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def built():
    """build the oracle and the product library once per session"""
    import __graft_entry__ as g
    g.build()
    return True


@pytest.fixture(scope="session")
def kat_rigs(built):
    from multicam_calibration_b200 import synth
    out = {}
    for model in ("radtan", "equidistant"):
        poses = np.load(os.path.join(GOLDEN, f"{model}_init_poses.npy"))
        out[model] = synth.kat_rig(model, poses)
    return out


@pytest.fixture(scope="session")
def small_rig(built):
    """3 mixed-model cameras, 24 frames, noisy: the workhorse of the unit-level parity tests"""
    from multicam_calibration_b200 import synth
    return synth.make_synthetic_rig(3, 24, 7, detect_prob=0.8)


@pytest.fixture(scope="session")
def c3_rig(built):
    from multicam_calibration_b200 import synth
    return synth.config_c3()
