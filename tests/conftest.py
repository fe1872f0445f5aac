import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200); run with -m gpu")


@pytest.fixture(scope="session")
def golden():
    import numpy as np
    return np.load(os.path.join(ROOT, "tests", "golden", "reference_vectors.npz"))


@pytest.fixture(scope="session")
def lib_built():
    from tmdiff_b200 import build
    return build.build()


@pytest.fixture(scope="session")
def golden_samplers():
    """Reference runs of the multistep / noise-type / classifier-free-guidance samplers (tests/golden/make_golden_samplers.py)."""
    import numpy as np
    return np.load(os.path.join(ROOT, "tests", "golden", "sampler_variants.npz"))


# tag: (bands, HW, sensor, steps, order, model_type, guidance, denoise_to_zero) -- as in make_golden_samplers.py
SAMPLER_CASES = {
    "ms3_noise_12": (4, 32, "QB", 12, 3, "noise", None, True),
    "ms2_cfg_8": (4, 32, "GF2", 8, 2, "noise", 2.0, False),
    "ms3_xstart_9": (8, 32, "WV3", 9, 3, "x_start", None, True),
}
