import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def golden_st():
    """student_t_lpdf vectors from the compiled reference (tests/golden/make_student_t_golden.py)."""
    import numpy as np
    return np.load(os.path.join(ROOT, "tests", "golden", "student_t_golden.npz"), allow_pickle=False)


@pytest.fixture(scope="session")
def golden_d2():
    """poisson_lpmf / bernoulli_lpmf vectors from the compiled reference (tests/golden/make_density2_golden.py)."""
    import numpy as np
    return np.load(os.path.join(ROOT, "tests", "golden", "density2_golden.npz"), allow_pickle=False)


@pytest.fixture(scope="session")
def golden():
    import numpy as np
    path = os.path.join(ROOT, "tests", "golden", "glm_golden.npz")
    return np.load(path, allow_pickle=False)
