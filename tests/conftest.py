import os
import sys
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def oracle():
    from oracle.oracle import Oracle
    return Oracle()


@pytest.fixture(scope="session")
def golden_conv():
    import numpy as np
    return np.load(os.path.join(ROOT, "tests", "golden", "conv_cases.npz"))


@pytest.fixture(scope="session")
def golden_ew():
    import numpy as np
    return np.load(os.path.join(ROOT, "tests", "golden", "elementwise_cases.npz"))


@pytest.fixture(scope="session")
def golden_mm():
    import numpy as np
    return np.load(os.path.join(ROOT, "tests", "golden", "matmul_cases.npz"))


@pytest.fixture(scope="session")
def golden_split():
    import numpy as np
    return np.load(os.path.join(ROOT, "tests", "golden", "split_cases.npz"))


@pytest.fixture(scope="session")
def golden_rgb():
    import numpy as np
    return np.load(os.path.join(ROOT, "tests", "golden", "conv_rgb_cases.npz"))
