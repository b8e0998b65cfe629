import sys
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parents[1]
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def gri():
    from loadbalancedchemistrymodel_b200.foam_io import Mechanism
    return Mechanism.load(ROOT / "tests/golden/gri30.npz")


@pytest.fixture(scope="session")
def ch4():
    from loadbalancedchemistrymodel_b200.foam_io import Mechanism
    return Mechanism.load(ROOT / "tests/golden/ch4_1step.npz")


@pytest.fixture(scope="session")
def flame():
    """The reference's 4000-cell counter-flow flame state (tests/Cases/Case-chemistry/0.2)."""
    z = np.load(ROOT / "tests/golden/case_chemistry_0p2.npz")
    return {k: z[k] for k in z.files}


H2_SPECIES = ["H2", "H", "O", "O2", "OH", "H2O", "HO2", "H2O2", "N2", "AR"]


@pytest.fixture(scope="session")
def h2(gri):
    return gri.subset(H2_SPECIES)
