import os, sys
import pytest

REPO = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
PKG = os.path.join(REPO, "vistaf-roboskin-vision-integrated-multimodal-sensor_b200")
for p in (PKG, os.path.join(REPO, "oracle"), REPO):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def golden():
    import json
    import numpy as np
    g = os.path.join(REPO, "tests", "golden")
    return json.load(open(os.path.join(g, "reference_outputs.json"))), np.load(os.path.join(g, "reference_outputs.npz"))
