"""pytest configuration: the `gpu` marker and shared helpers.

`-m "not gpu"` runs in the build container (no GPU): oracle vs golden fixtures, host
logic, C-ABI symbol checks.  `-m gpu` runs on a B200: CUDA path vs oracle, through
the C ABI.  Nothing here reads /root/reference.
"""
import sys
from pathlib import Path

import numpy as np
import pytest
import torch

ROOT = Path(__file__).resolve().parent.parent
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))

GOLDEN = ROOT / "tests" / "golden"


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def pytest_collection_modifyitems(config, items):
    if torch.cuda.is_available():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


@pytest.fixture(scope="session")
def golden():
    def load(name):
        return np.load(GOLDEN / f"{name}.npz")
    return load


def load_oracle_model(d, prefix="w."):
    """OracleModel with the weights stored in a golden fixture."""
    from oracle import reference as orc
    V, E, P, J, H, L = (int(x) for x in d["dims"])
    m = orc.OracleModel(V, E, P, J, E, H, L, blank=0)
    sd = {k[len(prefix):]: torch.from_numpy(d[k]) for k in d.files if k.startswith(prefix)}
    missing = m.load_state_dict(sd, strict=False)
    assert not missing.unexpected_keys, missing
    m.eval()
    return m
