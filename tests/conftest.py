import importlib
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu on the GPU box)")


def pkg(sub: str = ""):
    """The package directory name starts with a digit, so it is imported through importlib."""
    return importlib.import_module("3dreconstruction_b200" + ("." + sub if sub else ""))


@pytest.fixture(scope="session")
def golden():
    path = os.path.join(ROOT, "tests", "golden", "bf_golden.npz")
    z = np.load(path)
    out = {}
    for name in z["_names"]:
        name = str(name)
        norm, cc, ratio = z[f"{name}/params"]
        out[name] = dict(desc1=z[f"{name}/desc1"], desc2=z[f"{name}/desc2"], norm=int(norm),
                         cross_check=bool(cc), ratio=None if ratio < 0 else float(ratio),
                         q=z[f"{name}/q"], t=z[f"{name}/t"], d=z[f"{name}/d"])
    return out


def sfm_golden_sets():
    """tests/golden/sfm_golden.npz (oracle/make_golden_sfm.py): the reference's matcher + converter on small image sets."""
    z = np.load(os.path.join(ROOT, "tests", "golden", "sfm_golden.npz"))
    out = {}
    for name in z["_sets"]:
        name = str(name)
        norm, cc, ratio = z[f"{name}/params"]
        count = z[f"{name}/count"]
        pairs = {}
        for i in range(len(count)):
            for j in range(i + 1, len(count)):
                k = f"{name}/{i}_{j}"
                pairs[(i, j)] = dict(q=z[k + "/q"], t=z[k + "/t"], s=z[k + "/s"], c=z[k + "/c"])
        out[name] = dict(norm=int(norm), cross_check=bool(cc), ratio=None if ratio < 0 else float(ratio), count=count,
                         desc=z[f"{name}/desc"], xy=z[f"{name}/xy"], pairs=pairs)
    return out


def is_integer_valued(a) -> bool:
    return a.dtype == np.uint8 or bool(np.all(a == np.rint(a)))
