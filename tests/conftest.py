import glob
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu on the GPU box)")


def golden_event_cases():
    return sorted(os.path.basename(p)[len("events_"):-len(".npz")] for p in glob.glob(os.path.join(GOLDEN, "events_*.npz")))


def load_golden_events(name):
    z = np.load(os.path.join(GOLDEN, f"events_{name}.npz"))
    ev = {k: z[k] for k in ("x", "y", "t", "p")}
    hw = z["frame_hw"]
    frames, off = [], 0
    for h, w in hw:
        frames.append(z["frames"][off:off + h * w * 3].reshape(h, w, 3))
        off += h * w * 3
    return ev, frames, z["pix"]


@pytest.fixture(scope="session")
def golden_dir():
    return GOLDEN
