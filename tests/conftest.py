import os
import sys
import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def load_golden(name):
    return np.load(os.path.join(GOLDEN, name + ".npz"))


GOLDEN_CASES = ["g_tiny", "g_small", "g_loops", "g_outliers", "g_odom_only", "g_worse", "g_worse_lm", "g_converged", "c1"]
# exit codes the fixtures end with (OptimizerCpu.h:140-179): 1 "Error is getting worse", 2 "Plateau", 3 "CONVERGED"
GOLDEN_EXIT_CODES = {"g_worse": 1, "g_worse_lm": 1, "c1": 2, "g_converged": 3}


def wrap(a):
    return np.arctan2(np.sin(a), np.cos(a))


def state_diff(pose_a, lm_a, pose_b, lm_b):
    """max abs difference of two states, theta compared modulo 2*pi, relative to max(1, |ref|)."""
    d = pose_a - pose_b
    d[:, 2] = wrap(d[:, 2])
    scale_p = np.maximum(1.0, np.abs(pose_b))
    out = float(np.max(np.abs(d) / scale_p)) if len(d) else 0.0
    if len(lm_a):
        out = max(out, float(np.max(np.abs(lm_a - lm_b) / np.maximum(1.0, np.abs(lm_b)))))
    return out
