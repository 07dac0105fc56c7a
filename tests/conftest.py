"""Shared fixtures. `-m "not gpu"` runs here (no GPU); `-m gpu` runs on the B200 box.

The oracle (oracle/) is the CHECKER; the product is libmpccbf.so reached through its C ABI.
"""
import importlib.util
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu on the GPU box)")


def load_product():
    name = "mpc_cbf_b200"
    if name in sys.modules:
        return sys.modules[name]
    spec = importlib.util.spec_from_file_location(name, os.path.join(ROOT, "mpc-cbf_b200", "__init__.py"))
    mod = importlib.util.module_from_spec(spec)
    sys.modules[name] = mod
    spec.loader.exec_module(mod)
    return mod


@pytest.fixture(scope="session")
def M():
    mod = load_product()
    if not os.path.exists(mod.LIB_PATH):
        mod.build()
    return mod


@pytest.fixture(scope="session")
def O():
    import pyoracle
    pyoracle.build()
    return pyoracle


def swarm(S, N, seed=20261104, rmin=3.0, rmax=6.0, vmax=0.5):
    """cfg3-style synthetic swarm: agents on a circle, antipodal goals."""
    rng = np.random.default_rng(seed)
    R = rng.uniform(rmin, rmax, size=(S, 1))
    phi = rng.uniform(0, 2 * np.pi, size=(S, 1))
    ang = phi + 2 * np.pi * np.arange(N)[None, :] / N
    states = np.zeros((S, N, 6))
    states[:, :, 0] = R * np.cos(ang)
    states[:, :, 1] = R * np.sin(ang)
    states[:, :, 3:5] = rng.uniform(-vmax, vmax, size=(S, N, 2))
    targets = np.zeros((S, N, 3))
    targets[:, :, 0] = -states[:, :, 0]
    targets[:, :, 1] = -states[:, :, 1]
    return states, targets


def north_star_cfg(mod, **kw):
    """K=20, 4 pieces, d_min 0.8."""
    base = dict(k_hor=20, num_pieces=4, d_min=0.8, num_agents=16, max_scenarios=64)
    base.update(kw)
    return mod.default_config(**base)
