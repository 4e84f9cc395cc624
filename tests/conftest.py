"""pytest configuration: markers, import path, shared fixtures."""
from __future__ import annotations

import importlib.util
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLD = os.path.join(ROOT, "tests", "golden")
SCENES = os.path.join(GOLD, "scenes")
PKG_DIR = os.path.join(ROOT, "3dg-vol-renderer_b200")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def load_pkg_module(name: str):
    """Import 3dg-vol-renderer_b200/<name>.py (the directory name is not a Python identifier)."""
    key = "volrender_b200_" + name
    if key in sys.modules:
        return sys.modules[key]
    spec = importlib.util.spec_from_file_location(key, os.path.join(PKG_DIR, name + ".py"))
    mod = importlib.util.module_from_spec(spec)
    sys.modules[key] = mod
    spec.loader.exec_module(mod)
    return mod


@pytest.fixture(scope="session")
def golden():
    return np.load(os.path.join(GOLD, "golden_v1.npz"))


@pytest.fixture(scope="session")
def golden_march():
    """reference images of the validation integrators (oracle/make_golden_march.py)"""
    return np.load(os.path.join(GOLD, "golden_march_v1.npz"))


@pytest.fixture(scope="session")
def golden_r2():
    """per-sample radiance, black-environment sphere images, far / scaled scenes from the reference (oracle/make_golden_r2.py)"""
    return np.load(os.path.join(GOLD, "golden_r2.npz"))


@pytest.fixture(scope="session")
def scenegen():
    return load_pkg_module("scenegen")


@pytest.fixture(scope="session")
def ob():
    from oracle import oracle_binding
    oracle_binding.build()
    return oracle_binding


CAMS = {
    "pin": dict(kind="pinhole", pos=(0, 1, 6), look=(0, 1, 0), fov=0.78539816339),
    "pin2": dict(kind="pinhole", pos=(2.5, 1.8, 3.5), look=(0, 0.9, 0), fov=0.6),
    "ortho": dict(kind="ortho", pos=(0, 1, 6), look=(0, 1, 0), fov=0.78539816339),
    # the "pin" camera moved / scaled with the far and x50 variants of rand300 (oracle/make_golden_r2.py)
    "pin_far": dict(kind="pinhole", pos=(100, -49, 206), look=(100, -49, 200), fov=0.78539816339),
    "pin_x50": dict(kind="pinhole", pos=(0, 50, 300), look=(0, 50, 0), fov=0.78539816339),
}
ENV = (0.53, 0.81, 0.92)


def decode_record(raw: np.ndarray):
    """golden '<render>/record' -> list of per-pixel index arrays (ref_driver layout)."""
    npix = int(raw[0])
    out, off = [], 1
    for _ in range(npix):
        c = int(raw[off]); off += 1
        out.append(raw[off:off + c].astype(np.uint32)); off += c
    return out


_BIG_SCENES = {}


def big_scene(scenegen, name):
    """the seeded benchmark mixtures of SURVEY.md 8(d) (generation is a Python loop: 200k components take ~20 s, once per session)"""
    if name not in _BIG_SCENES:
        _BIG_SCENES[name] = {"c1": lambda: scenegen.generate(1000, 1),
                             "c2": lambda: scenegen.generate(10000, 2, nonuniform=True),
                             "c3": lambda: scenegen.generate(200000, 3, albedo_range=(0.9, 0.999))}[name]()
    return _BIG_SCENES[name]
