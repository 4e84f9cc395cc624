"""Host-side culling structures (no GPU): the binned-SAH hierarchy with its threaded subtree build and 16-bit
quantisation (csrc/vr_bvh.cpp) and the per-light direction grids (csrc/vr_lightgrid.cpp).  Both only propose
candidates to the reference's quadric test (gaussian.h:126-164), so the property that keeps results identical to the
reference's is conservativeness; csrc/tests/test_builders.cpp checks it on seeded random boxes."""
from __future__ import annotations

import os
import subprocess

import pytest

from conftest import ROOT

CSRC = os.path.join(ROOT, "3dg-vol-renderer_b200", "csrc")


@pytest.fixture(scope="module")
def builders_exe(tmp_path_factory):
    exe = str(tmp_path_factory.mktemp("builders") / "test_builders")
    subprocess.check_call(["g++", "-O2", "-std=c++17", "-pthread", os.path.join(CSRC, "tests", "test_builders.cpp"),
                           os.path.join(CSRC, "vr_bvh.cpp"), os.path.join(CSRC, "vr_lightgrid.cpp"), "-o", exe])
    return exe


@pytest.mark.parametrize("n,seed", [(1, 5), (2, 6), (300, 2), (5000, 1), (20000, 3)])   # < 4096: sequential build, above: threaded
def test_builders_are_conservative_and_deterministic(builders_exe, n, seed):
    r = subprocess.run([builders_exe, str(n), str(seed)], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout + r.stderr
    for name in ("bvh_deterministic", "bvh_partition_and_containment", "lightgrid_sorted", "lightgrid_conservative"):
        assert f"ok {name}" in r.stdout
