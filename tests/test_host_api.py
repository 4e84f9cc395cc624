"""The C++ drop-in API (3dg-vol-renderer_b200/host): Scene / Camera / Image / Integrator /
InverseIntegrator classes with the reference's names and signatures, over the C-ABI.

CPU part: host logic (loaders, cameras, image I/O, parameter packing, error behaviour) through
host/bin/test_host.  GPU part: host/bin/vol_render -- the replacement of the reference's
tests/main.cpp -- renders the golden scenes and must reproduce the REFERENCE's float images and
recorded Gaussian lists (same bars as tests/test_gpu_parity.py, which goes through ctypes)."""
from __future__ import annotations

import os
import subprocess

import numpy as np
import pytest

from conftest import CAMS, ROOT, SCENES, decode_record

HOST = os.path.join(ROOT, "3dg-vol-renderer_b200", "host")
BIN = os.path.join(HOST, "bin")


@pytest.fixture(scope="module")
def host_bins():
    from conftest import load_pkg_module
    vr = load_pkg_module("volrender")
    if not os.path.exists(vr.LIB_PATH):
        vr.build()
    subprocess.check_call(["make", "-C", HOST, "-j4"], stdout=subprocess.DEVNULL)
    return BIN


def run(args, **kw):
    return subprocess.run(args, capture_output=True, text=True, timeout=600, **kw)


def test_host_logic(host_bins, tmp_path):
    r = run([os.path.join(host_bins, "test_host"), SCENES, str(tmp_path)])
    assert r.returncode == 0, r.stdout + r.stderr
    for name in ("scene_gmm_text", "scene_spheres_xml", "cameras", "image", "pack_parameters", "errors"):
        assert f"ok {name}" in r.stdout


def test_cli_fails_loudly_without_gpu_or_scene(host_bins, tmp_path):
    r = run([os.path.join(host_bins, "vol_render"), "--scene", str(tmp_path / "missing.txt")])
    assert r.returncode == 1 and "Failed to open scene file" in r.stderr
    import torch
    if not torch.cuda.is_available():
        r = run([os.path.join(host_bins, "vol_render"), "--scene", os.path.join(SCENES, "rand50.txt"), "--spp", "4",
                 "--width", "8", "--height", "8"], cwd=str(tmp_path))
        assert r.returncode == 1 and "no CPU fallback" in r.stderr


def _cam_args(name):
    c = CAMS[name]
    return ["--camera", "pinhole" if c["kind"] == "pinhole" else "ortho", "--pos", ",".join(map(str, c["pos"])),
            "--look", ",".join(map(str, c["look"])), "--fov", repr(c["fov"])]


def frac_within(a, b, rtol, atol=1e-6):
    return float((np.abs(a - b) <= atol + rtol * np.abs(b)).all(axis=-1).mean())


@pytest.mark.gpu
@pytest.mark.parametrize("sc,cam,spp,integ", [("two_gauss", "pin", 16, "ms"), ("two_gauss", "ortho", 4, "ss"),
                                              ("rand50", "pin2", 9, "ms"), ("rand300", "pin", 16, "ss"),
                                              ("rand2k", "pin", 16, "ms")])
def test_vol_render_matches_reference(host_bins, golden, tmp_path, sc, cam, spp, integ):
    """>= 98% of pixels within 1e-4 relative of the reference's float image (north_star tolerance)."""
    raw = tmp_path / "img.f32"
    r = run([os.path.join(host_bins, "vol_render"), "--scene", os.path.join(SCENES, sc + ".txt"), "--integrator", integ,
             "--spp", str(spp), "--width", "32", "--height", "32", "--raw", str(raw), "--out", str(tmp_path / "o.ppm")]
            + _cam_args(cam))
    assert r.returncode == 0, r.stdout + r.stderr
    img = np.fromfile(raw, np.float32).reshape(32, 32, 3)
    ref = golden[f"render/{sc}/{cam}/{integ}/{spp}"]
    assert frac_within(img, ref, 1e-4) >= 0.98
    ppm = open(tmp_path / "o.ppm", "rb").read()
    assert ppm.startswith(b"P6\n32 32\n255\n") and len(ppm) == 13 + 32 * 32 * 3
    body = np.frombuffer(ppm[13:], np.uint8).reshape(32, 32, 3)
    assert np.array_equal(body, np.clip(img * np.float32(255.0), 0, 255).astype(np.uint8))   # image.h:65-67


@pytest.mark.gpu
def test_vol_render_recorded_lists(host_bins, golden, tmp_path):
    rec = tmp_path / "rec.u32"
    r = run([os.path.join(host_bins, "vol_render"), "--scene", os.path.join(SCENES, "rand50.txt"), "--spp", "16",
             "--width", "32", "--height", "32", "--record", str(rec), "--out", str(tmp_path / "o.ppm")] + _cam_args("pin"))
    assert r.returncode == 0, r.stdout + r.stderr
    mine = decode_record(np.fromfile(rec, np.uint32))
    ref = decode_record(golden["render/rand50/pin/ms/16/record"])
    same = sum(set(a.tolist()) == set(b.tolist()) for a, b in zip(mine, ref))
    assert len(mine) == len(ref) and same / len(ref) >= 0.98


@pytest.mark.gpu
def test_vol_render_two_devices_equals_one(host_bins, tmp_path):
    """VR_DEVICES="0,0": two contexts, interleaved rows, host threads -- the image must be bit-identical
    to the single-context one (samples depend only on (x, y, si))."""
    outs = []
    for devs in ("0", "0,0"):
        raw = tmp_path / f"img{len(outs)}.f32"
        r = run([os.path.join(host_bins, "vol_render"), "--scene", os.path.join(SCENES, "rand300.txt"), "--spp", "16",
                 "--width", "40", "--height", "33", "--raw", str(raw), "--out", str(tmp_path / "o.ppm")],
                env=dict(os.environ, VR_DEVICES=devs))
        assert r.returncode == 0, r.stdout + r.stderr
        outs.append(np.fromfile(raw, np.float32))
    assert np.array_equal(outs[0], outs[1])


@pytest.mark.gpu
def test_vol_render_spheres_xml(host_bins, golden, tmp_path):
    """config C0 through the XML reader: one sphere, orthographic camera, RayMarchingSpheres."""
    xml = tmp_path / "one.xml"
    xml.write_text("""<scene version="3.0.0">
  <sensor type="orthographic"><transform name="to_world"><lookat origin="0,1,6" target="0,1,0" up="0,1,0"/></transform>
    <film type="hdrfilm"><integer name="width" value="48"/><integer name="height" value="48"/></film></sensor>
  <emitter type="constant"><rgb name="radiance" value="0.53, 0.81, 0.92"/></emitter>
  <emitter type="point"><point name="position" x="0" y="4" z="0"/><rgb name="intensity" value="35.0"/></emitter>
  <medium id="vol" type="homogeneous"><rgb name="albedo" value="0.875"/><rgb name="sigma_t" value="0.8"/></medium>
  <shape type="sphere"><point name="center" x="0" y="1" z="0"/><float name="radius" value="1"/><ref name="interior" id="vol"/></shape>
</scene>""")
    raw = tmp_path / "img.f32"
    r = run([os.path.join(host_bins, "vol_render"), "--scene", str(xml), "--raw", str(raw), "--out", str(tmp_path / "o.ppm")])
    assert r.returncode == 0, r.stdout + r.stderr
    img = np.fromfile(raw, np.float32).reshape(48, 48, 3)
    gold = golden["spheres/one_sphere/ortho"]
    bg = (gold == np.array((0.53, 0.81, 0.92), np.float32)).all(axis=-1)
    assert np.array_equal(img[bg], gold[bg])
    assert abs(float(img.mean()) - float(gold.mean())) < 0.01 * float(gold.mean())


@pytest.mark.gpu
def test_vol_render_inverse_reduces_loss(host_bins, tmp_path):
    """StochasticFiniteDiffInverseIntegrator::optimize end to end: target = render of the scene, start =
    same scene with densities scaled; a few iterations must lower the mean L1 loss and write the PPMs."""
    exe = os.path.join(host_bins, "vol_render")
    cam = _cam_args("pin")
    tgt = tmp_path / "target.ppm"
    r = run([exe, "--scene", os.path.join(SCENES, "rand50.txt"), "--spp", "64", "--width", "32", "--height", "32", "--out", str(tgt)] + cam)
    assert r.returncode == 0, r.stderr
    lines = open(os.path.join(SCENES, "rand50.txt")).read().split("\n")
    pert = []
    for ln in lines:
        t = ln.split()
        if t and t[0] == "g":
            t[10] = repr(float(t[10]) * 1.6)
            ln = " ".join(t)
        pert.append(ln)
    start = tmp_path / "start.txt"
    start.write_text("\n".join(pert))
    out_dir = tmp_path / "sfd"
    r = run([exe, "--scene", str(start), "--inverse", str(tgt), "--spp", "16", "--iters", "31", "--save-every", "10", "--final-spp", "64",
             "--out-dir", str(out_dir), "--save-scene", str(tmp_path / "fit.txt")] + cam)
    assert r.returncode == 0, r.stdout + r.stderr
    losses = [float(l.split(":")[1]) for l in r.stdout.splitlines() if "mean loss" in l.lower()]
    assert len(losses) >= 4 and losses[-1] < losses[0], losses
    assert (out_dir / "iter_0000.ppm").exists() and (out_dir / "iter_0030.ppm").exists()
    assert sum(1 for l in open(tmp_path / "fit.txt") if l.startswith("g ")) == 50


def _inverse_params(exe, tmp_path, tag, extra=(), env=None, iters=5):
    cam = _cam_args("pin")
    tgt = tmp_path / "target_common.ppm"
    scene = os.path.join(SCENES, "rand50.txt")
    if not tgt.exists():
        assert run([exe, "--scene", scene, "--spp", "16", "--width", "24", "--height", "24", "--out", str(tgt)] + cam).returncode == 0
    out = tmp_path / f"{tag}.f32"
    r = run([exe, "--scene", scene, "--inverse", str(tgt), "--spp", "16", "--save-every", "2", "--final-spp", "16", "--quiet", "--iters", str(iters),
             "--out-dir", str(tmp_path / tag), "--save-params", str(out)] + cam + list(extra), env=env)
    assert r.returncode == 0, r.stdout + r.stderr
    return np.fromfile(out, np.float32)


@pytest.mark.gpu
def test_inverse_uses_the_forward_integrators_camera(host_bins, tmp_path):
    """The reference's StochasticFiniteDiffInverseIntegrator renders through forward_integrator->render, i.e. with the FORWARD
    integrator's camera (inverse_integrator.h:119,153); its own camera member is never read.  An inverse integrator built
    from a different camera object must give the same parameters, also with save renders (every 2nd iteration here)
    interleaved with the SFD renders."""
    exe = os.path.join(host_bins, "vol_render")
    a = _inverse_params(exe, tmp_path, "same")
    b = _inverse_params(exe, tmp_path, "other", extra=["--inverse-camera-pos", "3,2,-4"])
    assert a.size == 550 and np.array_equal(a, b)


@pytest.mark.gpu
def test_inverse_multi_device_allreduce_is_exact(host_bins, tmp_path):
    """VR_DEVICES with several contexts: rows are split, the int64 partial F_diff sums of all contexts are all-reduced on the
    devices (Session::allreduce_sfd -> vr_sfd_allreduce_all: one ncclAllReduce on distinct GPUs, peer copies + an integer add
    when contexts share a GPU) and every context applies the same Adam step: parameters bit-identical to one context."""
    exe = os.path.join(host_bins, "vol_render")
    one = _inverse_params(exe, tmp_path, "dev1", env=dict(os.environ, VR_DEVICES="0"))
    two = _inverse_params(exe, tmp_path, "dev00", env=dict(os.environ, VR_DEVICES="0,0"))
    assert np.array_equal(one, two)
    import subprocess
    ngpu = len(subprocess.run(["nvidia-smi", "-L"], capture_output=True, text=True).stdout.strip().splitlines())
    if ngpu >= 2:   # distinct GPUs: the NCCL path
        devs = ",".join(str(i) for i in range(min(ngpu, 4)))
        many = _inverse_params(exe, tmp_path, "devN", env=dict(os.environ, VR_DEVICES=devs))
        assert np.array_equal(one, many)


@pytest.mark.gpu
def test_inverse_checkpoint_resume_is_bit_exact(host_bins, tmp_path):
    """12 iterations in one go == 6 iterations, checkpoint, resume for 6 more: identical final parameters
    (state = parameters + Adam moments + step count + sign-vector generator)."""
    exe = os.path.join(host_bins, "vol_render")
    cam = _cam_args("pin")
    tgt = tmp_path / "target.ppm"
    scene = os.path.join(SCENES, "rand50.txt")
    assert run([exe, "--scene", scene, "--spp", "16", "--width", "24", "--height", "24", "--out", str(tgt)] + cam).returncode == 0
    common = [exe, "--scene", scene, "--inverse", str(tgt), "--spp", "16", "--save-every", "1000", "--final-spp", "16", "--quiet"] + cam
    r = run(common + ["--iters", "12", "--out-dir", str(tmp_path / "a"), "--save-params", str(tmp_path / "a.f32")])
    assert r.returncode == 0, r.stdout + r.stderr
    ck = tmp_path / "state.bin"
    r = run(common + ["--iters", "6", "--out-dir", str(tmp_path / "b"), "--checkpoint", str(ck), "--checkpoint-every", "6"])
    assert r.returncode == 0 and ck.exists(), r.stdout + r.stderr
    r = run(common + ["--iters", "12", "--out-dir", str(tmp_path / "c"), "--resume", str(ck), "--save-params", str(tmp_path / "c.f32")])
    assert r.returncode == 0, r.stdout + r.stderr
    a = np.fromfile(tmp_path / "a.f32", np.float32); c = np.fromfile(tmp_path / "c.f32", np.float32)
    assert a.size == 550 and np.array_equal(a, c)


@pytest.mark.gpu
@pytest.mark.parametrize("integ,key", [("test", "test"), ("pure", "march_pure"), ("march", "march_gauss")])
def test_vol_render_validation_integrators(host_bins, golden_march, tmp_path, integ, key):
    """TestIntegrator / PureRayMarching / RayMarchingGaussians classes of the drop-in API against the reference's images
    (black environment: the reference's random_device-seeded term vanishes, per-pixel comparison)."""
    wh, step, env_samples = golden_march["meta/rand50"]
    raw = tmp_path / "img.f32"
    r = run([os.path.join(host_bins, "vol_render"), "--scene", os.path.join(SCENES, "rand50.txt"), "--integrator", integ,
             "--width", str(int(wh)), "--height", str(int(wh)), "--step", repr(float(step)), "--env-samples", str(int(env_samples)),
             "--env", "0,0,0", "--raw", str(raw), "--out", str(tmp_path / "o.ppm")] + _cam_args("pin"))
    assert r.returncode == 0, r.stdout + r.stderr
    img = np.fromfile(raw, np.float32).reshape(int(wh), int(wh), 3)
    assert frac_within(img, golden_march[f"env0/rand50/{key}"], 1e-4, atol=1e-7) >= 0.98
