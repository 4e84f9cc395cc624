"""Seeded synthetic Gaussian-scene generator and text-format reader/writer.

Host-side tooling (not on the hot path).  Distributions follow the reference's
unseeded generators (tests/make_random.py:21-47, tests/make_nonuniform_random.py:
21-31): x,z ~ U(-1,1); y ~ U(0,2) or 2*u^2 ("nonuniform"); axis diameters
U(0.01,0.035) -> variances (d/2)^2; random rotation from the QR of a normal
3x3; density U(0.2,0.5); albedo U(0.25,0.95).  Values are rounded exactly like
the reference's f-strings (4 / 6 decimals) so files round-trip.  The 3-light
header is the one hand-added to the bundled random scenes
(scenes/gaussians/10k_random.txt:1-3).

Text format (include/scene.h:36-37,70-71):
    l x y z r g b
    g x y z cxx cxy cxz cyy cyz czz density albedo [er eg eb]
    s x y z radius sigma_a sigma_s
"""
from __future__ import annotations

import numpy as np

DEFAULT_LIGHTS = np.array(
    [[0.0, 5.0, 0.1, 50.0, 0.0, 0.0], [-3.0, 3.0, 0.3, 0.0, 30.0, 0.0], [3.0, 3.0, -0.2, 0.0, 0.0, 30.0]],
    dtype=np.float64,
)
DEFAULT_ENV = (0.53, 0.81, 0.92)  # include/scene.h:29


def _random_rotation(rs: np.random.RandomState) -> np.ndarray:
    q, _ = np.linalg.qr(rs.normal(size=(3, 3)))
    if np.linalg.det(q) < 0:
        q[:, 0] = -q[:, 0]
    return q


def generate(n: int, seed: int, nonuniform: bool = False, albedo_range=(0.25, 0.95), diameter_range=(0.01, 0.035)):
    """Return dict(mean[n,3], cov6[n,6], density[n], albedo[n], lights[k,6]) as float32 arrays."""
    rs = np.random.RandomState(seed)
    mean = np.empty((n, 3)); cov6 = np.empty((n, 6)); density = np.empty(n); albedo = np.empty(n)
    for i in range(n):
        x = rs.uniform(-1.0, 1.0)
        y = 2.0 * rs.uniform(0.0, 1.0) ** 2 if nonuniform else rs.uniform(0.0, 2.0)
        z = rs.uniform(-1.0, 1.0)
        d = rs.uniform(diameter_range[0], diameter_range[1], size=3)
        r = _random_rotation(rs)
        cov = r @ np.diag((d / 2.0) ** 2) @ r.T
        mean[i] = np.round([x, y, z], 4)
        cov6[i] = np.round([cov[0, 0], cov[0, 1], cov[0, 2], cov[1, 1], cov[1, 2], cov[2, 2]], 6)
        density[i] = np.round(rs.uniform(0.2, 0.5), 4)
        albedo[i] = np.round(rs.uniform(albedo_range[0], albedo_range[1]), 4)
        rs.uniform(0.0, 1.0, size=3)  # emission: parsed but never used by the reference
    return dict(mean=mean.astype(np.float32), cov6=cov6.astype(np.float32), density=density.astype(np.float32),
                albedo=albedo.astype(np.float32), lights=DEFAULT_LIGHTS.astype(np.float32))


EPS11 = np.array([0.02] * 3 + [0.10] * 3 + [0.05] * 3 + [0.25, 0.5], np.float32)   # gmm.h:677-706


def pack_parameters(scene: dict) -> np.ndarray:
    """11 floats per Gaussian like GaussianMixtureModel::pack_parameters (gmm.h:583-628): mean, Rodrigues
    vector of the (proper) eigenvector rotation, log sqrt(eigenvalues), log density, logit albedo."""
    n = len(scene["density"])
    c = np.asarray(scene["cov6"], np.float64)
    cov = np.empty((n, 3, 3))
    cov[:, 0, 0], cov[:, 0, 1], cov[:, 0, 2], cov[:, 1, 1], cov[:, 1, 2], cov[:, 2, 2] = c.T
    cov[:, 1, 0], cov[:, 2, 0], cov[:, 2, 1] = cov[:, 0, 1], cov[:, 0, 2], cov[:, 1, 2]
    lam, vec = np.linalg.eigh(cov)
    flip = np.linalg.det(vec) < 0
    vec[flip, :, 0] *= -1.0
    from scipy.spatial.transform import Rotation
    rod = Rotation.from_matrix(vec).as_rotvec()
    out = np.empty((n, 11), np.float32)
    out[:, 0:3] = scene["mean"]
    out[:, 3:6] = rod
    out[:, 6:9] = np.log(np.maximum(np.sqrt(np.maximum(lam, 0.0)), 1e-12))
    out[:, 9] = np.log(np.maximum(np.asarray(scene["density"], np.float64), 1e-12))
    a = np.clip(np.asarray(scene["albedo"], np.float64), 1e-7, 1.0 - 1e-7)
    out[:, 10] = np.log(a / (1.0 - a))
    return out.reshape(-1)


def orbit_cameras(n_views: int = 16, radius: float = 6.0, height: float = 1.0, look=(0.0, 1.0, 0.0)):
    """camera positions on the orbit of tests/main.cpp:94-101 (look-at + (r sin a, h, r cos a))"""
    ang = 2.0 * np.pi * np.arange(n_views) / n_views
    return [(look[0] + radius * np.sin(a), look[1] + height, look[2] + radius * np.cos(a)) for a in ang]


def write_gmm(path: str, scene: dict) -> None:
    with open(path, "w") as f:
        for l in scene.get("lights", []):
            f.write("l " + " ".join(f"{v:.4f}" for v in l) + "\n")
        f.write("\n")
        m, c, d, a = scene["mean"], scene["cov6"], scene["density"], scene["albedo"]
        for i in range(len(d)):
            f.write("g " + " ".join(f"{v:.4f}" for v in m[i]) + "  " + " ".join(f"{v:.6f}" for v in c[i])
                    + f"  {d[i]:.4f} {a[i]:.4f}\n")


def load_gmm(path: str) -> dict:
    """Token reader with the semantics of Scene::load_GMM (include/scene.h:72-120): unknown tags
    are skipped token by token; after albedo, three optional emission floats are consumed if the
    next character is not a newline."""
    lights, mean, cov6, density, albedo = [], [], [], [], []
    with open(path) as f:
        for line in f:
            tok = line.split()
            i = 0
            while i < len(tok):
                if tok[i] == "l" and i + 6 < len(tok) + 0:
                    lights.append([float(v) for v in tok[i + 1:i + 7]]); i += 7
                elif tok[i] == "g" and i + 11 < len(tok):
                    v = [float(t) for t in tok[i + 1:i + 12]]
                    mean.append(v[0:3]); cov6.append(v[3:9]); density.append(v[9]); albedo.append(v[10])
                    i += 12
                    i += min(3, len(tok) - i) if i < len(tok) and _is_float(tok[i]) else 0
                else:
                    i += 1
    return dict(mean=np.array(mean, np.float32).reshape(-1, 3), cov6=np.array(cov6, np.float32).reshape(-1, 6),
                density=np.array(density, np.float32), albedo=np.array(albedo, np.float32),
                lights=np.array(lights, np.float32).reshape(-1, 6))


def load_smm(path: str) -> dict:
    lights, sph = [], []
    with open(path) as f:
        tok = f.read().split()
    i = 0
    while i < len(tok):
        if tok[i] == "l":
            lights.append([float(v) for v in tok[i + 1:i + 7]]); i += 7
        elif tok[i] == "s":
            sph.append([float(v) for v in tok[i + 1:i + 7]]); i += 7
        else:
            i += 1
    s = np.array(sph, np.float32).reshape(-1, 6)
    return dict(center=s[:, 0:3].copy(), radius=s[:, 3].copy(), sigma_a=s[:, 4].copy(), sigma_s=s[:, 5].copy(),
                lights=np.array(lights, np.float32).reshape(-1, 6))


def _is_float(s: str) -> bool:
    try:
        float(s); return True
    except ValueError:
        return False


if __name__ == "__main__":
    import argparse
    ap = argparse.ArgumentParser(description=__doc__.splitlines()[0])
    ap.add_argument("num_gaussians", type=int); ap.add_argument("filename")
    ap.add_argument("--seed", type=int, default=1); ap.add_argument("--nonuniform", action="store_true")
    a = ap.parse_args()
    write_gmm(a.filename, generate(a.num_gaussians, a.seed, a.nonuniform))
