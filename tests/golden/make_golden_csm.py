#!/usr/bin/env python
"""Generates tests/golden/csm.npz from oracle/_ref/libref_csm.so — the cascade fit of
src/RenderEngine_Draw.cpp:149-293 run through the reference's own vendored glm (oracle/ref_csm.cpp; built by
`make -C oracle refcsm` where /root/reference exists). Run from the repo root: `python tests/golden/make_golden_csm.py`.
Inputs are stored with the outputs."""
from __future__ import annotations

import ctypes as C
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))
OUT = Path(__file__).resolve().parent

REF_CAM = np.dtype([("viewOnly", "<f4", 16), ("zNear", "<f4"), ("zFar", "<f4"), ("fov", "<f4"), ("w", "<u4"), ("h", "<u4")])
REF_LIGHT = np.dtype([("direction", "<f4", 3), ("castsShadows", "<u4"), ("numCascades", "<u4"), ("shadowCascades", "<f4", 4),
                      ("factor", "<f4")])


def ref_lib() -> C.CDLL:
    p = ROOT / "oracle" / "_ref" / "libref_csm.so"
    if not p.exists():
        raise FileNotFoundError(f"{p} missing: run `make -C oracle refcsm` in a container that has /root/reference")
    return C.CDLL(str(p))


def run_ref(cams: np.ndarray, lights: np.ndarray):
    slots = len(cams) * len(lights)
    vp = np.zeros((slots, 4, 16), np.float32)
    v, p = np.zeros_like(vp), np.zeros_like(vp)
    dist = np.zeros((len(lights), 4), np.float32)
    tanh = np.zeros(len(cams), np.float32)
    VP = C.c_void_p
    ref_lib().ref_csm(VP(cams.ctypes.data), C.c_uint32(len(cams)), VP(lights.ctypes.data), C.c_uint32(len(lights)), VP(vp.ctypes.data),
                      VP(v.ctypes.data), VP(p.ctypes.data), VP(dist.ctypes.data), VP(tanh.ctypes.data))
    return vp, v, p, dist, tanh


def random_view(rng) -> np.ndarray:
    """A rigid camera view matrix (inverse of a random pose), column-major, in float32."""
    q = rng.normal(size=4)
    q /= np.linalg.norm(q)
    w, x, y, z = q
    r = np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w)],
                  [2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w)],
                  [2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)]])
    t = rng.uniform(-200, 200, 3)
    m = np.eye(4)
    m[:3, :3] = r.T
    m[:3, 3] = -r.T @ t
    return m.T.astype(np.float32).reshape(16)  # column-major


def make_inputs(seed: int, n_cams: int, n_lights: int):
    rng = np.random.default_rng(seed)
    cams = np.zeros(n_cams, REF_CAM)
    sizes = [(1920, 1080), (1280, 720), (1024, 1024), (2560, 1080), (3840, 1080)]
    for i in range(n_cams):
        cams[i]["viewOnly"] = random_view(rng)
        cams[i]["zNear"] = rng.choice([0.1, 0.3, 1.0])
        cams[i]["zFar"] = rng.choice([100.0, 1000.0, 5000.0])
        cams[i]["fov"] = rng.choice([45.0, 60.0, 75.5, 90.0])
        cams[i]["w"], cams[i]["h"] = sizes[i % len(sizes)]
    lights = np.zeros(n_lights, REF_LIGHT)
    for i in range(n_lights):
        d = rng.normal(size=3)
        d /= np.linalg.norm(d)
        if i == 1:
            d = np.array([0.3, -0.9, 0.2]) / np.linalg.norm([0.3, -0.9, 0.2])
        lights[i]["direction"] = d
        lights[i]["castsShadows"] = 0 if i == 2 else 1
        lights[i]["numCascades"] = 1 + (i + seed) % 4
        lights[i]["shadowCascades"] = np.sort(rng.uniform(0.02, 0.9, 4)).astype(np.float32)
        lights[i]["factor"] = rng.choice([1.0, 0.5, 0.25])
    return cams, lights


def main():
    blobs = {}
    for name, (seed, nc, nl) in {"a": (1, 1, 1), "b": (2, 3, 4), "c": (3, 5, 2)}.items():
        cams, lights = make_inputs(seed, nc, nl)
        vp, v, p, dist, tanh = run_ref(cams, lights)
        assert np.isfinite(vp).all()
        blobs.update({f"{name}_cams": cams, f"{name}_lights": lights, f"{name}_viewproj": vp, f"{name}_view": v, f"{name}_proj": p,
                      f"{name}_dist": dist, f"{name}_tanhalf": tanh})
    np.savez_compressed(OUT / "csm.npz", **blobs)
    print("wrote", OUT / "csm.npz", {k: v.shape for k, v in blobs.items() if k.endswith("viewproj")})


if __name__ == "__main__":
    main()
