"""Synthetic scene data (numpy) for the configs in BASELINE.json — data generation only,
no hot-path arithmetic. Everything is seeded; the same arrays feed the CUDA path and the oracle.

Conventions follow the reference: column-major mat4 flattened as m[col*4+row]; reverse-Z
projection perspectiveRH_ZO(fovy, aspect, zFar, zNear) (include/RavEngine/mathtypes.hpp:141-149),
viewProj = proj * inverse(cameraWorld) (src/App_priv.hpp:409-418); image row 0 = NDC y=+1.
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field

import numpy as np


def colmajor(m_rc: np.ndarray) -> np.ndarray:
    """4x4 [row, col] -> 16 floats m[col*4+row] (float32)."""
    return np.ascontiguousarray(np.asarray(m_rc, dtype=np.float64).T.reshape(16)).astype(np.float32)


def perspective_reverse_z(fovy_deg: float, aspect: float, z_near: float, z_far: float) -> np.ndarray:
    """glm::perspectiveRH_ZO(fovy, aspect, zNear:=z_far, zFar:=z_near) as [row, col] float64."""
    t = math.tan(math.radians(fovy_deg) / 2.0)
    zn, zf = z_far, z_near  # swapped on purpose (reverse-Z)
    m = np.zeros((4, 4))
    m[0, 0] = 1.0 / (aspect * t)
    m[1, 1] = 1.0 / t
    m[2, 2] = zf / (zn - zf)
    m[3, 2] = -1.0
    m[2, 3] = -(zf * zn) / (zf - zn)
    return m


def ortho_reverse_z(l, r, b, t, z_near, z_far) -> np.ndarray:
    """glm::orthoRH_ZO(l,r,b,t, zNear:=z_far, zFar:=z_near) as [row, col]."""
    zn, zf = z_far, z_near
    m = np.eye(4)
    m[0, 0] = 2.0 / (r - l)
    m[1, 1] = 2.0 / (t - b)
    m[2, 2] = -1.0 / (zf - zn)
    m[0, 3] = -(r + l) / (r - l)
    m[1, 3] = -(t + b) / (t - b)
    m[2, 3] = -zn / (zf - zn)
    return m


def look_at(eye, target, up=(0, 1, 0)) -> np.ndarray:
    """View matrix (world -> view), right-handed, looking down -Z; [row, col]."""
    eye, target, up = (np.asarray(v, dtype=np.float64) for v in (eye, target, up))
    f = target - eye
    f /= np.linalg.norm(f)
    s = np.cross(f, up)
    s /= np.linalg.norm(s)
    u = np.cross(s, f)
    m = np.eye(4)
    m[0, :3], m[1, :3], m[2, :3] = s, u, -f
    m[0, 3], m[1, 3], m[2, 3] = -s @ eye, -u @ eye, f @ eye
    return m


@dataclass
class Camera:
    view: np.ndarray  # [row, col] float64
    proj: np.ndarray
    pos: np.ndarray
    z_near: float
    z_far: float

    @property
    def view_proj_cm(self) -> np.ndarray:
        return colmajor(self.proj @ self.view)

    @property
    def view_cm(self) -> np.ndarray:
        return colmajor(self.view)

    @property
    def inv_proj_cm(self) -> np.ndarray:
        return colmajor(np.linalg.inv(self.proj))


def default_camera(z_near=0.1, z_far=100.0, fovy=60.0, aspect=16 / 9, eye=(0, 0, 0), target=(0, 0, -1)) -> Camera:
    return Camera(look_at(eye, target), perspective_reverse_z(fovy, aspect, z_near, z_far), np.asarray(eye, np.float32),
                  z_near, z_far)


def random_transforms(rng: np.random.Generator, n: int, extent_xy: float, extent_y: float, smin=0.5, smax=2.0) -> np.ndarray:
    """n x 16 float32 column-major T*R*S (uniform scale), uniform random rotations."""
    q = rng.normal(size=(n, 4))
    q /= np.linalg.norm(q, axis=1, keepdims=True)
    x, y, z, w = q.T
    s = rng.uniform(smin, smax, size=n)
    m = np.zeros((n, 16))
    m[:, 0] = (1 - 2 * (y * y + z * z)) * s
    m[:, 1] = (2 * (x * y + w * z)) * s
    m[:, 2] = (2 * (x * z - w * y)) * s
    m[:, 4] = (2 * (x * y - w * z)) * s
    m[:, 5] = (1 - 2 * (x * x + z * z)) * s
    m[:, 6] = (2 * (y * z + w * x)) * s
    m[:, 8] = (2 * (x * z + w * y)) * s
    m[:, 9] = (2 * (y * z - w * x)) * s
    m[:, 10] = (1 - 2 * (x * x + y * y)) * s
    m[:, 12] = rng.uniform(-extent_xy, extent_xy, size=n)
    m[:, 13] = rng.uniform(-extent_y, extent_y, size=n)
    m[:, 14] = rng.uniform(-extent_xy, extent_xy, size=n)
    m[:, 15] = 1.0
    return m.astype(np.float32)


def synth_depth(width: int, height: int, cam: Camera, fovy=60.0, seed=7, n_occluders=200, ground_y=-60.0,
                occ_dist=(40.0, 600.0)) -> np.ndarray:
    """Reverse-Z depth image (height x width float32, row 0 = top): ground plane + screen-space
    rectangular occluders at constant view depth (analytic, no rasteriser)."""
    rng = np.random.default_rng(seed)
    zn, zf = cam.z_near, cam.z_far
    t = math.tan(math.radians(fovy) / 2.0)

    def ndc_z(d):
        d = np.asarray(d, dtype=np.float64)
        z = zn * (zf - d) / ((zf - zn) * d)
        return np.clip(z, 0.0, 1.0)

    y_ndc = 1.0 - 2.0 * (np.arange(height) + 0.5) / height
    dir_y = y_ndc * t
    row = np.zeros(height)
    below = dir_y < -1e-6
    d = np.full(height, np.inf)
    d[below] = (-ground_y) / (-dir_y[below])
    row[below] = np.where(d[below] < zf, ndc_z(d[below]), 0.0)
    img = np.repeat(row[:, None], width, axis=1)
    for _ in range(n_occluders):
        dist = rng.uniform(*occ_dist)
        cx, cy = rng.uniform(-1, 1), rng.uniform(-0.6, 0.3)
        hw, hh = rng.uniform(0.01, 0.12), rng.uniform(0.02, 0.25)
        x0, x1 = int((cx - hw + 1) / 2 * width), int((cx + hw + 1) / 2 * width)
        y0, y1 = int((1 - (cy + hh)) / 2 * height), int((1 - (cy - hh)) / 2 * height)
        x0, x1, y0, y1 = max(x0, 0), min(x1, width), max(y0, 0), min(y1, height)
        if x1 > x0 and y1 > y0:
            img[y0:y1, x0:x1] = np.maximum(img[y0:y1, x0:x1], float(ndc_z(dist)))
    return img.astype(np.float32)


# ---- static-mesh draw description (World::RenderData stand-in, World.hpp:339-349) ----
@dataclass
class LODMesh:
    index_count: int
    index_start: int
    base_vertex: int


@dataclass
class MeshCollection:
    radius: float
    lod_distances: list[float]  # metres, unsquared; single LOD = [inf] (MeshCollection.cpp:24)
    lods: list[LODMesh]


@dataclass
class DrawCommand:
    mesh: MeshCollection
    entity_ids: np.ndarray  # dense u32 list (BufferedVRAMSparseSet dense array)


@dataclass
class MaterialDraws:
    commands: list[DrawCommand] = field(default_factory=list)


CUBE_RADIUS = 1.7320508  # meshes/cube.obj: max |v|
SPHERE_RADIUS = 1.74


def simple_mesh(radius: float, lod_distances=(math.inf,), base=0) -> MeshCollection:
    lods = [LODMesh(index_count=36 * max(1, 4 >> i) * 3, index_start=base + 1000 * i, base_vertex=base + 500 * i)
            for i in range(len(lod_distances))]
    return MeshCollection(radius, [float(x) for x in lod_distances], lods)


def point_lights_bytes(positions: np.ndarray, intensities: np.ndarray) -> np.ndarray:
    """n x 512-byte PointLight structs (only position @448 and intensity @472 are read by the path)."""
    n = len(positions)
    raw = np.zeros((n, 128), dtype=np.float32)
    raw[:, 112:115] = positions
    raw[:, 115:118] = 1.0  # colour
    raw[:, 118] = intensities
    return raw.view(np.uint8).reshape(n, 512)


def spot_lights_bytes(positions: np.ndarray, intensities: np.ndarray) -> np.ndarray:
    """n x 168-byte SpotLight structs (worldTransform @64 column 3 = position, intensity @140)."""
    n = len(positions)
    raw = np.zeros((n, 42), dtype=np.float32)
    for d in (0, 5, 10, 15):
        raw[:, 16 + d] = 1.0  # identity rotation
    raw[:, 28:31] = positions
    raw[:, 35] = intensities
    return raw.view(np.uint8).reshape(n, 168)


def skinning_inputs(rng: np.random.Generator, n_verts: int, n_objects: int, n_bones: int, coherent_joints=False):
    """C3-style inputs: vertex streams, 4 joint influences per vertex, rigid joint palettes."""
    pos = rng.uniform(0, 1, size=(n_verts, 3)).astype(np.float32)

    def unit(n):
        v = rng.normal(size=(n, 3))
        return (v / np.linalg.norm(v, axis=1, keepdims=True)).astype(np.float32)

    nrm, tan, bit = unit(n_verts), unit(n_verts), unit(n_verts)
    uv = rng.uniform(0, 1, size=(n_verts, 2)).astype(np.float32)
    if coherent_joints:  # neighbouring vertices share a primary bone, like a real mesh
        primary = (np.arange(n_verts) * n_bones // max(n_verts, 1)).astype(np.int64)
        joints = (primary[:, None] + rng.integers(0, 3, size=(n_verts, 4))) % n_bones
    else:
        joints = rng.integers(0, n_bones, size=(n_verts, 4))
    w = rng.uniform(0.05, 1, size=(n_verts, 4))
    w /= w.sum(axis=1, keepdims=True)
    weights = np.zeros((n_verts, 8), dtype=np.uint32)
    weights[:, 0::2] = joints.astype(np.uint32)
    weights[:, 1::2] = w.astype(np.float32).view(np.uint32)
    nb = n_objects * n_bones
    ang = rng.uniform(0, math.radians(30), size=nb)
    axis = rng.normal(size=(nb, 3))
    axis /= np.linalg.norm(axis, axis=1, keepdims=True)
    c, s = np.cos(ang), np.sin(ang)
    x, y, z = axis.T
    pose = np.zeros((nb, 16))
    pose[:, 0] = c + x * x * (1 - c)
    pose[:, 1] = y * x * (1 - c) + z * s
    pose[:, 2] = z * x * (1 - c) - y * s
    pose[:, 4] = x * y * (1 - c) - z * s
    pose[:, 5] = c + y * y * (1 - c)
    pose[:, 6] = z * y * (1 - c) + x * s
    pose[:, 8] = x * z * (1 - c) + y * s
    pose[:, 9] = y * z * (1 - c) - x * s
    pose[:, 10] = c + z * z * (1 - c)
    pose[:, 12:15] = rng.uniform(-0.2, 0.2, size=(nb, 3))
    pose[:, 15] = 1.0
    return dict(pos=pos, nrm=nrm, tan=tan, bit=bit, uv=uv, weights=weights, pose=pose.astype(np.float32))
