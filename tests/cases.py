"""Hand-worked known-answer cull cases. Each builder returns a CullCase whose expected
per-slot instance counts (and optionally id slices) follow from shaders/defaultcull.csh by hand with
exactly representable numbers. Used by the CPU oracle tests and by the GPU parity tests."""
from __future__ import annotations

import math
from dataclasses import dataclass, field

import numpy as np

from ravengine_b200 import scenes

IDENT = np.eye(4, dtype=np.float32).T.reshape(16)


def trs(x, y, z, s=1.0) -> np.ndarray:
    m = np.zeros(16, np.float32)
    m[0] = m[5] = m[10] = s
    m[12:15] = (x, y, z)
    m[15] = 1.0
    return m


@dataclass
class CullCase:
    materials: list
    model: np.ndarray
    layers: np.ndarray
    flags: np.ndarray
    expect_counts: list
    expect_ids: list | None = None  # [(start slot in the culling buffer, ids)]
    view_proj: np.ndarray = field(default_factory=lambda: IDENT.copy())
    cam_pos: tuple = (0.0, 0.0, 0.0)
    cam_layers: int = 0xFFFFFFFF
    shadow: bool = False
    pyr_dim: int = 4
    pyr_levels: int = 3
    pyramid: np.ndarray = field(default_factory=lambda: np.zeros(21, np.float32))


def _one(positions, radius=0.5, lods=(math.inf,), flags=0xFFFD, layers=0xFFFFFFFF, scale=1.0, **kw) -> CullCase:
    n = len(positions)
    model = np.stack([trs(*p, s=scale) for p in positions])
    mesh = scenes.simple_mesh(radius, lods)
    mats = [scenes.MaterialDraws([scenes.DrawCommand(mesh, np.arange(n, dtype=np.uint32))])]
    fl = np.full(n, flags, np.uint16) if np.isscalar(flags) else np.asarray(flags, np.uint16)
    la = np.full(n, layers, np.uint32) if np.isscalar(layers) else np.asarray(layers, np.uint32)
    return CullCase(mats, model, la, fl, **kw)


# ---- frustum (defaultcull.csh:71-92, 218). Identity viewProj => frustum = the cube [-1,1]^3 ----
def on_plane_rejected():
    # centre x=1.5, r=0.5: distance to the right plane = -0.5, + r = 0 -> "> 0" is strict -> culled
    return _one([(1.5, 0, 0)], expect_counts=[0])


def just_inside_visible():
    return _one([(1.5, 0, 0)], radius=float(np.nextafter(np.float32(0.5), np.float32(1))), expect_counts=[1], expect_ids=[(0, [0])])


def scaled_radius_uses_largest_axis():
    # model scale 2 doubles the world radius (:203-214): r = 0.25*2 = 0.5 -> same as on_plane_rejected
    return _one([(1.5, 0, 0)], radius=0.25, scale=2.0, expect_counts=[0])


def frustum_bit_clear_skips_test():
    return _one([(1.5, 0, 0), (40, 40, 40)], flags=0xFFFC, expect_counts=[2], expect_ids=[(0, [0, 1])])


# ---- layer / shadow filters (:173-193) ----
def layer_mask_zero():
    return _one([(0, 0, 0), (0, 0, 0)], layers=[0, 0b01], cam_layers=0b10, expect_counts=[0])


def layer_mask_overlap():
    return _one([(0, 0, 0), (0, 0, 0), (0, 0, 0)], layers=[0b01, 0b10, 0b11], cam_layers=0b10, expect_counts=[2],
                expect_ids=[(0, [1, 2])])


def shadow_mode_needs_casts_shadows():
    return _one([(0, 0, 0), (0, 0, 0)], flags=[0xFFFD & ~4, 0xFFFD], shadow=True, expect_counts=[1], expect_ids=[(0, [1])])


def shadow_bit_ignored_outside_shadow_mode():
    return _one([(0, 0, 0), (0, 0, 0)], flags=[0xFFFD & ~4, 0xFFFD], shadow=False, expect_counts=[2])


# ---- LOD selection (:278-303) ----
def lod_thresholds():
    # lodDist {0,10,50}, camera at origin: d=9.99 -> 0; d=10 -> 0 (strict >); d=10.01 -> 1; d=50.01 -> 2
    return _one([(9.99, 0, 0), (10.0, 0, 0), (10.01, 0, 0), (50.01, 0, 0)], lods=(0.0, 10.0, 50.0), flags=0xFFFC,
                expect_counts=[2, 1, 1], expect_ids=[(0, [0, 1]), (4, [2]), (8, [3])])


def lod_unsquared_best_quirk():
    # currentBest starts as lodDist[0] UNsquared (=3). LOD 1 (1.5^2 = 2.25) would need 2.25 > 3: never chosen,
    # even at d=2 where LOD 0's own threshold (9) is not met. Both land in LOD 0.
    return _one([(2.0, 0, 0), (4.0, 0, 0)], lods=(3.0, 1.5), flags=0xFFFC, expect_counts=[2, 0], expect_ids=[(0, [0, 1])])


def lod_unsorted_thresholds():
    # "price is right" with unsorted distances {0, 50, 10}: d=20 -> LOD 2 (100), d=60 -> LOD 1 (2500)
    return _one([(20.0, 0, 0), (60.0, 0, 0), (5.0, 0, 0)], lods=(0.0, 50.0, 10.0), flags=0xFFFC, expect_counts=[1, 1, 1],
                expect_ids=[(0, [2]), (3, [1]), (6, [0])])


def lod_camera_offset():
    # distance is measured from camPos (:283), not from the origin
    return _one([(30.0, 0, 0)], lods=(0.0, 10.0, 50.0), flags=0xFFFC, cam_pos=(25.0, 0.0, 0.0), expect_counts=[1, 0, 0])


def single_lod_infinite_threshold():
    return _one([(0, 0, 0)], lods=(math.inf,), expect_counts=[1])


# ---- Hi-Z (:108-145, 221-274). centre (0,0,.5), r=.25, identity viewProj, 4x4 pyramid:
# corners x,y in +-0.25, z in {.25,.75} -> refZ .75; width .5*4 = 2 -> level 0; footprint texels [1..3]^2 ----
def _hiz(pyr, z=0.5, **kw):
    return _one([(0, 0, z)], radius=0.25, flags=0xFFFF, pyramid=np.asarray(pyr, np.float32), **kw)


def hiz_occluded():
    return _hiz(np.full(21, 0.8), expect_counts=[0])


def hiz_equal_depth_is_visible():
    p = np.full(21, 0.8, np.float32)
    p[3 * 4 + 3] = 0.75  # refZ >= minDepth is inclusive (:272)
    return _hiz(p, expect_counts=[1])


def hiz_texel_outside_footprint_ignored():
    p = np.full(21, 0.8, np.float32)
    p[0] = 0.0  # texel (0,0) is not in the footprint [1..3]^2
    return _hiz(p, expect_counts=[0])


def hiz_bit_clear_skips_test():
    return _one([(0, 0, 0.5)], radius=0.25, flags=0xFFFD, pyramid=np.full(21, 1.0, np.float32), expect_counts=[1])


def hiz_near_plane_straddle_is_visible():
    # z in {.65, 1.15}: refZ > 1 -> the depth test is skipped (:227)
    return _hiz(np.full(21, 1.0), z=0.9, expect_counts=[1])


def hiz_all_near_occludes():
    return _hiz(np.full(21, 1.0), z=0.5, expect_counts=[0])


def hiz_zero_pyramid_never_occludes():
    return _hiz(np.zeros(21), expect_counts=[1])  # first frame: cleared depth = far


def hiz_bbox_wider_than_screen_is_visible():
    # viewProj = diag(4,1,1,1), r=.4: NDC x in +-1.6 -> width 3.2*4 = 12.8 -> ceil(log2)=4 -> level 3 > levels-1 (:234-237)
    vp = IDENT.copy()
    vp[0] = 4.0
    return _one([(0, 0, 0.5)], radius=0.4, flags=0xFFFF, view_proj=vp, pyramid=np.full(21, 1.0, np.float32), expect_counts=[1])


def hiz_ndc_width_not_halved():
    # 8x8 pyramid: width = 0.5 (NDC units, not halved) * 8 = 4 -> level 1. Level 0 would occlude, level 1 does not.
    p = np.full(85, 1.0, np.float32)
    p[64:80] = 0.0
    return _one([(0, 0, 0.5)], radius=0.25, flags=0xFFFF, pyramid=p, pyr_dim=8, pyr_levels=4, expect_counts=[1])


# ---- CUBO prefix / slot layout (:150-168, 305-312) ----
def two_commands_slot_layout():
    n = 22
    model = np.stack([trs(0, 0, 0) for _ in range(n)])
    mesh_a, mesh_b = scenes.simple_mesh(0.5, (0.0, 5.0)), scenes.simple_mesh(0.5)
    ids_a, ids_b = np.array([10, 11, 12], np.uint32), np.array([20, 21], np.uint32)
    model[11] = trs(7.0, 0, 0)  # d=7 > 5 -> LOD 1 of command A
    mats = [scenes.MaterialDraws([scenes.DrawCommand(mesh_a, ids_a), scenes.DrawCommand(mesh_b, ids_b)])]
    # command A: slots 0,1, ids at 0.. (LOD 0) and 3.. (LOD 1); command B: slot 2, ids at cullingBufferOffset 6
    return CullCase(mats, model, np.full(n, 0xFFFFFFFF, np.uint32), np.full(n, 0xFFFC, np.uint16), expect_counts=[2, 1, 2],
                    expect_ids=[(0, [10, 12]), (3, [11]), (6, [20, 21])])


CULL_CASES = {f.__name__: f for f in [
    on_plane_rejected, just_inside_visible, scaled_radius_uses_largest_axis, frustum_bit_clear_skips_test, layer_mask_zero,
    layer_mask_overlap, shadow_mode_needs_casts_shadows, shadow_bit_ignored_outside_shadow_mode, lod_thresholds,
    lod_unsquared_best_quirk, lod_unsorted_thresholds, lod_camera_offset, single_lod_infinite_threshold, hiz_occluded,
    hiz_equal_depth_is_visible, hiz_texel_outside_footprint_ignored, hiz_bit_clear_skips_test,
    hiz_near_plane_straddle_is_visible, hiz_all_near_occludes, hiz_zero_pyramid_never_occludes,
    hiz_bbox_wider_than_screen_is_visible, hiz_ndc_width_not_halved, two_commands_slot_layout]}
