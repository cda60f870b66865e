"""Known-answer cases that pin the CPU oracle (SURVEY 8c "pins we must create").

The reference ships no test, golden vector or fixture for the render-prep path, so these cases are
worked by hand from the shader source (cited per test) with exactly representable numbers. They run
on the CPU; the same cases run through the CUDA path in tests/test_gpu_known_answers.py.
"""
import math

import numpy as np
import pytest

import oracle
from oracle import host_tables
from ravengine_b200 import GridAssignUBO, GridBuildUBO, SkinningPrepareUBO, SkinningUBO, scenes
from tests import cases
from tests.helpers import make_ubo, oracle_cull

IDENT = np.eye(4, dtype=np.float32).T.reshape(16)


def _counts(tables, out):
    return out[int(tables.cubos[0]["indirectOutputBufferBindlessHandle"])][:, 1].tolist()


def _ids(tables, out):
    return out[int(tables.cubos[0]["idOutputBufferBindlessHandle"])]


@pytest.mark.parametrize("name", sorted(cases.CULL_CASES))
def test_cull_known_answer(name):
    c = cases.CULL_CASES[name]()
    t = host_tables.build_static_tables(c.materials)
    ubo = make_ubo(c.view_proj, c.cam_pos, t.num_cubos, layers=c.cam_layers, shadow=c.shadow)
    out = oracle_cull(t, ubo, c.model, c.layers, c.flags, c.pyramid, c.pyr_dim, c.pyr_levels)
    assert _counts(t, out) == c.expect_counts, name
    if c.expect_ids is not None:
        got = _ids(t, out)
        for (start, ids) in c.expect_ids:
            assert sorted(got[start:start + len(ids)].tolist()) == sorted(ids), name


def test_ceil_log2_exact():
    l = oracle.lib()
    # defaultcull.csh:231 ceil(log2(x)) from the exponent: exact at and around powers of two
    for v, want in [(1.0, 0.0), (2.0, 1.0), (2.0000002, 2.0), (3.9999998, 2.0), (4.0, 2.0), (4.0000005, 3.0), (0.5, -1.0),
                    (0.75, 0.0), (1024.0, 10.0), (1025.0, 11.0)]:
        assert l.orc_ceil_log2(np.float32(v)) == want, v
    assert l.orc_ceil_log2(np.float32(0.0)) == -math.inf
    assert l.orc_ceil_log2(np.float32(math.inf)) == math.inf
    assert math.isnan(l.orc_ceil_log2(np.float32(-1.0)))


def test_sampler_footprint_and_clamp():
    l = oracle.lib()
    img = np.arange(16, dtype=np.float32).reshape(4, 4)  # img[y][x] = 4y + x
    s = lambda u, v: l.orc_sample_min(img.ctypes.data, 4, 4, np.float32(u), np.float32(v))
    # u*W-0.5 = 1.0 exactly -> texels {1,2} on each axis; min = img[1][1] = 5
    assert s(0.375, 0.375) == 5.0
    # just left of the texel centre -> floor(0.99..) = 0 -> texels {0,1}: min = img[0][0]
    assert s(0.3749, 0.3749) == 0.0
    # clamp to edge: u=0 -> x=-0.5 -> both taps texel 0; v=1 -> y=3.5 -> texels {3,3}
    assert s(0.0, 1.0) == 12.0
    assert s(1.0, 1.0) == 15.0
    assert s(-5.0, 7.0) == 12.0  # far outside still clamps


def test_pyramid_4x4_and_2x2():
    depth = (1.0 - np.arange(16, dtype=np.float32) / 16.0).reshape(4, 4)  # decreasing: min is bottom-right
    pyr, dim, levels = oracle.pyramid_build(depth)
    assert (dim, levels) == (4, 3) and pyr.size == 21
    # mip 0 (depthpyramidcopy.fsh:10-20): sampled at texel centres -> footprint {x,x+1}x{y,y+1} clamped
    want0 = np.array([[depth[min(y + 1, 3), min(x + 1, 3)] for x in range(4)] for y in range(4)], np.float32)
    np.testing.assert_array_equal(pyr[:16].reshape(4, 4), want0)
    # mips 1.. (depthpyramid.csh:9-28): exact 2x2 min of the previous level
    want1 = want0.reshape(2, 2, 2, 2).min(axis=(1, 3))
    np.testing.assert_array_equal(pyr[16:20].reshape(2, 2), want1)
    assert pyr[20] == want1.min() == depth[3, 3]
    d2 = np.array([[0.5, 0.25], [0.75, 0.125]], np.float32)
    p2, dim2, lv2 = oracle.pyramid_build(d2)
    assert (dim2, lv2) == (2, 2)
    np.testing.assert_array_equal(p2, np.array([0.125, 0.125, 0.125, 0.125, 0.125], np.float32))


@pytest.mark.parametrize("w,dim,levels", [(1, 1, 1), (2, 2, 2), (3, 2, 2), (1023, 512, 10), (1024, 1024, 11), (2160, 1024, 11), (4096, 1024, 11)])
def test_pyramid_dims(w, dim, levels):
    # src/DepthPyramid.cpp:10-17: width capped at 1024, rounded down to a power of two
    assert oracle.pyramid_dims(w) == (dim, levels)


def test_pyramid_nonsquare_source_subsamples():
    """3840x2160 -> 1024^2: mip 0 point-samples a 2x2 footprint of the source (NOT a full reduce, SURVEY a-1 quirk)."""
    rng = np.random.default_rng(3)
    depth = rng.random((27, 48), dtype=np.float32)  # same 16:9 shape, small
    pyr, dim, levels = oracle.pyramid_build(depth)
    assert dim == 16
    for (x, y) in [(0, 0), (5, 7), (15, 15)]:
        fx = math.floor((x + 0.5) / 16 * 48 - 0.5)
        fy = math.floor((y + 0.5) / 16 * 27 - 0.5)
        xs = [min(max(fx, 0), 47), min(max(fx + 1, 0), 47)]
        ys = [min(max(fy, 0), 26), min(max(fy + 1, 0), 26)]
        assert pyr[y * 16 + x] == min(depth[j, i] for j in ys for i in xs)


def test_skinning_identity_palette_dyadic_weights_exact():
    """skinning_cs.csh:98-137 with identity palettes and weights (1/2,1/4,1/8,1/8): out == in bit for bit."""
    rng = np.random.default_rng(1)
    d = scenes.skinning_inputs(rng, 100, 3, 8)
    d["pose"][:] = IDENT
    d["weights"][:, 1::2] = np.array([0.5, 0.25, 0.125, 0.125], np.float32).view(np.uint32)
    ins = [d["pos"], d["nrm"], d["tan"], d["bit"], d["uv"]]
    outs = [np.zeros((300, k), np.float32) for k in (3, 3, 3, 3, 2)]
    oracle.skin(SkinningUBO(3, 100, 8, 0, 0, 0), outs, ins, d["pose"], d["weights"])
    for o, i in zip(outs, ins):
        np.testing.assert_array_equal(o, np.tile(i, (3, 1)))


def test_skinning_single_joint_translation():
    """one joint with weight 1, pure translation: pos moves, normals/tangents do not (mat3 only, no renormalise)."""
    pose = np.tile(IDENT, (2, 1))
    pose[1, 12:15] = (1.0, 2.0, 3.0)
    w = np.zeros((1, 8), np.uint32)
    w[0, 0] = 1
    w[0, 1::2] = np.array([1.0, 0, 0, 0], np.float32).view(np.uint32)
    ins = [np.array([[0.5, 0.25, 0.125]], np.float32), np.array([[0, 1, 0]], np.float32), np.array([[1, 0, 0]], np.float32),
           np.array([[0, 0, 1]], np.float32), np.array([[0.3, 0.7]], np.float32)]
    outs = [np.zeros((1, k), np.float32) for k in (3, 3, 3, 3, 2)]
    oracle.skin(SkinningUBO(1, 1, 2, 0, 0, 0), outs, ins, pose, w)
    np.testing.assert_array_equal(outs[0], [[1.5, 2.25, 3.125]])
    for k in (1, 2, 3, 4):
        np.testing.assert_array_equal(outs[k], ins[k])


def test_skinned_drawcall_prepare_known_answer():
    # skinned_mesh_drawcall.csh:25-39
    ubo = SkinningPrepareUBO(1000, 77, 36, 24, 3, 5, 11)
    cmds = np.full((10, 5), 0xABCD, np.uint32)
    oracle.skinned_drawcall_prepare(ubo, cmds)
    np.testing.assert_array_equal(cmds[5:8], [[36, 0, 1000, 77, 11], [36, 0, 1036, 101, 12], [36, 0, 1072, 125, 13]])
    assert (cmds[:5] == 0xABCD).all() and (cmds[8:] == 0xABCD).all()


def _cluster_array(mins, maxs):
    cl = np.zeros((len(mins), 40), np.uint32)
    cl[:, 0:3] = np.asarray(mins, np.float32).view(np.uint32)
    cl[:, 3:6] = np.asarray(maxs, np.float32).view(np.uint32)
    cl[:, 6:] = 0xDEADBEEF  # stale counts / indices
    return cl


def test_cluster_assign_tangent_and_cap():
    """cluster_assign_lights.csh:28-35 (<= is inclusive) and :60-67 (first 16 by index)."""
    cl = _cluster_array([(0, 0, -2)], [(1, 1, -1)])
    i4 = np.float32(0.01) * np.float32(4.0)  # sqrt(I/0.01) == 2 exactly
    pos = [(3.0, 0.5, -1.5), (3.0 + 2.0 ** -20, 0.5, -1.5)] + [(0.5, 0.5, -1.5)] * 18
    inten = [i4, i4] + [i4] * 18
    pts = scenes.point_lights_bytes(np.asarray(pos, np.float32), np.asarray(inten, np.float32))
    a = GridAssignUBO()
    a.viewMat[:] = [float(x) for x in IDENT]
    a.pointLightCount, a.spotLightCount = 20, 0
    oracle.cluster_assign(a, cl, pts, None, 1)
    assert cl[0, 6] == 16 and cl[0, 7] == 0
    # light 0 is tangent (hit), light 1 misses by 2^-20, then 2..16 fill the list; 17..19 are dropped
    assert cl[0, 8:24].tolist() == [0] + list(range(2, 17))
    assert (cl[0, 24:] == 0xDEADBEEF).all()  # spot indices untouched


def test_cluster_assign_spots_capped_at_16():
    """the shader has no capacity check for spots (:80-83, OOB in the reference); defined here as capped."""
    cl = _cluster_array([(0, 0, -2)], [(1, 1, -1)])
    spots = scenes.spot_lights_bytes(np.tile(np.float32([0.5, 0.5, -1.5]), (20, 1)), np.full(20, 1.0, np.float32))
    a = GridAssignUBO()
    a.viewMat[:] = [float(x) for x in IDENT]
    a.pointLightCount, a.spotLightCount = 0, 20
    oracle.cluster_assign(a, cl, None, spots, 1)
    assert cl[0, 6] == 0 and cl[0, 7] == 16
    assert cl[0, 24:40].tolist() == list(range(16))


def test_cluster_build_known_answer():
    """cluster_build_grid.csh:41-79 with an identity inverse projection, 2x2x1 grid, near 1, far 4."""
    u = GridBuildUBO()
    u.invProj[:] = [float(x) for x in IDENT]
    u.gridSize[:] = (2, 2, 1)
    u.zNear, u.zFar = 1.0, 4.0
    u.screenDim[:] = (8, 4)
    cl = np.zeros((4, 40), np.uint32)
    oracle.cluster_build(u, cl)
    f = cl.view(np.float32)
    # identity invProj: view point = (ndc.x, ndc.y, 0) -> direction.z = 0 -> t = zDist / -0 = -inf:
    # the shader's math degenerates; what matters is that the oracle reproduces IEEE results, not that they are useful
    assert np.isinf(f[:, 0:6]).any() or np.isnan(f[:, 0:6]).any()
    # a real projection: slices are zFar*(zNear/zFar)^(k/gridZ); with gridZ=1: near plane z=-4, far plane z=-1
    cam = scenes.default_camera(z_near=1.0, z_far=4.0, fovy=90.0, aspect=1.0)
    u.invProj[:] = [float(x) for x in cam.inv_proj_cm]
    u.screenDim[:] = (4, 4)
    oracle.cluster_build(u, cl)
    f = cl.view(np.float32)
    # tile (0,0): screen [0,2]x[0,2] -> ndc [-1,0]^2; fov 90 => at depth z the frustum spans [-|z|,|z|]
    np.testing.assert_allclose(f[0, 0:3], [-4.0, -4.0, -4.0], rtol=1e-6)
    np.testing.assert_allclose(f[0, 3:6], [0.0, 0.0, -1.0], rtol=1e-6, atol=1e-7)
    np.testing.assert_allclose(f[3, 0:3], [0.0, 0.0, -4.0], rtol=1e-6, atol=1e-7)
    np.testing.assert_allclose(f[3, 3:6], [4.0, 4.0, -1.0], rtol=1e-6)
    assert (cl[:, 6:] == 0).all()  # counts / indices untouched by the build
