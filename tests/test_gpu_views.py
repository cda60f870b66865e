"""rve_cull_views (view-batched cull, SURVEY 8e partition B / 8f-2): one launch over V views must leave, in view v's
slice of every output buffer, exactly what the oracle produces for a single-view cull with view v's parameters."""
import math

import numpy as np
import pytest

import oracle
from oracle import host_tables
from ravengine_b200 import scenes
from ravengine_b200.api import cull_views
from tests.helpers import compare_cull_outputs, make_ubo, oracle_cull

pytestmark = pytest.mark.gpu


def _views():
    """(camera, depth size, shadow, camera layers): different positions / directions / projections per view."""
    out = []
    eyes = [((0, 0, 0), (0, 0, -1)), ((120, 10, 80), (0, 0, 0)), ((-200, 30, -50), (50, -10, 100)), ((0, 150, 0), (1, 0, 1)),
            ((300, 5, 300), (0, 0, 0)), ((10, 0, 10), (200, 0, 10)), ((0, -20, 0), (0, 0, 1))]
    for i, (eye, tgt) in enumerate(eyes):
        cam = scenes.default_camera(z_near=0.1, z_far=[1000.0, 400.0, 2000.0][i % 3], fovy=[60.0, 45.0, 90.0][i % 3], eye=eye, target=tgt)
        out.append((cam, [(640, 360), (512, 512), (300, 200)][i % 3], i == 3, 0xFFFFFFFF if i != 5 else 0b0110))
    return out


@pytest.mark.parametrize("n,shuffle", [(70_001, False), (20_000, True)])
def test_cull_views_matches_per_view_oracle(ctx, n, shuffle):
    rng = np.random.default_rng(21 + n)
    model = scenes.random_transforms(rng, n, 400.0, 60.0)
    layers = rng.integers(1, 16, size=n).astype(np.uint32)
    layers[rng.random(n) < 0.05] = 0
    flags = np.full(n, 0xFFFF, np.uint16)
    flags[rng.random(n) < 0.1] &= 0xFFFC
    flags[rng.random(n) < 0.3] &= 0xFFFB  # no shadows
    ids = np.arange(n, dtype=np.uint32)
    if shuffle:
        rng.shuffle(ids)
    cuts = [0, n // 5, n // 5 + 37, n // 2, n]  # four materials, one of them tiny; the third has two commands
    mats = []
    for i in range(4):
        sl = ids[cuts[i]:cuts[i + 1]]
        lods = [(0.0, 40.0, 120.0), (math.inf,), (0.0, 80.0), (0.0, 30.0, 90.0, 250.0)][i]
        mesh = scenes.simple_mesh(scenes.CUBE_RADIUS if i % 2 == 0 else scenes.SPHERE_RADIUS, lods, base=i * 1000)
        if i == 2:
            h = len(sl) // 3
            mats.append(scenes.MaterialDraws([scenes.DrawCommand(mesh, sl[:h]), scenes.DrawCommand(scenes.simple_mesh(2.5, (math.inf,)), sl[h:])]))
        else:
            mats.append(scenes.MaterialDraws([scenes.DrawCommand(mesh, sl)]))
    t = host_tables.build_static_tables(mats)
    views = _views()
    if shuffle:
        views = views * 6  # 42 views: more than one kernel launch (32 views each) behind one call
    V = len(views)

    # per-view expectation: the oracle on a fresh copy of the single-view tables
    ubos, refs, pyrs_host = [], [], []
    for cam, (w, h), shadow, cam_layers in views:
        depth = scenes.synth_depth(w, h, cam, n_occluders=25, seed=w + h)
        pyr, dim, levels = oracle.pyramid_build(depth, min(w, h))
        ubo = make_ubo(cam.view_proj_cm, cam.pos, t.num_cubos, layers=cam_layers, shadow=shadow)
        refs.append(oracle_cull(t, ubo, model, layers, flags, pyr, dim, levels, nthreads=4))
        ubos.append(ubo)
        pyrs_host.append((depth, min(w, h), pyr))

    # device: outputs replicated V times, view v displaced by v * (material numLODs, len(culling buffer))
    out_handles = {}
    for c in t.cubos:
        out_handles[int(c["idOutputBufferBindlessHandle"])] = "cull"
        out_handles[int(c["indirectOutputBufferBindlessHandle"])] = "ind"
    dev = []
    for i, b in enumerate(t.bindless):
        if i in out_handles:
            dev.append(ctx.buffer_from_numpy(np.concatenate([b.reshape(-1)] * V)))
        else:
            dev.append(ctx.buffer_from_numpy(b))
    cubos = t.cubos.copy()
    for f in ("idOutputBufferBindlessHandle", "entityIDInputBufferBindlessHandle", "indirectOutputBufferBindlessHandle",
              "lodDistanceBufferBindlessHandle"):
        cubos[f] = np.array([dev[int(h)].bindless for h in t.cubos[f]], dtype=np.uint32)
    strides = np.array([[t.bindless[int(c["indirectOutputBufferBindlessHandle"])].shape[0],
                         t.bindless[int(c["idOutputBufferBindlessHandle"])].size] for c in t.cubos], np.uint32)
    d_cubos = ctx.buffer_from_numpy(cubos.view(np.uint8))
    d_model, d_layers, d_flags = ctx.buffer_from_numpy(model), ctx.buffer_from_numpy(layers), ctx.buffer_from_numpy(flags)
    pyrs = []
    for depth, width, pyr_ref in pyrs_host:
        p = ctx.create_pyramid(width)
        ctx.build_pyramid(p, ctx.buffer_from_numpy(depth), depth.shape[1], depth.shape[0])
        pyrs.append(p)
    cull_views(ctx, ubos, pyrs, strides, d_cubos, d_model, d_layers, d_flags, (t.total_entities + 63) // 64)
    ctx.synchronize()
    for p, (_, _, pyr_ref) in zip(pyrs, pyrs_host):
        np.testing.assert_array_equal(p.texels.download(np.float32), pyr_ref)
    total_visible = 0
    for v in range(V):
        got = []
        for i, b in enumerate(t.bindless):
            if b is None:
                got.append(None)
                continue
            raw = dev[i].download(np.uint8).view(b.dtype)
            got.append(raw[v * b.size:(v + 1) * b.size].reshape(b.shape) if i in out_handles else raw[: b.size].reshape(b.shape))
        compare_cull_outputs(t, mats, refs[v], got)
        total_visible += sum(int(b.reshape(-1, 5)[:, 1].sum()) for i, b in enumerate(got) if out_handles.get(i) == "ind")
    assert total_visible > 0


def test_cull_views_single_view_equals_rve_cull(ctx):
    """one view through the batched entry point == rve_cull"""
    from tests.helpers import device_cull

    rng = np.random.default_rng(3)
    n = 30_000
    model = scenes.random_transforms(rng, n, 300.0, 50.0)
    layers = np.full(n, 0xFFFFFFFF, np.uint32)
    flags = np.full(n, 0xFFFF, np.uint16)
    mesh = scenes.simple_mesh(scenes.CUBE_RADIUS, (0.0, 40.0, 120.0))
    mats = [scenes.MaterialDraws([scenes.DrawCommand(mesh, np.arange(n, dtype=np.uint32))])]
    t = host_tables.build_static_tables(mats)
    cam = scenes.default_camera(z_far=1000.0)
    depth = scenes.synth_depth(640, 360, cam, n_occluders=30)
    ubo = make_ubo(cam.view_proj_cm, cam.pos, t.num_cubos)
    want, _ = device_cull(ctx, t, ubo, model, layers, flags, depth=depth, pyr_width=360)
    dev = [ctx.buffer_from_numpy(b) for b in t.bindless]
    cubos = t.cubos.copy()
    for f in ("idOutputBufferBindlessHandle", "entityIDInputBufferBindlessHandle", "indirectOutputBufferBindlessHandle",
              "lodDistanceBufferBindlessHandle"):
        cubos[f] = np.array([dev[int(h)].bindless for h in t.cubos[f]], dtype=np.uint32)
    p = ctx.create_pyramid(360)
    ctx.build_pyramid(p, ctx.buffer_from_numpy(depth), 640, 360)
    cull_views(ctx, [ubo], [p], np.array([[3, 3 * n]], np.uint32), ctx.buffer_from_numpy(cubos.view(np.uint8)), ctx.buffer_from_numpy(model),
               ctx.buffer_from_numpy(layers), ctx.buffer_from_numpy(flags), (n + 63) // 64)
    ctx.synchronize()
    got = [b.download(np.uint8).view(h.dtype)[: h.size].reshape(h.shape) for b, h in zip(dev, t.bindless)]
    compare_cull_outputs(t, mats, want, got)


def test_cascades_feed_the_view_batch_without_a_host_round_trip(ctx):
    """rve_csm_cascades -> rve_cull_views_matrices: the shadow cascades' lightViewProj stay in device memory and drive the
    view-batched cull (planes derived on the device). Must equal the same batch driven by host copies of the matrices,
    which in turn is checked against the oracle per view."""
    from ravengine_b200.api import CSM_CAMERA, CSM_LIGHT, csm_camera_params, csm_cascades, cull_views_matrices
    from tests.golden import make_golden_csm as mg

    rng = np.random.default_rng(77)
    n = 40_000
    model = scenes.random_transforms(rng, n, 150.0, 40.0)
    layers = np.full(n, 0xFFFFFFFF, np.uint32)
    flags = np.full(n, 0xFFFF, np.uint16)
    flags[rng.random(n) < 0.2] &= 0xFFFB  # not every instance casts shadows
    ids = np.arange(n, dtype=np.uint32)
    mats = [scenes.MaterialDraws([scenes.DrawCommand(scenes.simple_mesh(scenes.CUBE_RADIUS, (0.0, 60.0, 200.0)), ids[: n // 2])]),
            scenes.MaterialDraws([scenes.DrawCommand(scenes.simple_mesh(scenes.SPHERE_RADIUS, (math.inf,), base=500), ids[n // 2:])])]
    t = host_tables.build_static_tables(mats)

    # cascades of one camera x two lights on the device
    cam = scenes.default_camera(z_near=0.5, z_far=400.0, eye=(0, 20, 120), target=(0, 0, 0))
    cams = np.zeros(1, CSM_CAMERA)
    cams[0]["viewOnly"] = np.asarray(cam.view_cm if hasattr(cam, "view_cm") else mg.random_view(rng), np.float32)
    cams[0]["zNear"], cams[0]["zFar"] = 0.5, 400.0
    cams[0]["tanHalfFovy"], cams[0]["aspect"] = csm_camera_params(60.0, 1920, 1080)
    lights = np.zeros(2, CSM_LIGHT)
    lights["direction"] = np.array([[0.3, -0.9, 0.2], [-0.5, -0.7, -0.4]], np.float32) / np.linalg.norm([[0.3, -0.9, 0.2], [-0.5, -0.7, -0.4]], axis=1, keepdims=True).astype(np.float32)
    lights["castsShadows"] = 1
    lights["numCascades"] = [4, 2]
    lights["shadowCascades"] = np.array([[0.05, 0.15, 0.4, 1.0], [0.1, 0.5, 1.0, 1.0]], np.float32)
    lights["maxShadowDistanceFactor"] = 1.0
    d_vp = ctx.create_buffer(2 * 4 * 64)
    d_vp.zero()
    csm_cascades(ctx, ctx.buffer_from_numpy(cams.view(np.uint8)), 1, ctx.buffer_from_numpy(lights.view(np.uint8)), 2, d_vp)
    ctx.synchronize()
    vp_host = d_vp.download(np.float32).reshape(2, 4, 16)
    index = [0, 1, 2, 3, 4, 5]  # light 0: cascades 0..3 (matrices 0..3), light 1: cascades 0..1 (matrices 4, 5)
    assert all(np.isfinite(vp_host.reshape(8, 16)[i]).all() and np.any(vp_host.reshape(8, 16)[i]) for i in index)
    V = len(index)

    empty_dim, empty_levels = oracle.pyramid_dims(64)
    empty = np.zeros((4 * empty_dim * empty_dim - 1) // 3, np.float32)
    ubos, refs = [], []
    for i in index:
        ubo = make_ubo(vp_host.reshape(8, 16)[i], cam.pos, t.num_cubos, shadow=True)
        refs.append(oracle_cull(t, ubo, model, layers, flags, empty, empty_dim, empty_levels, nthreads=4))
        ubos.append(ubo)

    def run(device_matrices: bool):
        out_handles = {}
        for c in t.cubos:
            out_handles[int(c["idOutputBufferBindlessHandle"])] = "cull"
            out_handles[int(c["indirectOutputBufferBindlessHandle"])] = "ind"
        dev = [ctx.buffer_from_numpy(np.concatenate([b.reshape(-1)] * V)) if i in out_handles else ctx.buffer_from_numpy(b)
               for i, b in enumerate(t.bindless)]
        cubos = t.cubos.copy()
        for f in ("idOutputBufferBindlessHandle", "entityIDInputBufferBindlessHandle", "indirectOutputBufferBindlessHandle",
                  "lodDistanceBufferBindlessHandle"):
            cubos[f] = np.array([dev[int(h)].bindless for h in t.cubos[f]], dtype=np.uint32)
        strides = np.array([[t.bindless[int(c["indirectOutputBufferBindlessHandle"])].shape[0],
                             t.bindless[int(c["idOutputBufferBindlessHandle"])].size] for c in t.cubos], np.uint32)
        args = (ctx.buffer_from_numpy(cubos.view(np.uint8)), ctx.buffer_from_numpy(model), ctx.buffer_from_numpy(layers), ctx.buffer_from_numpy(flags),
                (t.total_entities + 63) // 64)
        pyrs = [ctx.create_pyramid(64)] * V
        if device_matrices:
            blank = [make_ubo(np.zeros(16, np.float32), cam.pos, t.num_cubos, shadow=True) for _ in index]  # viewProj must be ignored
            cull_views_matrices(ctx, blank, d_vp, index, pyrs, strides, *args)
        else:
            cull_views(ctx, ubos, pyrs, strides, *args)
        ctx.synchronize()
        per_view = []
        for v in range(V):
            got = []
            for i, b in enumerate(t.bindless):
                if b is None:
                    got.append(None)
                    continue
                raw = dev[i].download(np.uint8).view(b.dtype)
                got.append(raw[v * b.size:(v + 1) * b.size].reshape(b.shape) if i in out_handles else raw[: b.size].reshape(b.shape))
            per_view.append(got)
        return per_view

    visible = 0
    for device_matrices in (False, True):
        for v, got in enumerate(run(device_matrices)):
            compare_cull_outputs(t, mats, refs[v], got)
            visible += sum(int(b.reshape(-1, 5)[:, 1].sum()) for b in got if b is not None and b.ndim == 2 and b.shape[1] == 5)
    assert visible > 0
