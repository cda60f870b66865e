"""GPU parity: rve_cull / rve_depth_pyramid_build (through the C-ABI) vs the CPU oracle.
Bit-exact counts and commands; id slices equal after sorting (append order = atomic arrival)."""
import math

import numpy as np
import pytest

import oracle
from oracle import host_tables
from ravengine_b200 import scenes
from tests.helpers import compare_cull_outputs, device_cull, footmin_chain, make_ubo, oracle_cull

pytestmark = pytest.mark.gpu


def _scene(n, seed, n_mats=2, lods=(math.inf,), extent=200.0, extent_y=200.0, flags_val=0xFFFD, shuffle=False,
           frac_layer0=0.0, frac_nocull=0.0):
    rng = np.random.default_rng(seed)
    model = scenes.random_transforms(rng, n, extent, extent_y)
    layers = np.full(n, 0xFFFFFFFF, np.uint32)
    flags = np.full(n, flags_val, np.uint16)
    layers[rng.random(n) < frac_layer0] = 0
    flags[rng.random(n) < frac_nocull] &= 0xFFFC
    ids = np.arange(n, dtype=np.uint32)
    if shuffle:
        rng.shuffle(ids)
    mats = []
    bounds = np.linspace(0, n, n_mats + 1).astype(int)
    for i in range(n_mats):
        mesh = scenes.simple_mesh(scenes.CUBE_RADIUS if i % 2 == 0 else scenes.SPHERE_RADIUS, lods, base=i * 10000)
        mats.append(scenes.MaterialDraws([scenes.DrawCommand(mesh, ids[bounds[i]:bounds[i + 1]])]))
    return model, layers, flags, mats


def _run_both(ctx, model, layers, flags, mats, cam, depth=None, pyr_width=1024, shadow=False, cam_layers=0xFFFFFFFF,
              groups=None, nthreads=4, deferred=False):
    t = host_tables.build_static_tables(mats)
    ubo = make_ubo(cam.view_proj_cm, cam.pos, t.num_cubos, layers=cam_layers, shadow=shadow)
    if depth is not None:
        pyr, dim, levels = oracle.pyramid_build(depth, pyr_width)
    else:
        dim, levels = oracle.pyramid_dims(pyr_width)
        pyr = np.zeros((4 * dim * dim - 1) // 3, np.float32)
    ref = oracle_cull(t, ubo, model, layers, flags, pyr, dim, levels, groups=groups, nthreads=nthreads)
    got, pyr_dev = device_cull(ctx, t, ubo, model, layers, flags, depth=depth, pyr_width=pyr_width, groups=groups, deferred=deferred)
    np.testing.assert_array_equal(pyr, pyr_dev)
    compare_cull_outputs(t, mats, ref, got)
    return t, ref


def test_c1_frustum_lod_100k(ctx):
    model, layers, flags, mats = _scene(100_000, 1)
    t, ref = _run_both(ctx, model, layers, flags, mats, scenes.default_camera())
    vis = sum(int(b.reshape(-1, 5)[:, 1].sum()) for b in ref if b is not None and b.dtype == np.uint32 and b.ndim == 2)
    assert 0 < vis < 100_000


def test_c1_shuffled_ids(ctx):
    model, layers, flags, mats = _scene(50_000, 2, shuffle=True)
    _run_both(ctx, model, layers, flags, mats, scenes.default_camera())


@pytest.mark.parametrize("n", [0, 1, 31, 32, 33, 63, 64, 65, 1000, 4097])
def test_ragged_sizes(ctx, n):
    model, layers, flags, mats = _scene(max(n, 1), 3 + n, n_mats=1)
    if n == 0:
        mats[0].commands[0].entity_ids = np.zeros(0, np.uint32)
    _run_both(ctx, model, layers, flags, mats, scenes.default_camera())


def test_c2_small_hiz_3lods(ctx):
    cam = scenes.default_camera(z_far=1000.0)
    depth = scenes.synth_depth(960, 540, cam)
    model, layers, flags, mats = _scene(200_000, 5, n_mats=4, lods=(0.0, 40.0, 120.0), extent=500.0, extent_y=50.0,
                                        flags_val=0xFFFF, frac_layer0=0.05, frac_nocull=0.05)
    t, ref = _run_both(ctx, model, layers, flags, mats, cam, depth=depth, pyr_width=540)
    counts = np.concatenate([b.reshape(-1, 5)[:, 1] for b in ref if b is not None and b.dtype == np.uint32 and b.ndim == 2])
    assert (counts > 0).sum() >= 8  # several LOD slots populated


def test_c2_fullres_pyramid_hiz(ctx):
    cam = scenes.default_camera(z_far=1000.0)
    depth = scenes.synth_depth(3840, 2160, cam)
    model, layers, flags, mats = _scene(300_000, 6, n_mats=4, lods=(0.0, 40.0, 120.0), extent=500.0, extent_y=50.0,
                                        flags_val=0xFFFF, frac_layer0=0.05, frac_nocull=0.05, shuffle=True)
    _run_both(ctx, model, layers, flags, mats, cam, depth=depth, pyr_width=2160)


def test_shadow_mode_and_layers(ctx):
    model, layers, flags, mats = _scene(20_000, 7, flags_val=0xFFFF)
    rng = np.random.default_rng(70)
    flags[rng.random(len(flags)) < 0.5] &= 0xFFFB  # clear casts-shadows
    layers[:] = rng.integers(0, 16, size=len(layers)).astype(np.uint32)
    _run_both(ctx, model, layers, flags, mats, scenes.default_camera(), shadow=True, cam_layers=0b0101)


def test_many_small_cubos(ctx):
    """CUBO boundaries inside warps: 300 commands of 1..40 entities, mixed LOD counts, one empty command."""
    rng = np.random.default_rng(8)
    sizes = rng.integers(1, 41, size=300)
    sizes[17] = 0
    n = int(sizes.sum())
    model = scenes.random_transforms(rng, n, 60, 60)
    layers = np.full(n, 0xFFFFFFFF, np.uint32)
    flags = np.full(n, 0xFFFF, np.uint16)
    ids = rng.permutation(n).astype(np.uint32)
    mats, off = [], 0
    for i, s in enumerate(sizes):
        lods = [(math.inf,), (0.0, 20.0), (0.0, 15.0, 45.0, 90.0)][i % 3]
        cmd = scenes.DrawCommand(scenes.simple_mesh(1.0 + (i % 5) * 0.3, lods), ids[off:off + s])
        off += s
        if i % 4 == 0 or not mats:
            mats.append(scenes.MaterialDraws([cmd]))
        else:
            mats[-1].commands.append(cmd)
    cam = scenes.default_camera()
    depth = scenes.synth_depth(512, 512, cam, n_occluders=30)
    _run_both(ctx, model, layers, flags, mats, cam, depth=depth, pyr_width=512)


def test_dispatch_smaller_than_total(ctx):
    """groups*64 < sum(numObjects): invocations past the dispatch never run (skinned-cull quirk, :1214)."""
    model, layers, flags, mats = _scene(5000, 9, n_mats=2)
    _run_both(ctx, model, layers, flags, mats, scenes.default_camera(), groups=40)


def test_zero_pyramid_occludes_nothing(ctx):
    """First frame: pyramid all zeros = far plane in reverse-Z, so Hi-Z never rejects."""
    model, layers, flags, mats = _scene(30_000, 10, flags_val=0xFFFF)
    cam = scenes.default_camera()
    t = host_tables.build_static_tables(mats)
    ubo = make_ubo(cam.view_proj_cm, cam.pos, t.num_cubos)
    got_occ, _ = device_cull(ctx, t, ubo, model, layers, flags)
    flags2 = flags & 0xFFFD
    got_nocc, _ = device_cull(ctx, t, ubo, model, layers, flags2)
    compare_cull_outputs(t, mats, got_occ, got_nocc)


@pytest.mark.parametrize("size", [(1, 1), (2, 2), (5, 3), (64, 64), (100, 37), (1920, 1080), (3840, 2160), (4096, 4096)])
def test_pyramid_sizes(ctx, size):
    w, h = size
    rng = np.random.default_rng(w * 7 + h)
    depth = rng.random((h, w), dtype=np.float32)
    ref, dim, levels = oracle.pyramid_build(depth)
    pyr = ctx.create_pyramid(min(w, h))
    assert (pyr.dim, pyr.levels) == (dim, levels)
    d = ctx.buffer_from_numpy(depth)
    for _ in range(2):  # second build checks the ticket counter re-arms
        ctx.build_pyramid(pyr, d, w, h)
    ctx.synchronize()
    np.testing.assert_array_equal(ref, pyr.texels.download(np.float32))


def test_broken_id_runs_mix_all_staging_paths(ctx):
    """Entity-id lists made of dense runs of 1..300 ids at arbitrary (unaligned) starts, in shuffled run order, over a
    sparse entity space: one dispatch mixes 64-row TMA rounds, single dense halves, gathers, CUBO boundaries inside
    rounds and a partial last round (csrc/cull.cu issue_round / load_round_ids)."""
    rng = np.random.default_rng(11)
    n_entities = 120_000
    model = scenes.random_transforms(rng, n_entities, 300.0, 300.0)
    layers = np.full(n_entities, 0xFFFFFFFF, np.uint32)
    flags = np.full(n_entities, 0xFFFF, np.uint16)
    layers[rng.random(n_entities) < 0.05] = 0
    flags[rng.random(n_entities) < 0.1] &= 0xFFFC
    runs, pos = [], 3  # entity 0..2 unused: the first run starts unaligned
    while pos < n_entities - 400:
        ln = int(rng.integers(1, 301))
        runs.append(np.arange(pos, pos + ln, dtype=np.uint32))
        pos += ln + int(rng.integers(0, 5))  # holes in the entity space
    order = rng.permutation(len(runs))
    cut = [0, len(runs) // 3, len(runs) // 3 + 1, 2 * len(runs) // 3, len(runs)]
    mats = []
    for i in range(4):
        ids = np.concatenate([runs[j] for j in order[cut[i]:cut[i + 1]]])
        mesh = scenes.simple_mesh(scenes.CUBE_RADIUS, (0.0, 60.0, 150.0) if i % 2 == 0 else (math.inf,), base=i * 1000)
        mats.append(scenes.MaterialDraws([scenes.DrawCommand(mesh, ids)]))
    cam = scenes.default_camera(z_far=1000.0)
    depth = scenes.synth_depth(640, 360, cam, n_occluders=30)
    _run_both(ctx, model, layers, flags, mats, cam, depth=depth, pyr_width=360)


def test_c2_full_size_10m_matches_oracle(ctx):
    """BASELINE config C2 at its full size: 10 M instances, 4 materials x 3 LODs, Hi-Z against the 3840x2160 pyramid,
    scene generated by the shared (seed, index) generator; counts and sorted id slices against the oracle."""
    from ravengine_b200 import SceneGenParams

    n = 10_000_000
    params = SceneGenParams(20240, 500.0, 50.0, 0.5, 2.0, 0.05, 0.05, 0)
    model = np.zeros((n, 16), np.float32)
    layers, flags, ids = np.zeros(n, np.uint32), np.zeros(n, np.uint16), np.zeros(n, np.uint32)
    oracle.gen_instances(params, 0, n, model, layers, flags, ids)
    mats = []
    for i in range(4):
        mesh = scenes.simple_mesh(1.7320508 if i % 2 == 0 else 1.74, (0.0, 40.0, 120.0))
        mats.append(scenes.MaterialDraws([scenes.DrawCommand(mesh, ids[n * i // 4: n * (i + 1) // 4])]))
    cam = scenes.default_camera(z_near=0.1, z_far=1000.0)
    depth = scenes.synth_depth(3840, 2160, cam)
    t, ref = _run_both(ctx, model, layers, flags, mats, cam, depth=depth, pyr_width=2160, nthreads=oracle.default_workers())
    counts = np.concatenate([b.reshape(-1, 5)[:, 1] for b in ref if b is not None and b.dtype == np.uint32 and b.ndim == 2])
    assert int(counts.sum()) == 2405721  # the bench's visible count for this seed


@pytest.mark.parametrize("n", [909_313, 1_000_003])
def test_dynamic_tail_rounds_mixed_paths(ctx, n):
    """Launches just past the threshold where the cull kernel hands its last rounds out through a global counter
    (rounds > 4 x resident warps, csrc/cull.cu next_issue): ragged total, broken id runs in the part of the dispatch the
    tail covers, a dispatch shorter than the CUBO sum. Every round must be processed exactly once."""
    rng = np.random.default_rng(n)
    model = scenes.random_transforms(rng, n, 400.0, 80.0)
    layers = np.full(n, 0xFFFFFFFF, np.uint32)
    layers[rng.random(n) < 0.03] = 0
    flags = np.full(n, 0xFFFF, np.uint16)
    flags[rng.random(n) < 0.1] &= 0xFFFC
    ids = np.arange(n, dtype=np.uint32)
    tail = n - 60_000  # the last 60k invocations: reversed blocks of 50 ids -> no 64-id run, gathers and dense halves
    blocks = [ids[i:i + 50] for i in range(tail, n, 50)]
    ids[tail:] = np.concatenate(blocks[::-1])
    cuts = [0, n // 3, n // 3 + 11, n]
    mats = []
    for i in range(3):
        mesh = scenes.simple_mesh(scenes.CUBE_RADIUS, (0.0, 50.0, 140.0) if i != 1 else (math.inf,), base=i * 500)
        mats.append(scenes.MaterialDraws([scenes.DrawCommand(mesh, ids[cuts[i]:cuts[i + 1]])]))
    cam = scenes.default_camera(z_far=1000.0)
    depth = scenes.synth_depth(640, 360, cam, n_occluders=25)
    t, ref = _run_both(ctx, model, layers, flags, mats, cam, depth=depth, pyr_width=360, nthreads=oracle.default_workers())
    if n == 1_000_003:  # and with the dispatch cut short (defaultcull.csh:166-171 bounds by the dispatch as well)
        _run_both(ctx, model, layers, flags, mats, cam, depth=depth, pyr_width=360, groups=(n - 70_000) // 64, nthreads=oracle.default_workers())


@pytest.mark.parametrize("size", [(3840, 2160), (1920, 1080), (1280, 720), (2560, 1440), (7680, 4320), (1024, 1024), (1028, 4000), (36, 32), (64, 64), (200, 130)])
def test_pyramid_build_paths_equal_each_other_and_the_oracle(ctx, size):
    """The three builds - one persistent launch (TMA ring, register / shuffle tiles, grid barrier), two launches with one
    tensor copy per CTA, two launches with LSU-fed tiles - against the oracle's texels and the footprint-min chain derived
    from them, at sizes with different source / pyramid ratios (1x .. 7.5x) and edge clamping on both axes. Three builds
    in a row per path: the grid barrier re-arms itself."""
    import os

    w, h = size
    rng = np.random.default_rng(w + 3 * h)
    depth = rng.random((h, w), dtype=np.float32)
    ref, dim, levels = oracle.pyramid_build(depth)
    fm_ref = footmin_chain(ref, dim, levels)
    d = ctx.buffer_from_numpy(depth)
    junk = ctx.buffer_from_numpy(rng.random((h, w), dtype=np.float32))
    saved = {k: os.environ.get(k) for k in ("RVE_PYRAMID_NO_TMA", "RVE_PYRAMID_PERSISTENT")}
    try:
        for env in ({"RVE_PYRAMID_PERSISTENT": "1"}, {}, {"RVE_PYRAMID_NO_TMA": "1"}):
            for k in saved:
                os.environ.pop(k, None)
            os.environ.update(env)
            pyr = ctx.create_pyramid(min(w, h))
            ctx.build_pyramid(pyr, junk, w, h)  # other contents first: every texel and chain entry must be rewritten
            for _ in range(2):
                ctx.build_pyramid(pyr, d, w, h)
            ctx.synchronize()
            np.testing.assert_array_equal(ref, pyr.texels.download(np.float32), err_msg=str(env))
            np.testing.assert_array_equal(fm_ref, ctx.pyramid_footmin(pyr).download(np.float32)[: fm_ref.size], err_msg=str(env))
    finally:
        for k, v in saved.items():
            os.environ.pop(k, None)
            if v is not None:
                os.environ[k] = v


def test_garbage_in_the_unwritten_cubo_tail_is_never_followed(ctx):
    """The reference writes one CUBO per command but sets numCubos to the LOD count (src/RenderEngine_Draw.cpp:1238): the
    tail of its cuboBuffer is recycled staging memory. rve_cull must give the same result with garbage there - stale
    handles past the table, handles of destroyed buffers, huge numObjects past the dispatch - and must drop (and count) a
    selectable entry whose handles are dead instead of following them."""
    model, layers, flags, mats = _scene(30_080, 3, n_mats=2, lods=(0.0, 40.0, 120.0))  # a multiple of 64: the dispatch ends where the entities end
    cam = scenes.default_camera()
    t = host_tables.build_static_tables(mats)
    assert t.num_cubos == 6 and len(t.cubos) == 6  # 2 written entries, 4 zero tail entries in the oracle's table
    ubo = make_ubo(cam.view_proj_cm, cam.pos, t.num_cubos)
    depth = scenes.synth_depth(320, 180, cam, n_occluders=10)
    pyr, dim, levels = oracle.pyramid_build(depth, 180)
    ref = oracle_cull(t, ubo, model, layers, flags, pyr, dim, levels, nthreads=2)
    import dataclasses

    dirty = t.cubos.copy()
    rng = np.random.default_rng(0)
    for i in range(2, 6):  # never selectable: empty, or starting past the dispatch (= total entities)
        dirty[i]["numObjects"] = 0 if i == 2 else int(rng.integers(1, 1 << 30))
        dirty[i]["numLODs"] = int(rng.integers(2, 5))
        dirty[i]["radius"] = float(rng.random())
        for j, f in enumerate(("idOutputBufferBindlessHandle", "entityIDInputBufferBindlessHandle", "indirectOutputBufferBindlessHandle",
                               "lodDistanceBufferBindlessHandle")):
            dirty[i][f] = [0xFFFFFFF0, 123456, 0x00FFFFFF, 0x7FFFFFFF][(i + j) % 4]
    before = ctx.lib.rve_ctx_bad_cubo_handles(ctx.h)
    got, _ = device_cull(ctx, dataclasses.replace(t, cubos=dirty), ubo, model, layers, flags, depth=depth, pyr_width=180)
    compare_cull_outputs(t, mats, ref, got)
    assert ctx.lib.rve_ctx_bad_cubo_handles(ctx.h) == before  # nothing selectable was dropped
    # a SELECTABLE entry with a dead handle: dropped and counted; the other command's result is unchanged
    broken = t.cubos.copy()
    broken[1]["lodDistanceBufferBindlessHandle"] = 0xFFFFFF00
    got3, _ = device_cull(ctx, dataclasses.replace(t, cubos=broken), ubo, model, layers, flags, depth=depth, pyr_width=180)
    assert ctx.lib.rve_ctx_bad_cubo_handles(ctx.h) == before + 1
    h0 = int(t.cubos[0]["indirectOutputBufferBindlessHandle"])
    np.testing.assert_array_equal(got3[h0], ref[h0])
    h1 = int(t.cubos[1]["indirectOutputBufferBindlessHandle"])
    assert not got3[h1][:, 1].any()  # the broken command drew nothing


@pytest.mark.parametrize("n,size", [(40, (640, 360)), (5000, (640, 360)), (300_000, (640, 360)), (300_000, (3840, 2160)), (2_000_000, (1920, 1080)),
                                    (5000, (36, 32))])
def test_deferred_pyramid_build_is_finished_by_the_cull_that_follows(ctx, n, size):
    """rve_depth_pyramid_build_deferred + rve_cull: the build stops after the tile levels, the cull kernel derives mips >= 6 and
    footprint-min levels >= 1 in its start-up (one worker CTA for the smallest dispatch, 3 x 148 for the largest) - same
    pyramid, same cull as the oracle's."""
    model, layers, flags, mats = _scene(n, 17, n_mats=2, lods=(0.0, 40.0, 120.0), extent=400.0, extent_y=60.0, flags_val=0xFFFF, frac_nocull=0.1)
    cam = scenes.default_camera(z_far=1000.0)
    depth = scenes.synth_depth(size[0], size[1], cam, n_occluders=30)
    _run_both(ctx, model, layers, flags, mats, cam, depth=depth, pyr_width=min(size), nthreads=oracle.default_workers(), deferred=True)


def test_deferred_pyramid_build_is_finished_for_every_other_consumer(ctx):
    """A deferring build followed by something that is not a single-view cull: the footprint-min accessor and a view batch each
    complete the pyramid with a launch of their own first (pyramid_finish_upper); a second build supersedes a pending one."""
    from ravengine_b200.world import World

    rng = np.random.default_rng(5)
    w_, h_ = 1280, 720
    cam = scenes.default_camera(z_far=1000.0)
    depth_a, depth_b = rng.random((h_, w_), dtype=np.float32), scenes.synth_depth(w_, h_, cam, n_occluders=25)
    d_a, d_b = ctx.buffer_from_numpy(depth_a), ctx.buffer_from_numpy(depth_b)
    pyr = ctx.create_pyramid(h_)
    ref_a, dim, levels = oracle.pyramid_build(depth_a, h_)
    ref_b, _, _ = oracle.pyramid_build(depth_b, h_)
    # a download of the texel buffer through the library
    ctx.build_pyramid(pyr, d_a, w_, h_, deferred=True)
    np.testing.assert_array_equal(ref_a, pyr.texels.download(np.float32))
    # accessor
    ctx.build_pyramid(pyr, d_a, w_, h_, deferred=True)
    fm = ctx.pyramid_footmin(pyr).download(np.float32)
    np.testing.assert_array_equal(footmin_chain(ref_a, dim, levels), fm[: footmin_chain(ref_a, dim, levels).size])
    np.testing.assert_array_equal(ref_a, pyr.texels.download(np.float32))
    # pending build superseded by another deferring build, then a view batch
    ctx.build_pyramid(pyr, d_a, w_, h_, deferred=True)
    ctx.build_pyramid(pyr, d_b, w_, h_, deferred=True)
    n = 20_000
    model, layers, flags, mats = _scene(n, 23, n_mats=2, lods=(0.0, 40.0, 120.0), extent=400.0, extent_y=60.0, flags_val=0xFFFF)
    outs = []
    for deferred in (True, False):
        w = World(ctx)
        bufs = [ctx.buffer_from_numpy(model), ctx.buffer_from_numpy(layers), ctx.buffer_from_numpy(flags)]
        w.set_entity_buffers(*bufs)
        mids = []
        for m in mats:
            mid = w.material()
            for c in m.commands:
                k = len(c.mesh.lod_distances)
                mesh = w.mesh_collection(c.mesh.radius, c.mesh.lod_distances, [36, 24, 12][:k], [0, 36, 60][:k], [0, 24, 40][:k])
                ids = ctx.buffer_from_numpy(c.entity_ids)
                bufs.append(ids)
                w.add_command(mid, mesh, ids, len(c.entity_ids))
            mids.append(mid)
        if not deferred:
            ctx.build_pyramid(pyr, d_b, w_, h_)
        views = [(cam.view_proj_cm, cam.pos, 0xFFFFFFFF, False)] * 2
        w.cull_static_views(views, [pyr, pyr])
        ctx.synchronize()
        outs.append([(w.material_outputs(m)[1].download(np.uint32), np.sort(w.material_outputs(m)[0].download(np.uint32))) for m in mids])
        np.testing.assert_array_equal(ref_b, pyr.texels.download(np.float32))
        w.destroy()
    for (i0, s0), (i1, s1) in zip(*outs):
        np.testing.assert_array_equal(i0, i1)
        np.testing.assert_array_equal(s0, s1)
