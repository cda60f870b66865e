"""GPU tests of the C++ shim (include/rve_shim.h) — the mirror of RenderEngine::Draw's encode blocks — against the
oracle's restatement of the same host tables (oracle/host_tables.py) and the oracle kernels. The shim is driven
exactly like the engine drives its lambdas: cull_static per pass; prepare -> pose -> prepare-culling -> cull for
skinned meshes; bin_lights per lit view."""
import math

import numpy as np
import pytest

import oracle
from oracle import host_tables
from oracle.host_tables import CUBO_DTYPE, SkinnedCommandDesc, SkinnedMeshDesc
from ravengine_b200 import GridAssignUBO, GridBuildUBO, SkinningPrepareUBO, SkinningUBO, scenes
from ravengine_b200.world import MAT_LIT_OPAQUE, MAT_UNLIT_OPAQUE, QUIRK_INDIRECT_SLOT, World
from tests.helpers import footmin_chain, make_ubo, oracle_cull

pytestmark = pytest.mark.gpu


def _static_scene(seed=3, n=6000):
    rng = np.random.default_rng(seed)
    model = scenes.random_transforms(rng, n, 120.0, 60.0)
    layers = np.full(n, 0xFFFFFFFF, np.uint32)
    flags = np.full(n, 0xFFFF, np.uint16)
    flags[rng.random(n) < 0.1] &= 0xFFFB
    ids = rng.permutation(n).astype(np.uint32)
    a = scenes.simple_mesh(1.5, (0.0, 30.0, 90.0), base=0)
    b = scenes.simple_mesh(2.5, (math.inf,), base=7000)
    c = scenes.simple_mesh(0.7, (0.0, 50.0), base=9000)
    mats = [scenes.MaterialDraws([scenes.DrawCommand(a, ids[0:2500]), scenes.DrawCommand(b, ids[2500:3000])]),
            scenes.MaterialDraws([scenes.DrawCommand(c, ids[3000:5000]), scenes.DrawCommand(a, ids[5000:6000])])]
    return model, layers, flags, mats


def _world_from(ctx, model, layers, flags, mats, kinds=None, wflags=0):
    w = World(ctx, wflags)
    bufs = [ctx.buffer_from_numpy(model), ctx.buffer_from_numpy(layers), ctx.buffer_from_numpy(flags)]
    w.set_entity_buffers(*bufs)
    mesh_ids, mat_ids = {}, []
    for mi, m in enumerate(mats):
        mat = w.material(kinds[mi] if kinds else MAT_LIT_OPAQUE)
        mat_ids.append(mat)
        for c in m.commands:
            if id(c.mesh) not in mesh_ids:
                mesh_ids[id(c.mesh)] = w.mesh_collection(c.mesh.radius, c.mesh.lod_distances, [l.index_count for l in c.mesh.lods],
                                                         [l.index_start for l in c.mesh.lods], [l.base_vertex for l in c.mesh.lods])
            bufs.append(ctx.buffer_from_numpy(c.entity_ids) if len(c.entity_ids) else ctx.create_buffer(16))
            w.add_command(mat, mesh_ids[id(c.mesh)], bufs[-1], len(c.entity_ids))
    return w, mat_ids, bufs


def _compare_static(ctx, w, mat_ids, mats, t, ref, check_cubos=True):
    if check_cubos:
        cb_buf, num_cubos, total = w.last_cull_info()
        assert (num_cubos, total) == (t.num_cubos, t.total_entities)
        got = cb_buf.download(np.uint8)[: num_cubos * 36].view(CUBO_DTYPE)
        for f in ("indirectBufferOffset", "numObjects", "cullingBufferOffset", "radius", "numLODs"):
            np.testing.assert_array_equal(got[f], t.cubos[f], err_msg=f)
    ci = 0
    for mat, m in zip(mat_ids, mats):
        cull_b, ind_b, lods, ents = w.material_outputs(mat)
        cb = t.cubos[ci]
        r_ind, r_ids = ref[int(cb["indirectOutputBufferBindlessHandle"])], ref[int(cb["idOutputBufferBindlessHandle"])]
        g_ind = ind_b.download(np.uint32).reshape(-1, 5)[:lods]
        g_ids = cull_b.download(np.uint32)[: max(lods * ents, 1)]
        np.testing.assert_array_equal(g_ind, r_ind)
        rng_c = range(ci, ci + len(m.commands))
        for x, y in zip(host_tables.sorted_slices(r_ind, r_ids, t.cubos, rng_c), host_tables.sorted_slices(g_ind, g_ids, t.cubos, rng_c)):
            np.testing.assert_array_equal(x, y)
        ci += len(m.commands)


@pytest.mark.parametrize("shadow", [False, True])
def test_cull_static_matches_oracle_tables(ctx, shadow):
    model, layers, flags, mats = _static_scene()
    cam = scenes.default_camera(z_far=500.0)
    depth = scenes.synth_depth(512, 288, cam, n_occluders=30)
    pyr_h, dim, levels = oracle.pyramid_build(depth, 288)
    t = host_tables.build_static_tables(mats)
    ubo = make_ubo(cam.view_proj_cm, cam.pos, t.num_cubos, shadow=shadow)
    ref = oracle_cull(t, ubo, model, layers, flags, pyr_h, dim, levels, nthreads=4)
    w, mat_ids, keep = _world_from(ctx, model, layers, flags, mats)
    pyr = ctx.create_pyramid(288)
    ctx.build_pyramid(pyr, ctx.buffer_from_numpy(depth), 512, 288)
    for _ in range(2):  # second frame: same buffers, indirect counters re-initialised by the shim
        w.cull_static(cam.view_proj_cm, cam.pos, pyr, shadow=shadow)
    ctx.synchronize()
    _compare_static(ctx, w, mat_ids, mats, t, ref)
    w.destroy()


def test_cull_static_material_filter_and_growth(ctx):
    """filterRenderData: a pass only sees materials of its kind; entity lists growing between frames reallocate."""
    model, layers, flags, mats = _static_scene(seed=5)
    cam = scenes.default_camera(z_far=500.0)
    w, mat_ids, keep = _world_from(ctx, model, layers, flags, mats, kinds=[MAT_LIT_OPAQUE, MAT_UNLIT_OPAQUE])
    pyr = ctx.create_pyramid(64)  # all-zero pyramid: nothing occluded
    w.cull_static(cam.view_proj_cm, cam.pos, pyr, mask=MAT_UNLIT_OPAQUE)
    ctx.synchronize()
    t = host_tables.build_static_tables(mats[1:])
    ubo = make_ubo(cam.view_proj_cm, cam.pos, t.num_cubos)
    zero = np.zeros((4 * 64 * 64 - 1) // 3, np.float32)
    ref = oracle_cull(t, ubo, model, layers, flags, zero, 64, 7)
    _compare_static(ctx, w, mat_ids[1:], mats[1:], t, ref)
    assert w.material_outputs(mat_ids[0])[0] is None  # lit material untouched by the unlit pass
    # shrink command 0 of the unlit material and cull again
    mats[1].commands[0].entity_ids = mats[1].commands[0].entity_ids[:700]
    nb = ctx.buffer_from_numpy(mats[1].commands[0].entity_ids)
    w.set_command_entities(mat_ids[1], 0, nb, 700)
    w.cull_static(cam.view_proj_cm, cam.pos, pyr, mask=MAT_UNLIT_OPAQUE)
    ctx.synchronize()
    t = host_tables.build_static_tables(mats[1:])
    ref = oracle_cull(t, make_ubo(cam.view_proj_cm, cam.pos, t.num_cubos), model, layers, flags, zero, 64, 7)
    _compare_static(ctx, w, mat_ids[1:], mats[1:], t, ref)
    w.destroy()


def test_cull_static_reference_slot_quirk(ctx):
    """RVS_QUIRK_INDIRECT_SLOT reproduces the (meshID+lodID) indirect-init slot of :1314 byte for byte."""
    model, layers, flags, mats = _static_scene(seed=6)
    cam = scenes.default_camera(z_far=500.0)
    t = host_tables.build_static_tables(mats, quirk_slot=True)
    zero = np.zeros((4 * 64 * 64 - 1) // 3, np.float32)
    ref = oracle_cull(t, make_ubo(cam.view_proj_cm, cam.pos, t.num_cubos), model, layers, flags, zero, 64, 7)
    w, mat_ids, keep = _world_from(ctx, model, layers, flags, mats, wflags=QUIRK_INDIRECT_SLOT)
    w.cull_static(cam.view_proj_cm, cam.pos, ctx.create_pyramid(64))
    ctx.synchronize()
    _compare_static(ctx, w, mat_ids, mats, t, ref)
    w.destroy()


def test_skinned_path_matches_oracle(ctx):
    rng = np.random.default_rng(9)
    n_ent = 400
    model = scenes.random_transforms(rng, n_ent, 60.0, 20.0)
    layers = np.full(n_ent, 0xFFFFFFFF, np.uint32)
    flags = np.full(n_ent, 0xFFFF, np.uint16)
    # shared arena: mesh A at vertex 40 (100 verts), mesh B at vertex 400 (30 verts)
    arena_n = 512
    arena = [rng.random((arena_n, k), dtype=np.float32) for k in (3, 3, 3, 3, 2)]
    mesh_a = SkinnedMeshDesc(100, 300, 5000, 40, 1.0)
    mesh_b = SkinnedMeshDesc(30, 60, 9000, 400, 2.0)
    ids = rng.permutation(n_ent).astype(np.uint32)
    cmds = [[SkinnedCommandDesc(mesh_a, 8, ids[0:70]), SkinnedCommandDesc(mesh_b, 4, ids[70:200])],
            [SkinnedCommandDesc(mesh_b, 4, ids[200:330])]]
    wts = {id(mesh_a): scenes.skinning_inputs(rng, 100, 1, 8)["weights"], id(mesh_b): scenes.skinning_inputs(rng, 30, 1, 4)["weights"]}
    pals = [[scenes.skinning_inputs(rng, 1, len(c.entity_ids), c.num_bones)["pose"] for c in m] for m in cmds]
    st = host_tables.build_skinned_tables(cmds)

    # ---- oracle side ----
    total_verts = sum(c.mesh.num_verts * len(c.entity_ids) for m in cmds for c in m)
    outs_ref = [np.zeros((total_verts, k), np.float32) for k in (3, 3, 3, 3, 2)]
    pose_all = np.concatenate([p for m in pals for p in m])
    flat = [c for m in cmds for c in m]
    for c, u in zip(flat, st.skinning_ubos):
        oracle.skin(SkinningUBO(*u), outs_ref, arena, pose_all, wts[id(c.mesh)])
    cam = scenes.default_camera(z_far=300.0)
    zero = np.zeros((4 * 64 * 64 - 1) // 3, np.float32)
    ref_ind, ref_ids = [], []
    for mi, m in enumerate(cmds):
        cb0 = st.cubos[mi][0]
        bl = [None if b is None else b.copy() for b in st.bindless]
        ind = bl[int(cb0["indirectOutputBufferBindlessHandle"])]
        # prepareSkeletalCullingBuffer writes first, cullSkeletalMeshes then overwrites with its staging table
        for pu in st.prepare_ubos[mi]:
            oracle.skinned_drawcall_prepare(SkinningPrepareUBO(*pu), ind)
        ind[:] = st.indirect_reset[mi]
        ubo = make_ubo(cam.view_proj_cm, cam.pos, len(m), single=True)
        oracle.cull(ubo, st.cubos[mi], model, layers, flags, zero, 64, 7, bl, st.dispatch_groups[mi], 1)
        ref_ind.append(ind)
        ref_ids.append(bl[int(cb0["idOutputBufferBindlessHandle"])])

    # ---- shim side ----
    w = World(ctx)
    keep = [ctx.buffer_from_numpy(model), ctx.buffer_from_numpy(layers), ctx.buffer_from_numpy(flags)]
    w.set_entity_buffers(*keep)
    w.set_mesh_arena([ctx.buffer_from_numpy(a) for a in arena])
    mesh_h = {}
    for mdesc in (mesh_a, mesh_b):
        wb = ctx.buffer_from_numpy(wts[id(mdesc)])
        keep.append(wb)
        mesh_h[id(mdesc)] = w.skinned_mesh(mdesc.num_verts, mdesc.num_indices, mdesc.index_start, mdesc.vertex_start, mdesc.radius, wb)
    mat_h = []
    for mi, m in enumerate(cmds):
        mat = w.skinned_material()
        mat_h.append(mat)
        for ci, c in enumerate(m):
            idb = ctx.buffer_from_numpy(c.entity_ids)
            keep.append(idb)
            w.add_skinned_command(mat, mesh_h[id(c.mesh)], c.num_bones, idb, len(c.entity_ids), pals[mi][ci])
    assert w.prepare_skeletal_mesh_buffers()
    w.pose_skeletal_meshes()
    w.prepare_skeletal_culling_buffer()
    w.cull_skeletal(cam.view_proj_cm, cam.pos, ctx.create_pyramid(64))
    ctx.synchronize()
    outs, pose_buf = w.skinned_outputs()
    for b, r in zip(outs, outs_ref):
        np.testing.assert_array_equal(b.download(np.float32)[: r.size].reshape(r.shape), r)
    np.testing.assert_array_equal(pose_buf.download(np.float32)[: pose_all.size].reshape(pose_all.shape), pose_all)
    for mi, mat in enumerate(mat_h):
        cull_b, ind_b, ents = w.skinned_material_outputs(mat)
        np.testing.assert_array_equal(ind_b.download(np.uint32).reshape(-1, 5)[:ents], ref_ind[mi])
        got_ids = cull_b.download(np.uint32)[:ents]
        # single-instance mode: a visible entity writes its id at cullingBufferOffset + currentEntity; the oracle's
        # buffer starts as 0xFFFFFFFF, so the written positions are exactly the ones that differ from the fill
        written = ref_ids[mi][:ents] != 0xFFFFFFFF
        assert 0 < written.sum() < ents
        if len(cmds[mi]) == 1:
            np.testing.assert_array_equal(got_ids[written], ref_ids[mi][:ents][written])
        # With several commands in one skinned material the reference's slot arithmetic (indirectBufferOffset advances
        # by 1 per COMMAND, :1181, while the shader adds currentEntity) makes entity e of command k share a counter with
        # entity e-1 of command k+1: instanceCount reaches 2 and the id position depends on atomic arrival order in the
        # reference itself. Only the order-independent part (the counters, compared above) is defined there.
    w.destroy()


def test_bin_lights_matches_oracle(ctx):
    rng = np.random.default_rng(12)
    cam = scenes.default_camera()
    grid, screen = (12, 12, 24), (1920, 1080)
    z = -rng.uniform(0.2, 100.0, size=500)
    pos = np.stack([rng.uniform(-1, 1, size=500) * (-z), rng.uniform(-1, 1, size=500) * (-z) * 0.6, z], 1).astype(np.float32)
    inten = np.exp(rng.uniform(np.log(0.05), np.log(5.0), size=500)).astype(np.float32)
    points, spots = scenes.point_lights_bytes(pos[:400], inten[:400]), scenes.spot_lights_bytes(pos[400:], inten[400:])
    view = scenes.colmajor(np.eye(4))
    u = GridBuildUBO()
    u.invProj[:] = [float(x) for x in cam.inv_proj_cm]
    u.gridSize[:] = grid
    u.zNear, u.zFar = cam.z_near, cam.z_far
    u.screenDim[:] = screen
    n = grid[0] * grid[1] * grid[2]
    ref = np.zeros((n, 40), np.uint32)
    oracle.cluster_build(u, ref)
    a = GridAssignUBO()
    a.viewMat[:] = [float(x) for x in view]
    a.pointLightCount, a.spotLightCount = 400, 100
    oracle.cluster_assign(a, ref, points, spots, n, nthreads=4)
    w = World(ctx)
    out = w.bin_lights(cam.inv_proj_cm, view, grid, screen, cam.z_near, cam.z_far, ctx.buffer_from_numpy(points), 400,
                       ctx.buffer_from_numpy(spots), 100)
    ctx.synchronize()
    np.testing.assert_array_equal(out.download(np.uint32).reshape(n, 40), ref)
    # no lights: the reference skips both dispatches (:930); the buffer keeps its previous content
    w.bin_lights(cam.inv_proj_cm, view, grid, screen, cam.z_near, cam.z_far, None, 0, None, 0)
    ctx.synchronize()
    np.testing.assert_array_equal(out.download(np.uint32).reshape(n, 40), ref)
    w.destroy()


@pytest.mark.parametrize("batch_min", ["1", "100"])
def test_cull_static_views_matches_per_view_culls(ctx, batch_min, monkeypatch):
    """rvs_cull_static_views: all views of a frame (main camera + shadow passes) == the reference's one cullTheRenderData
    per view; view v occupies the v-th copy of every material's output layout. Both schedules: one view-batched launch
    (RVS_VIEW_BATCH_MIN=1) and the pass-by-pass fallback small batches take by default (forced with 100)."""
    monkeypatch.setenv("RVS_VIEW_BATCH_MIN", batch_min)
    model, layers, flags, mats = _static_scene(seed=9)
    t = host_tables.build_static_tables(mats)
    cams = [scenes.default_camera(z_far=500.0), scenes.default_camera(z_far=500.0, eye=(80, 20, 60), target=(0, 0, 0)),
            scenes.default_camera(z_far=800.0, fovy=90.0, eye=(-50, 100, 0), target=(0, 0, -10)),
            scenes.default_camera(z_far=300.0, eye=(0, 0, -100), target=(0, 0, 0))]
    shadows = [False, True, True, False]
    refs, pyrs, views = [], [], []
    for cam, sh in zip(cams, shadows):
        depth = scenes.synth_depth(512, 288, cam, n_occluders=20)
        pyr_h, dim, levels = oracle.pyramid_build(depth, 288)
        ubo = make_ubo(cam.view_proj_cm, cam.pos, t.num_cubos, shadow=sh)
        refs.append(oracle_cull(t, ubo, model, layers, flags, pyr_h, dim, levels, nthreads=4))
        p = ctx.create_pyramid(288)
        ctx.build_pyramid(p, ctx.buffer_from_numpy(depth), 512, 288)
        pyrs.append(p)
        views.append((cam.view_proj_cm, cam.pos, 0xFFFFFFFF, sh))
    w, mat_ids, keep = _world_from(ctx, model, layers, flags, mats)
    launches = ctx.launch_count
    for _ in range(2):
        w.cull_static_views(views, pyrs)
    ctx.synchronize()
    assert ctx.launch_count - launches == (2 * 2 if batch_min == "1" else 2 * 2 * len(views))  # upload + cull, per batch or per view
    ci = 0
    for mat, m in zip(mat_ids, mats):
        cull_b, ind_b, lods, ents = w.material_outputs(mat)
        cb = t.cubos[ci]
        all_ind = ind_b.download(np.uint32).reshape(-1, 5)
        all_ids = cull_b.download(np.uint32)
        rng_c = range(ci, ci + len(m.commands))
        for v, ref in enumerate(refs):
            r_ind, r_ids = ref[int(cb["indirectOutputBufferBindlessHandle"])], ref[int(cb["idOutputBufferBindlessHandle"])]
            g_ind = all_ind[v * lods:(v + 1) * lods]
            g_ids = all_ids[v * lods * ents:(v + 1) * lods * ents]
            np.testing.assert_array_equal(g_ind, r_ind)
            for x, y in zip(host_tables.sorted_slices(r_ind, r_ids, t.cubos, rng_c), host_tables.sorted_slices(g_ind, g_ids, t.cubos, rng_c)):
                np.testing.assert_array_equal(x, y)
        ci += len(m.commands)
    # back to a single-view pass on the same world
    w.cull_static(cams[0].view_proj_cm, cams[0].pos, pyrs[0])
    ctx.synchronize()
    _compare_static(ctx, w, mat_ids, mats, t, refs[0])
    w.destroy()


def test_frame_encode_equals_the_reference_pass_schedule(ctx):
    """rvs_frame_encode (one submission: pyramids, one view-batched cull over all material classes, lights binned once
    per camera) against the reference's schedule run pass by pass through the same shim: per camera one
    cullTheRenderData per material class + light binning, per shadow view one shadow cull."""
    model, layers, flags, mats = _static_scene(seed=21)
    kinds = [MAT_LIT_OPAQUE, MAT_UNLIT_OPAQUE]
    cams = [scenes.default_camera(z_far=500.0), scenes.default_camera(z_far=400.0, fovy=75.0, eye=(90, 30, 40), target=(0, 0, 0))]
    shadow_cams = [scenes.default_camera(z_far=600.0, eye=(0, 200, 1), target=(0, 0, 0)), scenes.default_camera(z_far=300.0, eye=(-120, 40, 0), target=(0, 0, 0)),
                   scenes.default_camera(z_far=300.0, eye=(0, 10, 150), target=(0, 0, 0))]
    rng = np.random.default_rng(5)
    z = -rng.uniform(0.2, 100.0, size=300)
    pos = np.stack([rng.uniform(-1, 1, size=300) * (-z), rng.uniform(-1, 1, size=300) * (-z) * 0.6, z], 1).astype(np.float32)
    inten = np.exp(rng.uniform(np.log(0.05), np.log(5.0), size=300)).astype(np.float32)
    points, spots = ctx.buffer_from_numpy(scenes.point_lights_bytes(pos[:200], inten[:200])), ctx.buffer_from_numpy(scenes.spot_lights_bytes(pos[200:], inten[200:]))
    grid, screen = (16, 9, 24), (1280, 720)
    depths = [ctx.buffer_from_numpy(scenes.synth_depth(640, 360, c, n_occluders=15, seed=3 + i)) for i, c in enumerate(cams)]
    empty = ctx.create_pyramid(64)  # shadow passes: zeroed pyramid = nothing occluded

    # ---- the reference's schedule, pass by pass ----
    w1, mat1, keep1 = _world_from(ctx, model, layers, flags, mats, kinds=kinds)
    want = {}  # (view, material) -> (indirect, ids)
    want_clusters = []
    views_all = [(c, False) for c in cams] + [(c, True) for c in shadow_cams]
    for v, (cam, shadow) in enumerate(views_all):
        pyr = empty
        if not shadow:
            pyr = ctx.create_pyramid(360)
            ctx.build_pyramid(pyr, depths[v], 640, 360)
        for mat, kind in zip(mat1, kinds):
            w1.cull_static(cam.view_proj_cm, cam.pos, pyr, shadow=shadow, mask=kind)
            ctx.synchronize()
            cull_b, ind_b, lods, ents = w1.material_outputs(mat)
            want[(v, mat)] = (ind_b.download(np.uint32).reshape(-1, 5)[:lods].copy(), cull_b.download(np.uint32)[:lods * ents].copy())
        if not shadow:
            out = w1.bin_lights(cam.inv_proj_cm, scenes.colmajor(np.eye(4)), grid, screen, cam.z_near, cam.z_far, points, 200, spots, 100)
            ctx.synchronize()
            want_clusters.append(out.download(np.uint32).copy())

    # ---- one submission ----
    w2, mat2, keep2 = _world_from(ctx, model, layers, flags, mats, kinds=kinds)
    pyrs = [ctx.create_pyramid(360) for _ in cams]
    cam_descs = [dict(view_proj=c.view_proj_cm, view=scenes.colmajor(np.eye(4)), inv_proj=c.inv_proj_cm, cam_pos=c.pos, z_near=c.z_near, z_far=c.z_far,
                      pyramid=p, depth=d, depth_wh=(640, 360), screen=screen) for c, p, d in zip(cams, pyrs, depths)]
    launches0 = ctx.launch_count
    w2.frame_encode(cam_descs, [(c.view_proj_cm, c.pos, 0xFFFFFFFF, True) for c in shadow_cams], [empty] * len(shadow_cams), points, 200, spots, 100, grid)
    ctx.synchronize()
    frame_launches = ctx.launch_count - launches0
    t = host_tables.build_static_tables(mats)
    ci = 0
    for mat, m in zip(mat2, mats):
        cull_b, ind_b, lods, ents = w2.material_outputs(mat)
        all_ind, all_ids = ind_b.download(np.uint32).reshape(-1, 5), cull_b.download(np.uint32)
        rng_c = range(ci, ci + len(m.commands))
        for v in range(len(views_all)):
            r_ind, r_ids = want[(v, mat)]
            g_ind, g_ids = all_ind[v * lods:(v + 1) * lods], all_ids[v * lods * ents:(v + 1) * lods * ents]
            np.testing.assert_array_equal(g_ind, r_ind)
            # slices of one material's buffers: its cubos carry offsets relative to that material, as in the tables
            for x, y in zip(host_tables.sorted_slices(r_ind, r_ids, t.cubos, rng_c), host_tables.sorted_slices(g_ind, g_ids, t.cubos, rng_c)):
                np.testing.assert_array_equal(x, y)
        ci += len(m.commands)
    for c in range(len(cams)):
        # index entries past a cluster's counts are never written (stale in a reused buffer, as in the reference)
        got, ref = w2.frame_cluster_buffer(c).download(np.uint32).reshape(-1, 40), want_clusters[c].reshape(-1, 40)
        np.testing.assert_array_equal(got[:, :8], ref[:, :8])
        assert ref[:, 6].max() > 0 and ref[:, 7].max() > 0
        live = np.concatenate([np.arange(16)[None, :] < ref[:, 6:7], np.arange(16)[None, :] < ref[:, 7:8]], axis=1)
        np.testing.assert_array_equal(got[:, 8:][live], ref[:, 8:][live])
    # 2 pyramid builds x 2 launches + the culls + 2 x (cluster build + assign prepare/assign); the culls are one (table upload,
    # cull) pair per view below 8 views (pass-by-pass fallback), one view-batched launch + upload from 8 views on
    nv = len(views_all)
    assert frame_launches <= 4 + (2 * nv if nv < 8 else 4) + 2 * 4, frame_launches
    w1.destroy()
    w2.destroy()


@pytest.mark.parametrize("n,size", [(6000, (512, 288)), (300_000, (3840, 2160)), (6000, (100, 37)), (6000, (101, 50)), (6000, (30, 20))])
def test_fused_pyramid_and_cull_encode_matches_oracle(ctx, n, size):
    """rvs_cull_static_with_pyramid (tables -> pyramid kernels -> cull overlapping the pyramid's tail, its programmatic
    dependency wait moved to the first Hi-Z batch) gives the oracle's pyramid and the oracle's cull, frame after frame
    with a depth buffer that changes between frames."""
    model, layers, flags, mats = _static_scene(seed=11, n=n)
    cam = scenes.default_camera(z_far=500.0)
    wd, hd = size
    w, mat_ids, keep = _world_from(ctx, model, layers, flags, mats)
    pyr = ctx.create_pyramid(min(size))
    t = host_tables.build_static_tables(mats)
    ubo = make_ubo(cam.view_proj_cm, cam.pos, t.num_cubos)
    for frame in range(3):
        depth = scenes.synth_depth(wd, hd, cam, n_occluders=30, seed=frame + 1)
        pyr_h, dim, levels = oracle.pyramid_build(depth, min(size))
        ref = oracle_cull(t, ubo, model, layers, flags, pyr_h, dim, levels, nthreads=4)
        d = ctx.buffer_from_numpy(depth)
        w.cull_static_with_pyramid(cam.view_proj_cm, cam.pos, pyr, d, wd, hd)
        ctx.synchronize()
        np.testing.assert_array_equal(pyr_h, pyr.texels.download(np.float32))
        fm_ref = footmin_chain(pyr_h, dim, levels)  # the upper levels are finished inside the cull kernel (deferred build)
        np.testing.assert_array_equal(fm_ref, ctx.pyramid_footmin(pyr).download(np.float32)[: fm_ref.size])
        _compare_static(ctx, w, mat_ids, mats, t, ref)
        d.destroy()
    w.destroy()


def test_captured_frame_replays_with_new_per_entity_data(ctx):
    """rve_ctx_capture_begin / _end: a frame (tables -> pyramid -> cull, then an 8-view batch and the light binning of
    rvs_frame_encode) recorded once and replayed with ONE launch per frame gives the oracle's result for the data that is
    in the buffers AT REPLAY TIME: transforms and the depth image change between replays."""
    model, layers, flags, mats = _static_scene(seed=21)
    cam = scenes.default_camera(z_far=500.0)
    t = host_tables.build_static_tables(mats)
    ubo = make_ubo(cam.view_proj_cm, cam.pos, t.num_cubos)
    w, mat_ids, keep = _world_from(ctx, model, layers, flags, mats)
    pyr = ctx.create_pyramid(288)
    depth = scenes.synth_depth(512, 288, cam, n_occluders=30, seed=1)
    d_depth = ctx.buffer_from_numpy(depth)
    w.cull_static_with_pyramid(cam.view_proj_cm, cam.pos, pyr, d_depth, 512, 288)  # warm-up: buffers, scratch, bindless table
    ctx.synchronize()
    launches = ctx.launch_count
    graph = ctx.capture(lambda: w.cull_static_with_pyramid(cam.view_proj_cm, cam.pos, pyr, d_depth, 512, 288))
    recorded = ctx.launch_count - launches
    assert recorded == 2  # pyramid tiles (carrying the table upload), cull (which finishes the upper levels) - recorded, not executed
    rng = np.random.default_rng(3)
    for frame in range(3):
        model2 = scenes.random_transforms(rng, len(model), 120.0, 60.0)
        depth2 = scenes.synth_depth(512, 288, cam, n_occluders=30, seed=10 + frame)
        keep[0].upload(model2)
        d_depth.upload(depth2)
        graph.launch()
        ctx.synchronize()
        pyr_h, dim, levels = oracle.pyramid_build(depth2, 288)
        np.testing.assert_array_equal(pyr_h, pyr.texels.download(np.float32))
        ref = oracle_cull(t, ubo, model2, layers, flags, pyr_h, dim, levels, nthreads=4)
        _compare_static(ctx, w, mat_ids, mats, t, ref)
    graph.destroy()
    # a view batch (host view table, strides: staged into pinned blocks owned by the graph)
    cams = [scenes.default_camera(z_far=500.0, eye=(10.0 * i, 5.0, 40.0 - 9.0 * i), target=(0.0, 0.0, -20.0)) for i in range(9)]
    views = [(c.view_proj_cm, c.pos, 0xFFFFFFFF, i % 3 == 1) for i, c in enumerate(cams)]
    w.cull_static_views(views, [pyr] * 9)
    ctx.synchronize()
    graph = ctx.capture(lambda: w.cull_static_views(views, [pyr] * 9))
    model3 = scenes.random_transforms(rng, len(model), 120.0, 60.0)
    keep[0].upload(model3)
    graph.launch()
    ctx.synchronize()
    pyr_h = pyr.texels.download(np.float32)
    dim, levels = oracle.pyramid_dims(288)
    ci = 0
    for mat, m in zip(mat_ids, mats):
        cull_b, ind_b, lods, ents = w.material_outputs(mat)
        all_ind, all_ids = ind_b.download(np.uint32).reshape(-1, 5), cull_b.download(np.uint32)
        cb = t.cubos[ci]
        rng_c = range(ci, ci + len(m.commands))
        for v, (c, (_, _, _, sh)) in enumerate(zip(cams, views)):
            ref = oracle_cull(t, make_ubo(c.view_proj_cm, c.pos, t.num_cubos, shadow=sh), model3, layers, flags, pyr_h, dim, levels, nthreads=4)
            r_ind, r_ids = ref[int(cb["indirectOutputBufferBindlessHandle"])], ref[int(cb["idOutputBufferBindlessHandle"])]
            g_ind, g_ids = all_ind[v * lods:(v + 1) * lods], all_ids[v * lods * ents:(v + 1) * lods * ents]
            np.testing.assert_array_equal(g_ind, r_ind)
            for x, y in zip(host_tables.sorted_slices(r_ind, r_ids, t.cubos, rng_c), host_tables.sorted_slices(g_ind, g_ids, t.cubos, rng_c)):
                np.testing.assert_array_equal(x, y)
        ci += len(m.commands)
    graph.destroy()
    w.destroy()
