"""GPU parity for skinning, skinned draw-call prep, cluster build/assign and the scene generator."""
import ctypes as C

import numpy as np
import pytest

import oracle
from ravengine_b200 import (GridAssignUBO, GridBuildUBO, SceneGenParams, SkinningPrepareUBO, SkinningUBO, scenes)

pytestmark = pytest.mark.gpu


def _skin_case(ctx, n_verts, n_obj, n_bones, seed, read_off=0, write_off=0, bone_off=0, identity=False, coherent=False):
    rng = np.random.default_rng(seed)
    d = scenes.skinning_inputs(rng, n_verts, n_obj, n_bones, coherent_joints=coherent)
    if identity:
        d["pose"][:] = np.eye(4, dtype=np.float32).reshape(16)
    pad = lambda a, k: np.concatenate([rng.random((read_off, k), dtype=np.float32), a])
    ins = [pad(d["pos"], 3), pad(d["nrm"], 3), pad(d["tan"], 3), pad(d["bit"], 3), pad(d["uv"], 2)]
    pose = np.concatenate([rng.random((bone_off, 16), dtype=np.float32), d["pose"]])
    total = write_off + n_obj * n_verts
    outs_ref = [np.full((total, k), -7.0, np.float32) for k in (3, 3, 3, 3, 2)]
    ubo = SkinningUBO(n_obj, n_verts, n_bones, bone_off, write_off, read_off)
    oracle.skin(ubo, outs_ref, ins, pose, d["weights"], nthreads=4)
    d_outs = [ctx.buffer_from_numpy(np.full((total, k), -7.0, np.float32)) for k in (3, 3, 3, 3, 2)]
    d_ins = [ctx.buffer_from_numpy(a) for a in ins]
    d_pose, d_w = ctx.buffer_from_numpy(pose), ctx.buffer_from_numpy(d["weights"])
    ctx.skin(ubo, d_outs, d_ins, d_pose, d_w)
    ctx.synchronize()
    outs = [b.download(np.float32)[: r.size].reshape(r.shape) for b, r in zip(d_outs, outs_ref)]
    for b in d_outs + d_ins + [d_pose, d_w]:
        b.destroy()
    return ubo, ins, pose, d["weights"], outs_ref, outs


@pytest.mark.parametrize("shape", [(1, 1, 1), (33, 3, 5), (1000, 17, 64), (20000, 24, 64), (257, 9, 300), (64, 2, 5000)])
def test_skinning_matches_oracle(ctx, shape):
    nv, no, nb = shape
    ubo, ins, pose, w, ref, got = _skin_case(ctx, nv, no, nb, seed=nv + no, read_off=7, write_off=13, bone_off=3)
    for r, g in zip(ref, got):
        np.testing.assert_array_equal(r, g)  # same op order, no FMA: bit-exact (stronger than the 1e-5 bar)
    # and within 1e-5 relative of a float64 evaluation (north_star tolerance)
    total = ref[0].shape[0]
    f64 = [np.zeros((total, 3), np.float64) for _ in range(4)]
    oracle.skin_f64(ubo, f64, ins, pose, w)
    sl = slice(13, None)
    for r64, g in zip(f64, got[:4]):
        scale = np.maximum(np.abs(r64[sl]).max(), 1e-30)
        assert np.abs(g[sl] - r64[sl]).max() / scale <= 1e-5


def test_skinning_more_object_groups_than_one_grid_holds(ctx):
    """600 000 objects x 1 bone = 75 000 groups of 8: past grid.y's 65 535, so rve_skin takes a second launch."""
    launches = ctx.launch_count
    ubo, ins, pose, w, ref, got = _skin_case(ctx, 3, 600_000, 1, seed=2)
    assert ctx.launch_count - launches == 2
    for r, g in zip(ref, got):
        np.testing.assert_array_equal(r, g)


def test_skinning_identity_palette_is_exact_copy(ctx):
    ubo, ins, pose, w, ref, got = _skin_case(ctx, 500, 4, 16, seed=5, identity=True)
    for o in range(4):
        np.testing.assert_array_equal(got[4][o * 500:(o + 1) * 500], ins[4])
        # weights sum to 1 only up to rounding, so compare against the oracle rather than the input
    for r, g in zip(ref, got):
        np.testing.assert_array_equal(r, g)


def test_skinned_drawcall_prepare(ctx):
    ubo = SkinningPrepareUBO(1000, 77, 36, 24, 333, 5, 11)
    ref = np.full((400, 5), 0xABCD, np.uint32)
    oracle.skinned_drawcall_prepare(ubo, ref)
    d = ctx.buffer_from_numpy(np.full((400, 5), 0xABCD, np.uint32))
    ctx.skinned_drawcall_prepare(ubo, d)
    ctx.synchronize()
    np.testing.assert_array_equal(ref, d.download(np.uint32).reshape(400, 5))


def _grid_ubo(cam, grid, screen=(3840, 2160)):
    u = GridBuildUBO()
    u.invProj[:] = [float(x) for x in cam.inv_proj_cm]
    u.gridSize[:] = grid
    u.zNear, u.zFar = cam.z_near, cam.z_far
    u.screenDim[:] = screen
    return u


@pytest.mark.parametrize("grid", [(12, 12, 24), (16, 9, 24), (7, 5, 3), (1, 1, 1)])
def test_cluster_build(ctx, grid):
    cam = scenes.default_camera()
    u = _grid_ubo(cam, grid)
    n = grid[0] * grid[1] * grid[2]
    rng = np.random.default_rng(1)
    init = rng.integers(0, 255, size=(n, 160), dtype=np.uint8)
    ref = init.copy()
    oracle.cluster_build(u, ref)
    d = ctx.buffer_from_numpy(init)
    ctx.cluster_build(u, d)
    ctx.synchronize()
    np.testing.assert_array_equal(ref, d.download(np.uint8).reshape(n, 160))


@pytest.mark.parametrize("case", [(16, 9, 24, 4096, 0), (12, 12, 24, 3000, 200), (16, 9, 24, 65536, 64), (4, 4, 4, 0, 0),
                                  (5, 3, 2, 17, 40)])
def test_cluster_assign(ctx, case):
    gx, gy, gz, npnt, nspot = case
    cam = scenes.default_camera()
    u = _grid_ubo(cam, (gx, gy, gz))
    n = gx * gy * gz
    rng = np.random.default_rng(npnt + nspot)
    init = rng.integers(0, 255, size=(n, 160), dtype=np.uint8)
    ref = init.copy()
    oracle.cluster_build(u, ref)

    def lights(k):
        z = -rng.uniform(0.2, 100.0, size=k)
        x = rng.uniform(-1, 1, size=k) * (-z) * 1.03
        y = rng.uniform(-1, 1, size=k) * (-z) * 0.6
        inten = np.exp(rng.uniform(np.log(0.05), np.log(5.0), size=k))
        return np.stack([x, y, z], 1).astype(np.float32), inten.astype(np.float32)

    view = scenes.look_at((1.0, 2.0, 3.0), (0.5, 1.5, -4.0))
    inv_view = np.linalg.inv(view)
    pv, pi = lights(npnt)
    sv, si = lights(nspot)
    to_world = lambda v: (np.c_[v, np.ones(len(v))] @ inv_view.T)[:, :3].astype(np.float32)
    points = scenes.point_lights_bytes(to_world(pv), pi) if npnt else None
    spots = scenes.spot_lights_bytes(to_world(sv), si) if nspot else None
    a = GridAssignUBO()
    a.viewMat[:] = [float(x) for x in scenes.colmajor(view)]
    a.pointLightCount, a.spotLightCount = npnt, nspot
    oracle.cluster_assign(a, ref, points, spots, n, nthreads=4)
    d = ctx.buffer_from_numpy(init)
    ctx.cluster_build(u, d)
    dp = ctx.buffer_from_numpy(points) if npnt else None
    ds = ctx.buffer_from_numpy(spots) if nspot else None
    ctx.cluster_assign(a, d, dp, ds, n)
    ctx.synchronize()
    got = d.download(np.uint8).reshape(n, 160)
    np.testing.assert_array_equal(ref, got)  # whole struct: AABB, counts, index lists, stale tail entries


def test_gen_instances_matches_oracle(ctx):
    p = SceneGenParams(1234, 500.0, 50.0, 0.5, 2.0, 0.05, 0.05, 100)
    n, first = 50_000, 123_456_789_012
    total = 100 + n
    model, layers = np.zeros((total, 16), np.float32), np.zeros(total, np.uint32)
    flags, ids = np.zeros(total, np.uint16), np.zeros(n, np.uint32)
    oracle.gen_instances(p, first, n, model, layers, flags, ids)
    bm, bl = ctx.create_buffer(total * 64), ctx.create_buffer(total * 4)
    bf, bi = ctx.create_buffer(total * 2), ctx.create_buffer(n * 4)
    ctx.gen_instances(p, first, n, bm, bl, bf, bi)
    ctx.synchronize()
    np.testing.assert_array_equal(model, bm.download(np.float32).reshape(total, 16))
    np.testing.assert_array_equal(layers, bl.download(np.uint32))
    np.testing.assert_array_equal(flags, bf.download(np.uint16))
    np.testing.assert_array_equal(ids, bi.download(np.uint32))


def test_shard_offsets_and_count_gather(ctx):
    """8(e) helpers: instanceCount gather from the 20-B-stride commands and the exclusive scan over ranks."""
    rng = np.random.default_rng(5)
    world, slots = 8, 12
    counts = rng.integers(0, 1_000_000, size=(world, slots)).astype(np.uint32)
    b_counts, b_off, b_tot = ctx.buffer_from_numpy(counts), ctx.create_buffer(world * slots * 4), ctx.create_buffer(slots * 4)
    ctx.shard_offsets(b_counts, world, slots, b_off, b_tot)
    cmds = rng.integers(0, 2**31, size=(slots, 5)).astype(np.uint32)
    b_cmds, b_out = ctx.buffer_from_numpy(cmds), ctx.create_buffer(64 + slots * 4)
    ctx.gather_instance_counts(b_cmds, 3, slots - 3, b_out, 16)
    ctx.synchronize()
    c64 = counts.astype(np.uint64)
    np.testing.assert_array_equal(b_off.download(np.uint32).reshape(world, slots), (np.cumsum(c64, axis=0) - c64).astype(np.uint32))
    np.testing.assert_array_equal(b_tot.download(np.uint32), c64.sum(axis=0).astype(np.uint32))
    np.testing.assert_array_equal(b_out.download(np.uint32)[4:4 + slots - 3], cmds[3:, 1])


def test_sharded_cull_single_rank_exchange(ctx):
    """ShardedCull with world_size 1: the count gather + scan path runs on the context's stream and reports the
    indirect buffers' instance counts as totals with zero offsets."""
    import torch

    from ravengine_b200.sharding import ShardedCull
    from ravengine_b200.world import World

    rng = np.random.default_rng(21)
    n = 20_000
    model = scenes.random_transforms(rng, n, 200.0, 50.0)
    w = World(ctx)
    keep = [ctx.buffer_from_numpy(model), ctx.buffer_from_numpy(np.full(n, 0xFFFFFFFF, np.uint32)), ctx.buffer_from_numpy(np.full(n, 0xFFFF, np.uint16))]
    w.set_entity_buffers(*keep)
    mats = []
    for i in range(2):
        ids = ctx.buffer_from_numpy(np.arange(i * n // 2, (i + 1) * n // 2, dtype=np.uint32))
        keep.append(ids)
        mesh = w.mesh_collection(1.5, [0.0, 40.0, 120.0], [36, 24, 12], [0, 36, 60], [0, 24, 40])
        m = w.material()
        w.add_command(m, mesh, ids, n // 2)
        mats.append(m)
    cam = scenes.default_camera(z_far=500.0)
    w.cull_static(cam.view_proj_cm, cam.pos, ctx.create_pyramid(64))
    stream = torch.cuda.ExternalStream(ctx.stream) if ctx.stream else torch.cuda.current_stream()
    sh = ShardedCull(ctx, w, mats, [3, 3], 1, stream)
    want = None
    for frame in range(3):  # double-buffered: frames alternate buffer sets
        if frame:
            w.cull_static(cam.view_proj_cm, cam.pos, ctx.create_pyramid(64))
        b = sh.exchange()
        assert b == frame % 2
        sh.finish()
        ctx.synchronize()
        torch.cuda.synchronize()
        want = np.concatenate([w.material_outputs(m)[1].download(np.uint32).reshape(-1, 5)[:3, 1] for m in mats])
        np.testing.assert_array_equal(sh.t_tot[b].cpu().numpy().view(np.uint32), want)
        assert int(sh.t_off[b].abs().sum()) == 0 and want.sum() > 0
    # the multi-buffer gather equals the per-buffer one
    out1, out2 = ctx.create_buffer(6 * 4), ctx.create_buffer(6 * 4)
    inds = [w.material_outputs(m)[1] for m in mats]
    ctx.gather_instance_counts_multi(inds, [0, 0], [3, 3], out1)
    for i, ind in enumerate(inds):
        ctx.gather_instance_counts(ind, 0, 3, out2, i * 12)
    ctx.synchronize()
    np.testing.assert_array_equal(out1.download(np.uint32), out2.download(np.uint32))
    np.testing.assert_array_equal(out1.download(np.uint32), want)
    w.destroy()


@pytest.mark.gpu
def test_p2p_count_exchange_two_ranks_on_one_device(ctx):
    """csrc/p2p.cu: two 'ranks' (two streams of one device, mailboxes in the same memory) exchange per-slot counts in
    one kernel each - peer stores, flags, exclusive scan - over several epochs; against numpy."""
    import torch

    from ravengine_b200 import Context

    world, slots = 2, 7
    s2 = torch.cuda.Stream()
    ctxs = [ctx, Context(ctx.device, s2.cuda_stream)]
    nbytes = ctx.mailbox_bytes(world, slots)
    boxes = [ctx.create_buffer(nbytes) for _ in range(world)]
    for b in boxes:
        b.zero()
    ctx.synchronize()
    rng = np.random.default_rng(5)
    # the other context sees the same device memory through wrapped pointers
    def view(c, b):
        import ctypes as C
        out = C.c_uint32()
        assert c.lib.rve_buffer_wrap(c.h, C.c_void_p(b.ptr), b.nbytes, C.byref(out)) == 0
        from ravengine_b200.api import Buffer
        return Buffer(c, out.value, b.nbytes)

    per_rank = []
    for r, c in enumerate(ctxs):
        mb = boxes if r == 0 else [view(c, b) for b in boxes]
        per_rank.append(dict(c=c, mb=mb, dense=c.create_buffer(slots * 4), off=c.create_buffer(world * slots * 4),
                             tot=c.create_buffer(slots * 4), status=c.create_buffer(16)))
        per_rank[-1]["status"].zero()
    for epoch in range(1, 6):
        counts = rng.integers(0, 1000, size=(world, slots)).astype(np.uint32)
        for r, pr in enumerate(per_rank):
            pr["dense"].upload(counts[r])
        for r, pr in enumerate(per_rank):
            pr["c"].exchange_counts_p2p(pr["dense"], slots, pr["mb"], r, world, epoch, pr["off"], pr["tot"], pr["status"])
        expect_off = np.cumsum(counts, axis=0, dtype=np.uint32) - counts
        for pr in per_rank:
            pr["c"].synchronize()
            assert pr["status"].download(np.uint32)[0] == 0
            assert np.array_equal(pr["off"].download(np.uint32).reshape(world, slots), expect_off)
            assert np.array_equal(pr["tot"].download(np.uint32), counts.sum(axis=0, dtype=np.uint32))
    # the same exchange with the kernel gathering the counts out of indirect command buffers (two "materials")
    split = [3, 4]
    ind = [[pr["c"].create_buffer((first + n) * 20) for first, n in zip((2, 0), split)] for pr in per_rank]
    for epoch in range(6, 9):
        counts = rng.integers(0, 1000, size=(world, slots)).astype(np.uint32)
        for r, pr in enumerate(per_rank):
            k = 0
            for j, (first, n) in enumerate(zip((2, 0), split)):
                cmds = rng.integers(0, 2 ** 31, size=(first + n, 5)).astype(np.uint32)
                cmds[first:, 1] = counts[r, k:k + n]
                ind[r][j].upload(cmds)
                k += n
        for r, pr in enumerate(per_rank):
            pr["c"].exchange_counts_p2p_indirect(ind[r], [2, 0], split, pr["mb"], r, world, epoch, pr["off"], pr["tot"], pr["status"])
        expect_off = np.cumsum(counts, axis=0, dtype=np.uint32) - counts
        for pr in per_rank:
            pr["c"].synchronize()
            assert pr["status"].download(np.uint32)[0] == 0
            assert np.array_equal(pr["off"].download(np.uint32).reshape(world, slots), expect_off)
            assert np.array_equal(pr["tot"].download(np.uint32), counts.sum(axis=0, dtype=np.uint32))
    ctxs[1].close()
