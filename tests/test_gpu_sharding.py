"""GPU tests of the multi-GPU paths that can run on one device: the device-side id gather into a consumer's
global culling buffers (csrc/p2p.cu, two 'ranks' = two streams of one device) against the single-GPU cull of the whole
scene, ShardedCull's gather at world size 1, and partition B (view ranges per rank) against the one-GPU view batch."""
import math

import numpy as np
import pytest

from ravengine_b200 import scenes, sharding
from ravengine_b200.sharding import SlotPlan
from ravengine_b200.world import World

pytestmark = pytest.mark.gpu

E = 30_000
LODS = [(0.0, 40.0, 120.0), (math.inf,), (0.0, 60.0)]
RADII = [1.7, 2.2, 0.9]
MAT_OF_CMD = [0, 0, 1]
FRACTIONS = [0.55, 0.12, 0.33]


def _scene():
    rng = np.random.default_rng(91)
    model = scenes.random_transforms(rng, E, 300.0, 40.0)
    layers = np.full(E, 0xFFFFFFFF, np.uint32)
    layers[rng.random(E) < 0.05] = 0
    flags = np.full(E, 0xFFFF, np.uint16)
    flags[rng.random(E) < 0.1] &= 0xFFFC
    cam = scenes.default_camera(z_far=1000.0)
    depth = scenes.synth_depth(320, 180, cam, n_occluders=20)
    return rng, model, layers, flags, cam, depth


def _lists(rng, lo, hi):
    ids = (lo + rng.permutation(hi - lo)).astype(np.uint32)
    n = hi - lo
    cuts = [0, int(n * FRACTIONS[0]), int(n * (FRACTIONS[0] + FRACTIONS[1])), n]
    return [ids[cuts[i]:cuts[i + 1]] for i in range(3)]


def _plans(lists):
    return [SlotPlan((len(lists[0]), len(lists[1])), (3, 1)), SlotPlan((len(lists[2]),), (2,))]


def _world(c, model, layers, flags, lists, keep):
    w = World(c)
    bufs = [c.buffer_from_numpy(model), c.buffer_from_numpy(layers), c.buffer_from_numpy(flags)]
    keep += bufs
    w.set_entity_buffers(*bufs)
    mats = [w.material(), w.material()]
    for i in range(3):
        k = len(LODS[i])
        mesh = w.mesh_collection(RADII[i], LODS[i], [36, 24, 12][:k], [0, 36, 60][:k], [0, 24, 40][:k])
        ids = c.buffer_from_numpy(lists[i])
        keep.append(ids)
        w.add_command(mats[MAT_OF_CMD[i]], mesh, ids, len(lists[i]))
    return w, mats


def _cull(c, w, cam, depth, keep):
    pyr = c.create_pyramid(180)
    d = c.buffer_from_numpy(depth)
    keep.append(d)
    c.build_pyramid(pyr, d, depth.shape[1], depth.shape[0])
    w.cull_static(cam.view_proj_cm, cam.pos, pyr)


def _view(c, b):
    return c.wrap_pointer(b.ptr, b.nbytes, keepalive=b)


@pytest.mark.parametrize("consumers", [0b01, 0b11])
def test_device_gather_two_ranks_on_one_device_equals_single_gpu_cull(ctx, consumers):
    import torch

    from ravengine_b200 import Context

    rng, model, layers, flags, cam, depth = _scene()
    world = 2
    ranges = sharding.split_range(E, world)
    lists = [_lists(rng, lo, hi) for lo, hi in ranges]
    global_lists = [np.concatenate([lists[r][i] for r in range(world)]) for i in range(3)]
    keep = []
    # reference: the whole scene on one GPU
    wref, mref = _world(ctx, model, layers, flags, global_lists, keep)
    _cull(ctx, wref, cam, depth, keep)
    ctx.synchronize()
    ref = [(wref.material_outputs(m)[1].download(np.uint32).reshape(-1, 5), wref.material_outputs(m)[0].download(np.uint32)) for m in mref]

    s2 = torch.cuda.Stream()
    ctxs = [ctx, Context(ctx.device, s2.cuda_stream)]
    gplans, ranks = [], []
    slots = 6
    boxes = [ctx.create_buffer(ctx.mailbox_bytes(world, slots)) for _ in range(world)]
    for b in boxes:
        b.zero()
    gp0 = sharding.gather_plan(_plans([l - ranges[0][0] for l in lists[0]]), _plans(global_lists))
    dst_own = [[ctxs[q].create_buffer(max(4 * n, 16)) for n in gp0.global_slots] if (consumers >> q) & 1 else None for q in range(world)]
    for q in range(world):
        for b in dst_own[q] or []:
            b.upload(np.full(b.nbytes // 4, 0xFFFFFFFF, np.uint32))
    for r, c in enumerate(ctxs):
        lo, hi = ranges[r]
        local_lists = [l - lo for l in lists[r]]
        w, mats = _world(c, model[lo:hi], layers[lo:hi], flags[lo:hi], local_lists, keep)
        _cull(c, w, cam, depth, keep)
        gplans.append(sharding.gather_plan(_plans(local_lists), _plans(global_lists)))
        mb = boxes if r == 0 else [_view(c, b) for b in boxes]
        dst = [[(x if q == r else _view(c, x)) for x in dst_own[q]] if dst_own[q] else None for q in range(world)]
        ranks.append(dict(c=c, w=w, mats=mats, mb=mb, dst=dst, off=c.create_buffer(world * slots * 4), tot=c.create_buffer(slots * 4),
                          status=c.create_buffer(16)))
        ranks[-1]["status"].zero()
    for c in ctxs:
        c.synchronize()
    for epoch in (1, 2):  # two frames: flags and the ticket re-arm
        for r, pr in enumerate(ranks):
            ind = [pr["w"].material_outputs(m)[1] for m in pr["mats"]]
            pr["c"].exchange_counts_p2p_indirect(ind, [0, 0], [4, 2], pr["mb"], r, world, epoch, pr["off"], pr["tot"], pr["status"])
            src = [pr["w"].material_outputs(m)[0] for m in pr["mats"]]
            pr["c"].gather_ids_p2p(src, gplans[r], pr["dst"], consumers, ranges[r][0], pr["off"], pr["tot"], pr["mb"], r, world, epoch, pr["status"])
        for pr in ranks:
            pr["c"].synchronize()
            assert pr["status"].download(np.uint32)[0] == 0
        totals = ranks[0]["tot"].download(np.uint32)
        assert np.array_equal(totals, np.concatenate([ind[:, 1] for ind, _ in ref]))
        for q in range(world):
            if not (consumers >> q) & 1:
                continue
            s = 0
            for m, plan in enumerate(_plans(global_lists)):
                got = dst_own[q][m].download(np.uint32)[: plan.culling_slots]
                for start in plan.slot_starts():
                    n = int(totals[s])
                    assert np.array_equal(np.sort(got[start:start + n]), np.sort(ref[m][1][start:start + n])), (q, m, s)
                    s += 1
                # nothing outside the visible slices was written
                assert int((got != 0xFFFFFFFF).sum()) == int(sum(totals[s - len(plan.slot_starts()):s]))
    ctxs[1].close()
    wref.destroy()


def test_sharded_cull_gather_world_size_1(ctx):
    """ShardedCull.setup_gather / gather with one rank: the 'global' buffers receive exactly the local visible slices."""
    import torch

    from ravengine_b200.sharding import ShardedCull

    rng, model, layers, flags, cam, depth = _scene()
    lists = _lists(rng, 0, E)
    keep = []
    w, mats = _world(ctx, model, layers, flags, lists, keep)
    _cull(ctx, w, cam, depth, keep)
    stream = torch.cuda.ExternalStream(ctx.stream) if ctx.stream else torch.cuda.current_stream()
    sh = ShardedCull(ctx, w, mats, [4, 2], 1, stream)
    sh.setup_gather(_plans(lists), _plans(lists), [0], 0)
    for g in sh.global_ids:
        g.upload(np.full(g.nbytes // 4, 0xFFFFFFFF, np.uint32))
    b = sh.exchange()
    sh.gather(b)
    sh.finish()
    ctx.synchronize()
    torch.cuda.synchronize()
    assert not sh.timed_out() and sh.status.download(np.uint32)[0] == 0
    totals = sh.t_tot[b].cpu().numpy().view(np.uint32)
    s = 0
    for m, plan in enumerate(_plans(lists)):
        local = w.material_outputs(mats[m])[0].download(np.uint32)
        got = sh.global_ids[m].download(np.uint32)
        for start in plan.slot_starts():
            n = int(totals[s])
            assert np.array_equal(got[start:start + n], local[start:start + n])
            s += 1
    assert totals.sum() > 0
    w.destroy()


def test_two_frame_slots_keep_their_own_counts(ctx):
    """ShardedCull.add_frame_slot (two frames in flight: a second world = a second set of indirect / culling buffers): frames
    alternate between the slots with different cameras and no wait on the other slot's exchange; every exchange reports the
    counts of the slot it was given, and the slot's buffers are not touched by the other slot's frame."""
    import torch

    from ravengine_b200.sharding import ShardedCull

    rng, model, layers, flags, cam, depth = _scene()
    lists = _lists(rng, 0, E)
    keep = []
    worlds = [_world(ctx, model, layers, flags, lists, keep) for _ in range(2)]
    cams = [cam, scenes.default_camera(z_far=1000.0, eye=(120.0, 10.0, 60.0), target=(0.0, 0.0, -80.0))]
    pyr = ctx.create_pyramid(180)
    d = ctx.buffer_from_numpy(depth)
    ctx.build_pyramid(pyr, d, depth.shape[1], depth.shape[0])
    stream = torch.cuda.ExternalStream(ctx.stream) if ctx.stream else torch.cuda.current_stream()
    sh = ShardedCull(ctx, worlds[0][0], worlds[0][1], [4, 2], 1, stream)
    assert sh.add_frame_slot(worlds[1][0], worlds[1][1]) == 1
    sh._solo_mailbox()
    sh.mode = "p2p"  # one rank: the peer-memory kernel with itself as the only peer (the fused, wait-free exchange path)
    expected = []
    for slot in range(2):  # what each camera sees, from a plain cull
        w, mats = worlds[slot]
        w.cull_static(cams[slot].view_proj_cm, cams[slot].pos, pyr)
        ctx.synchronize()
        expected.append(np.concatenate([w.material_outputs(m)[1].download(np.uint32).reshape(-1, 5)[:, 1] for m in mats]))
    assert not np.array_equal(expected[0], expected[1])
    for frame in range(6):
        slot = frame & 1
        w, mats = worlds[slot]
        sh.begin_cull(slot)
        w.cull_static(cams[slot].view_proj_cm, cams[slot].pos, pyr)
        b = sh.exchange(slot=slot)
        sh.finish()
        ctx.synchronize()
        torch.cuda.synchronize()
        assert not sh.timed_out()
        totals = sh.t_tot[b].cpu().numpy().view(np.uint32)
        assert np.array_equal(totals, expected[slot]), (frame, slot)
        other = worlds[1 - slot]
        got_other = np.concatenate([other[0].material_outputs(m)[1].download(np.uint32).reshape(-1, 5)[:, 1] for m in other[1]])
        assert np.array_equal(got_other, expected[1 - slot])
    sh.close()
    for w, _ in worlds:
        w.destroy()


@pytest.mark.parametrize("world,policy", [(1, "range"), (2, "strided"), (4, "strided"), (4, "range")])
def test_views_per_rank_equal_the_single_gpu_view_batch(ctx, world, policy):
    """Partition B: rank r culls its share of the views with ShardedViews; its per-view layouts must equal the corresponding
    views of one rve_cull_views batch over all views on a single GPU (which tests/test_gpu_views.py pins against the oracle)."""
    rng, model, layers, flags, cam, depth = _scene()
    lists = _lists(rng, 0, E)
    keep = []
    w, mats = _world(ctx, model, layers, flags, lists, keep)
    pyr = ctx.create_pyramid(180)
    d = ctx.buffer_from_numpy(depth)
    ctx.build_pyramid(pyr, d, depth.shape[1], depth.shape[0])
    V = 10
    cams = [scenes.default_camera(z_near=0.1, z_far=1000.0, eye=(40.0 * i - 200.0, 5.0, 30.0 * (i % 3)), target=(0.0, 0.0, -50.0)) for i in range(V)]
    views = [(c.view_proj_cm, c.pos, 0xFFFFFFFF if i % 4 else 0x1, i % 5 == 2) for i, c in enumerate(cams)]
    pyrs = [pyr] * V
    w.cull_static_views(views, pyrs)
    ctx.synchronize()
    full = []
    for m in mats:
        ids, ind, lods, ents = w.material_outputs(m)
        full.append((ind.download(np.uint32).reshape(V, lods, 5), ids.download(np.uint32).reshape(V, ents * lods), lods, ents))
    seen = []
    for r in range(world):
        sv = sharding.ShardedViews(w, views, pyrs, r, world, policy)
        assert sv.owned == sharding.assign_views(V, world, policy)[r]
        for i in range(len(sv.chunks)):
            owned = sv.cull_chunk(i)
            ctx.synchronize()
            for m, (find, fids, lods, ents) in zip(mats, full):
                ids, ind, _, _ = w.material_outputs(m)
                gi = ind.download(np.uint32).reshape(-1, lods, 5)[: len(owned)]
                gids = ids.download(np.uint32).reshape(-1, ents * lods)[: len(owned)]
                plan_starts = []
                off = 0
                for cmd_n, cmd_l in ((len(lists[j]), len(LODS[j])) for j in range(3) if mats[MAT_OF_CMD[j]] == m):
                    plan_starts += [(off + cmd_n * k) for k in range(cmd_l)]
                    off += cmd_n * cmd_l
                for v, g in enumerate(owned):
                    assert np.array_equal(gi[v], find[g])
                    for slot, start in enumerate(plan_starts):
                        n = int(gi[v, slot, 1])
                        assert np.array_equal(np.sort(gids[v, start:start + n]), np.sort(fids[g, start:start + n]))
            seen += owned
    assert sorted(seen) == list(range(V))
    w.destroy()
