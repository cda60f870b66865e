"""world_size-2 `gloo` test (CPU) of the N>1 path's host logic: every rank culls its instance range with
the oracle standing in for the kernel, per-slot counts are all-gathered, the exclusive scan over ranks gives the
offsets, the compacted slices are gathered into the global layout — and the result must equal a single-process cull
of the whole scene (counts exactly, id slices after sorting)."""
import math
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from ravengine_b200 import scenes, sharding
from ravengine_b200.sharding import SlotPlan

E = 9000  # entities
LODS = [(0.0, 40.0, 120.0), (math.inf,), (0.0, 60.0)]
RADII = [1.7, 2.2, 0.9]
MAT_OF_CMD = [0, 0, 1]
FRACTIONS = [0.55, 0.12, 0.33]


def _scene():
    rng = np.random.default_rng(77)
    model = scenes.random_transforms(rng, E, 300.0, 40.0)
    layers = np.full(E, 0xFFFFFFFF, np.uint32)
    layers[rng.random(E) < 0.05] = 0
    flags = np.full(E, 0xFFFF, np.uint16)
    flags[rng.random(E) < 0.1] &= 0xFFFC
    cam = scenes.default_camera(z_far=1000.0)
    depth = scenes.synth_depth(320, 180, cam, n_occluders=20)
    return rng, model, layers, flags, cam, depth


def _rank_lists(rng, world):
    """per rank: the three commands' GLOBAL id lists, all inside the rank's contiguous entity range"""
    out = []
    for lo, hi in sharding.split_range(E, world):
        ids = (lo + rng.permutation(hi - lo)).astype(np.uint32)
        n = hi - lo
        cuts = [0, int(n * FRACTIONS[0]), int(n * (FRACTIONS[0] + FRACTIONS[1])), n]
        out.append([ids[cuts[i]:cuts[i + 1]] for i in range(3)])
    return out


def _materials(lists):
    meshes = [scenes.simple_mesh(RADII[i], LODS[i], base=1000 * i) for i in range(3)]
    mats = [scenes.MaterialDraws([]), scenes.MaterialDraws([])]
    for i in range(3):
        mats[MAT_OF_CMD[i]].commands.append(scenes.DrawCommand(meshes[i], lists[i]))
    return mats


def _plans(lists):
    return [SlotPlan((len(lists[0]), len(lists[1])), (3, 1)), SlotPlan((len(lists[2]),), (2,))]


def _cull(mats, model, layers, flags, cam, depth):
    import oracle
    from oracle import host_tables
    from tests.helpers import make_ubo, oracle_cull

    t = host_tables.build_static_tables(mats)
    pyr, dim, levels = oracle.pyramid_build(depth, 180)
    out = oracle_cull(t, make_ubo(cam.view_proj_cm, cam.pos, t.num_cubos), model, layers, flags, pyr, dim, levels)
    res, ci = [], 0
    for m in mats:
        cb = t.cubos[ci]
        res.append((out[int(cb["indirectOutputBufferBindlessHandle"])], out[int(cb["idOutputBufferBindlessHandle"])]))
        ci += len(m.commands)
    return res


def _worker(rank, world, port, q):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        rng, model, layers, flags, cam, depth = _scene()
        lists = _rank_lists(rng, world)
        lo, hi = sharding.split_range(E, world)[rank]
        # this rank's slice: arrays indexed by LOCAL id = global id - lo
        local_lists = [l - lo for l in lists[rank]]
        local = _cull(_materials(local_lists), model[lo:hi], layers[lo:hi], flags[lo:hi], cam, depth)
        local_counts = torch.from_numpy(np.concatenate([ind[:, 1] for ind, _ in local]).astype(np.int32))
        counts_all = sharding.exchange_counts(local_counts, world)
        # the exclusive scan over ranks (device kernel rve_shard_offsets in production; numpy here)
        c = counts_all.numpy().astype(np.int64)
        offsets = torch.from_numpy((np.cumsum(c, axis=0) - c).astype(np.int32))
        totals = c.sum(axis=0)
        bases = [r[0] for r in sharding.split_range(E, world)]
        global_lists = [np.concatenate([lists[r][i] for r in range(world)]) for i in range(3)]
        # the device-side gather (rve_gather_ids_p2p) emulated with numpy from the same plan: every rank's slot slices land
        # in the consumer's global buffers at dst_start[s] + offsets[rank][s], ids shifted by the rank's id base
        plans_by_rank = [None] * world
        dist.all_gather_object(plans_by_rank, (sharding.gather_plan(_plans(local_lists), _plans(global_lists)), [ids for _, ids in local]))
        gathered = None
        if rank == 0:
            gp0 = plans_by_rank[0][0]
            gathered = [np.full(max(n, 1), 0xFFFFFFFF, np.uint32) for n in gp0.global_slots]
            offs = offsets.numpy().astype(np.int64)
            for r, (gp, ids_r) in enumerate(plans_by_rank):
                assert gp.dst_start == gp0.dst_start and gp.slot_buffer == gp0.slot_buffer
                for s_, (b, src, dst) in enumerate(zip(gp.slot_buffer, gp.src_start, gp.dst_start)):
                    n = int(c[r, s_])
                    gathered[b][dst + offs[r, s_]: dst + offs[r, s_] + n] = ids_r[b][src:src + n] + np.uint32(bases[r])
            gathered = [torch.from_numpy(g.view(np.int32)) for g in gathered]
        if rank == 0:
            ref = _cull(_materials(global_lists), model, layers, flags, cam, depth)
            ref_counts = np.concatenate([ind[:, 1] for ind, _ in ref]).astype(np.int64)
            ok = bool((ref_counts == totals).all())
            s = 0
            for m, plan in enumerate(_plans(global_lists)):
                got = gathered[m].numpy().view(np.uint32)
                for start in plan.slot_starts():
                    n = int(totals[s])
                    ok = ok and bool((np.sort(got[start:start + n]) == np.sort(ref[m][1][start:start + n])).all())
                    s += 1
                ok = ok and int((got != 0xFFFFFFFF).sum()) == int((ref[m][1] != 0xFFFFFFFF).sum())
            q.put((ok, int(totals.sum()), totals.tolist()))
    finally:
        dist.destroy_process_group()


def test_gather_plan_view_split_and_2d_grid():
    lp = [SlotPlan((5, 2), (3, 1)), SlotPlan((3,), (2,))]
    gp = [SlotPlan((10, 4), (3, 1)), SlotPlan((7,), (2,))]
    g = sharding.gather_plan(lp, gp)
    assert g.slot_buffer == (0, 0, 0, 0, 1, 1)
    assert g.src_start == (0, 5, 10, 15, 0, 3) and g.dst_start == (0, 10, 20, 30, 0, 7)
    assert g.global_slots == (56, 14)
    assert sharding.split_views(512, 8)[3] == (192, 256) and sharding.split_views(5, 2) == [(0, 2), (2, 5)]
    assert sharding.assign_views(7, 3) == [[0, 3, 6], [1, 4], [2, 5]] and sharding.assign_views(5, 2, "range") == [[0, 1], [2, 3, 4]]
    assert sorted(v for r in sharding.assign_views(512, 8) for v in r) == list(range(512))
    grid = sharding.Grid2D(8, 2)  # 2 view groups x 4 instance groups
    assert grid.instance_groups == 4 and grid.coords(5) == (1, 1) and grid.peers(5) == [4, 5, 6, 7]
    assert grid.owned(5, 512, 1_000_000_000) == ((256, 512), (250_000_000, 500_000_000))
    # every (view, instance) pair is owned exactly once
    cover = np.zeros((16, 40), np.int32)
    for r in range(8):
        (v0, v1), (i0, i1) = grid.owned(r, 16, 40)
        cover[v0:v1, i0:i1] += 1
    assert (cover == 1).all()
    with pytest.raises(AssertionError):
        sharding.Grid2D(8, 3)


def test_split_range_and_slot_plan():
    assert sharding.split_range(10, 4) == [(0, 2), (2, 5), (5, 7), (7, 10)]
    assert sum(hi - lo for lo, hi in sharding.split_range(1_000_000_007, 8)) == 1_000_000_007
    p = SlotPlan((10, 4), (3, 1))
    assert p.num_slots == 4 and p.slot_starts() == [0, 10, 20, 30] and p.culling_slots == 56
    plans = sharding.shard_plans([(10, 4), (7,)], [(3, 1), (2,)], 2)
    assert plans[0][0].counts == (5, 2) and plans[1][0].counts == (5, 2) and plans[0][1].counts == (3,) and plans[1][1].counts == (4,)


@pytest.mark.timeout(300)
def test_sharded_cull_world_size_2_gloo():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    ok, visible, totals = q.get(timeout=240)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert ok, totals
    assert 0 < visible < E
