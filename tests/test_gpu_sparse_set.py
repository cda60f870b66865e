"""Entity lists as the engine mutates them: the shim's BufferedVRAMSparseSet stand-in against a restatement
of the reference containers over random emplace / erase sequences, its private buffer after every sync, and a cull over a
churned list against the oracle."""
import math

import numpy as np
import pytest

import oracle
from oracle import host_tables
from oracle.host_tables import SparseSetRef
from ravengine_b200 import scenes
from ravengine_b200.vram_vector import SYNC_PACKED, SYNC_PULL, SYNC_RANGES, SparseSet
from ravengine_b200.world import World
from tests.helpers import make_ubo, oracle_cull

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("mode", [SYNC_RANGES, SYNC_PACKED, SYNC_PULL])
def test_sparse_set_matches_the_restated_reference_containers(ctx, mode):
    rng = np.random.default_rng(mode + 10)
    s, ref = SparseSet(ctx), SparseSetRef()
    universe = 5000
    for step in range(30):
        alive = np.array([i for i in range(universe) if ref.has(i)], np.uint32)
        n_erase = int(rng.integers(0, max(len(alive) // 3, 1) + 1)) if len(alive) else 0
        erase = rng.choice(alive, size=n_erase, replace=False) if n_erase else np.empty(0, np.uint32)
        add = rng.choice(universe, size=int(rng.integers(1, 400)), replace=False).astype(np.uint32)
        for i in erase:
            ref.erase(int(i))
        for i in add:  # emplace of an existing index is a no-op on both sides
            ref.emplace(int(i), int(i) ^ 0x5A5A)
        if step % 2:
            s.erase_many(erase)
            s.emplace_many(add, add ^ 0x5A5A)
        else:
            for i in erase:
                s.erase(int(i))
            for i in add:
                s.emplace(int(i), int(i) ^ 0x5A5A)
        assert len(s) == len(ref.dense)
        assert np.array_equal(s.dense(), np.array(ref.dense, np.uint32))
        assert np.array_equal(s.reverse_map(), np.array(ref.reverse, np.uint32))
        for probe in rng.integers(0, universe + 50, size=20):
            assert s.has(int(probe)) == ref.has(int(probe))
        s.encode_sync(mode, 16)
        ctx.synchronize()
        got = s.private_buffer.download(np.uint32)[: len(s)]
        assert np.array_equal(got, np.array(ref.dense, np.uint32))
    with pytest.raises(Exception):
        s.erase(universe + 7)  # the reference asserts
    s.destroy()


def test_cull_over_churned_entity_lists_matches_oracle(ctx):
    """Two commands whose entity lists went through swap-remove churn (10 % erased, re-inserted at the end): the cull walks
    the sets' private buffers; compared with the oracle over the same dense arrays, twice (a second churn between frames)."""
    rng = np.random.default_rng(4)
    n = 40_000
    model = scenes.random_transforms(rng, n, 200.0, 60.0)
    layers = np.full(n, 0xFFFFFFFF, np.uint32)
    flags = np.full(n, 0xFFFF, np.uint16)
    cam = scenes.default_camera(z_far=500.0)
    depth = scenes.synth_depth(512, 288, cam, n_occluders=30)
    pyr_h, dim, levels = oracle.pyramid_build(depth, 288)
    w = World(ctx)
    keep = [ctx.buffer_from_numpy(model), ctx.buffer_from_numpy(layers), ctx.buffer_from_numpy(flags)]
    w.set_entity_buffers(*keep)
    meshes = [scenes.simple_mesh(1.5, (0.0, 30.0, 90.0)), scenes.simple_mesh(2.5, (math.inf,), base=7000)]
    sets, mats = [], []
    for k, mesh in enumerate(meshes):
        s = SparseSet(ctx)
        s.emplace_many(np.arange(k * n // 2, (k + 1) * n // 2, dtype=np.uint32))
        sets.append(s)
        mat = w.material()
        mid = w.mesh_collection(mesh.radius, mesh.lod_distances, [l.index_count for l in mesh.lods], [l.index_start for l in mesh.lods],
                                [l.base_vertex for l in mesh.lods])
        w.add_command_set(mat, mid, s)
        mats.append(mat)
    pyr = ctx.create_pyramid(288)
    ctx.build_pyramid(pyr, ctx.buffer_from_numpy(depth), 512, 288)
    for frame in range(2):
        for k, s in enumerate(sets):  # churn: erase 10 % at random, re-insert them (they land at the end, out of order)
            alive = s.dense()
            gone = rng.choice(alive, size=len(alive) // 10, replace=False)
            s.erase_many(gone)
            s.emplace_many(rng.permutation(gone))
        lists = [s.dense() for s in sets]
        assert not np.array_equal(lists[0], np.sort(lists[0]))
        omats = [scenes.MaterialDraws([scenes.DrawCommand(mesh, l)]) for mesh, l in zip(meshes, lists)]
        t = host_tables.build_static_tables(omats)
        ref = oracle_cull(t, make_ubo(cam.view_proj_cm, cam.pos, t.num_cubos), model, layers, flags, pyr_h, dim, levels, nthreads=4)
        w.cull_static(cam.view_proj_cm, cam.pos, pyr)
        ctx.synchronize()
        ci = 0
        for mat, m in zip(mats, omats):
            cull_b, ind_b, lods, ents = w.material_outputs(mat)
            cb = t.cubos[ci]
            r_ind, r_ids = ref[int(cb["indirectOutputBufferBindlessHandle"])], ref[int(cb["idOutputBufferBindlessHandle"])]
            g_ind, g_ids = ind_b.download(np.uint32).reshape(-1, 5)[:lods], cull_b.download(np.uint32)[: lods * ents]
            np.testing.assert_array_equal(g_ind, r_ind)
            for x, y in zip(host_tables.sorted_slices(r_ind, r_ids, t.cubos, range(ci, ci + 1)), host_tables.sorted_slices(g_ind, g_ids, t.cubos, range(ci, ci + 1))):
                np.testing.assert_array_equal(x, y)
            ci += 1
    w.destroy()
    for s in sets:
        s.destroy()


@pytest.mark.parametrize("count,id_limit", [(1, 1), (1, 70_000), (33, 64), (5000, 5000), (100_000, 3_000_000), (1_000_003, 33_554_433)])
def test_sort_entity_ids_is_numpy_sort(ctx, count, id_limit):
    """rve_sort_entity_ids (bitmap pass: set bits, then block counts + look-back scan + expand in one launch) against np.sort,
    over id spaces from one bitmap word to several thousand look-back blocks; twice on the same context (the pass leaves its
    bitmap clean)."""
    rng = np.random.default_rng(count)
    for _ in range(2):
        ids = rng.choice(id_limit, size=count, replace=False).astype(np.uint32)
        src, dst = ctx.buffer_from_numpy(ids), ctx.create_buffer(4 * count)
        ctx.sort_entity_ids(src, count, id_limit, dst)
        ctx.synchronize()
        assert np.array_equal(dst.download(np.uint32)[:count], np.sort(ids))
        src.destroy()
        dst.destroy()


def test_sort_entity_ids_batch_is_numpy_sort_per_list(ctx):
    """rve_sort_entity_ids_batch: several lists of one id space in two launches per 16 lists (bits of every list set by one
    launch; counts + look-back scan + expansion by the other). Lists of very different lengths, an empty one, more than 16,
    ids past the limit (dropped), and batches of different geometry back to back (the scratch is re-zeroed when it changes)."""
    rng = np.random.default_rng(77)
    for id_limit, counts in ((3_000_000, [250_000, 1, 0, 40_000, 999_999]), (64, [3, 64]), (5_000_000, [20_000] * 19), (1_000, [1000]),
                             (3_000_000, [700_000, 5])):
        lists = [rng.choice(id_limit, size=c, replace=False).astype(np.uint32) for c in counts]
        if id_limit == 1_000:
            lists[0][::50] += 1_000  # 20 ids past the limit
        srcs = [ctx.buffer_from_numpy(l) if len(l) else ctx.create_buffer(16) for l in lists]
        outs = [ctx.create_buffer(max(4 * len(l), 16)) for l in lists]
        for o in outs:
            o.upload(np.full(o.nbytes // 4, 0xFFFFFFFF, np.uint32))
        l0 = ctx.launch_count
        ctx.sort_entity_ids_batch(srcs, counts, id_limit, outs)
        assert ctx.launch_count - l0 == 2 * ((sum(1 for c in counts if c) + 15) // 16 if len(counts) <= 16 else 2)
        ctx.synchronize()
        for l, o in zip(lists, outs):
            got = o.download(np.uint32)
            keep = np.sort(l[l < id_limit])
            assert np.array_equal(got[: len(keep)], keep)
            assert np.all(got[len(keep):] == 0xFFFFFFFF)  # nothing written past the ids that were kept
        for b in srcs + outs:
            b.destroy()


def test_sorted_entity_lists_give_the_same_cull_and_follow_list_changes(ctx):
    """RVS_SORT_ENTITY_LISTS: the cull reads an ascending copy of every list. Same oracle result as the unsorted world over
    churned lists, the copy is rebuilt exactly when a list changes (launch counts), for set-backed and buffer-backed
    commands."""
    from ravengine_b200.world import SORT_ENTITY_LISTS

    rng = np.random.default_rng(8)
    n = 30_000
    model = scenes.random_transforms(rng, n, 200.0, 60.0)
    layers = np.full(n, 0xFFFFFFFF, np.uint32)
    flags = np.full(n, 0xFFFF, np.uint16)
    cam = scenes.default_camera(z_far=500.0)
    depth = scenes.synth_depth(512, 288, cam, n_occluders=30)
    pyr_h, dim, levels = oracle.pyramid_build(depth, 288)
    w = World(ctx, SORT_ENTITY_LISTS)
    keep = [ctx.buffer_from_numpy(model), ctx.buffer_from_numpy(layers), ctx.buffer_from_numpy(flags)]
    w.set_entity_buffers(*keep)
    meshes = [scenes.simple_mesh(1.5, (0.0, 30.0, 90.0)), scenes.simple_mesh(2.5, (math.inf,), base=7000)]
    mids = [w.mesh_collection(m.radius, m.lod_distances, [l.index_count for l in m.lods], [l.index_start for l in m.lods],
                              [l.base_vertex for l in m.lods]) for m in meshes]
    s = SparseSet(ctx)
    s.emplace_many(rng.permutation(np.arange(0, n // 2, dtype=np.uint32)))
    raw_ids = rng.permutation(np.arange(n // 2, n, dtype=np.uint32))
    raw_buf = ctx.buffer_from_numpy(raw_ids)
    mats = [w.material(), w.material()]
    w.add_command_set(mats[0], mids[0], s)
    w.add_command(mats[1], mids[1], raw_buf, len(raw_ids))
    pyr = ctx.create_pyramid(288)
    ctx.build_pyramid(pyr, ctx.buffer_from_numpy(depth), 512, 288)

    def check():
        lists = [s.dense(), raw_ids]
        omats = [scenes.MaterialDraws([scenes.DrawCommand(mesh, l)]) for mesh, l in zip(meshes, lists)]
        t = host_tables.build_static_tables(omats)
        ref = oracle_cull(t, make_ubo(cam.view_proj_cm, cam.pos, t.num_cubos), model, layers, flags, pyr_h, dim, levels, nthreads=4)
        for ci, mat in enumerate(mats):
            cull_b, ind_b, lods, ents = w.material_outputs(mat)
            cb = t.cubos[ci]
            r_ind, r_ids = ref[int(cb["indirectOutputBufferBindlessHandle"])], ref[int(cb["idOutputBufferBindlessHandle"])]
            g_ind, g_ids = ind_b.download(np.uint32).reshape(-1, 5)[:lods], cull_b.download(np.uint32)[: lods * ents]
            np.testing.assert_array_equal(g_ind, r_ind)
            for x, y in zip(host_tables.sorted_slices(r_ind, r_ids, t.cubos, range(ci, ci + 1)), host_tables.sorted_slices(g_ind, g_ids, t.cubos, range(ci, ci + 1))):
                np.testing.assert_array_equal(x, y)

    l0 = ctx.launch_count
    w.cull_static(cam.view_proj_cm, cam.pos, pyr)
    ctx.synchronize()
    first = ctx.launch_count - l0
    check()
    l0 = ctx.launch_count
    w.cull_static(cam.view_proj_cm, cam.pos, pyr)  # nothing changed: no re-sort, no list upload
    ctx.synchronize()
    steady = ctx.launch_count - l0
    assert steady == 2 and first == steady + 4  # upload + cull; first frame: + one two-launch batch per kind of list (caller-owned: scratch bitmap; set-backed: its own bitmap)
    check()
    gone = rng.choice(s.dense(), size=1500, replace=False)  # the set changes: only its copy is rebuilt
    s.erase_many(gone)
    s.emplace_many(rng.permutation(gone)[:700])
    l0 = ctx.launch_count
    w.cull_static(cam.view_proj_cm, cam.pos, pyr)
    ctx.synchronize()
    assert 2 + 2 <= ctx.launch_count - l0 <= 2 + 2 + 2  # upload + cull + a one-list batch (+ the set's scatter kernels)
    check()
    raw_ids = rng.permutation(raw_ids)[:-300]  # the caller-owned list changes: set_command_entities announces it
    raw_buf2 = ctx.buffer_from_numpy(raw_ids)
    w.set_command_entities(mats[1], 0, raw_buf2, len(raw_ids))
    l0 = ctx.launch_count
    w.cull_static(cam.view_proj_cm, cam.pos, pyr)
    ctx.synchronize()
    assert ctx.launch_count - l0 == 2 + 2
    check()
    w.destroy()
    s.destroy()



def test_sorted_copy_of_a_set_backed_list_follows_per_frame_churn_by_deltas(ctx):
    """A sparse-set-backed command under RVS_SORT_ENTITY_LISTS, many frames with a few hundred erases / inserts each: after the
    first frame the ascending copy is maintained through the command's own bitmap - only the ids that entered / left the set are
    applied (rvs_sorted_list_stats), an id erased and re-inserted in one frame nets to nothing - and the cull matches the oracle
    in every frame. A burst that replaces more than a quarter of the list, and growth of the list, fall back to a full rebuild."""
    from ravengine_b200.vram_vector import SparseSet
    from ravengine_b200.world import SORT_ENTITY_LISTS

    rng = np.random.default_rng(12)
    n = 40_000
    model = scenes.random_transforms(rng, n, 200.0, 60.0)
    layers = np.full(n, 0xFFFFFFFF, np.uint32)
    flags = np.full(n, 0xFFFF, np.uint16)
    cam = scenes.default_camera(z_far=500.0)
    depth = scenes.synth_depth(512, 288, cam, n_occluders=30, seed=4)
    pyr_h, dim, levels = oracle.pyramid_build(depth, 288)
    mesh = scenes.simple_mesh(scenes.CUBE_RADIUS, (0.0, 40.0, 120.0))
    w = World(ctx, SORT_ENTITY_LISTS)
    w.set_entity_buffers(ctx.buffer_from_numpy(model), ctx.buffer_from_numpy(layers), ctx.buffer_from_numpy(flags))
    mid = w.mesh_collection(mesh.radius, mesh.lod_distances, [l.index_count for l in mesh.lods], [l.index_start for l in mesh.lods],
                            [l.base_vertex for l in mesh.lods])
    s = SparseSet(ctx)
    inside = rng.permutation(n)[:30_000].astype(np.uint32)
    s.emplace_many(inside)
    mat = w.material()
    w.add_command_set(mat, mid, s)
    pyr = ctx.create_pyramid(288)
    ctx.build_pyramid(pyr, ctx.buffer_from_numpy(depth), 512, 288)

    def check():
        w.cull_static(cam.view_proj_cm, cam.pos, pyr)
        ctx.synchronize()
        omats = [scenes.MaterialDraws([scenes.DrawCommand(mesh, s.dense())])]
        t = host_tables.build_static_tables(omats)
        ref = oracle_cull(t, make_ubo(cam.view_proj_cm, cam.pos, t.num_cubos), model, layers, flags, pyr_h, dim, levels, nthreads=4)
        cull_b, ind_b, lods, ents = w.material_outputs(mat)
        cb = t.cubos[0]
        r_ind, r_ids = ref[int(cb["indirectOutputBufferBindlessHandle"])], ref[int(cb["idOutputBufferBindlessHandle"])]
        g_ind, g_ids = ind_b.download(np.uint32).reshape(-1, 5)[:lods], cull_b.download(np.uint32)[: lods * ents]
        np.testing.assert_array_equal(g_ind, r_ind)
        for x, y in zip(host_tables.sorted_slices(r_ind, r_ids, t.cubos, range(1)), host_tables.sorted_slices(g_ind, g_ids, t.cubos, range(1))):
            np.testing.assert_array_equal(x, y)

    check()
    assert w.sorted_list_stats() == (1, 0)
    for frame in range(8):
        members = s.dense().copy()
        gone = rng.choice(members, size=300, replace=False)
        s.erase_many(gone)
        outside = np.setdiff1d(np.arange(n, dtype=np.uint32), s.dense())
        s.emplace_many(np.concatenate([rng.choice(outside, size=250, replace=False), gone[:50]]).astype(np.uint32))  # same size; 50 ids left and came back
        if frame % 2:
            s.encode_sync()  # the application syncs the set itself: the shim's own sync then finds nothing to do and must still see the change
        check()
        assert w.sorted_list_stats() == (1, frame + 1)
    s.emplace_many(rng.choice(np.setdiff1d(np.arange(n, dtype=np.uint32), s.dense()), size=500, replace=False).astype(np.uint32))  # the list grows
    check()
    full, delta = w.sorted_list_stats()
    assert (full, delta) in ((1, 9), (2, 8))  # a delta into the re-created copy buffer, or a rebuild: either is right
    members = s.dense().copy()
    s.erase_many(rng.choice(members, size=len(members) // 2, replace=False))  # a burst: more than a quarter of the list
    check()
    assert w.sorted_list_stats()[0] == full + 1
    w.destroy()
    s.destroy()
