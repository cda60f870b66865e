"""CPU tests of the host-side table construction restated in oracle/host_tables.py against values worked by hand
from src/RenderEngine_Draw.cpp:1223-1373 (static) and :546-574, 597-625, 1067-1218 (skinned), plus oracle
self-consistency (serial == threaded == taskflow)."""
import math

import numpy as np

import oracle
from oracle import host_tables
from oracle.host_tables import SkinnedCommandDesc, SkinnedMeshDesc
from ravengine_b200 import scenes
from tests.helpers import make_ubo, oracle_cull


def _mats():
    a = scenes.simple_mesh(1.5, (0.0, 10.0, 50.0), base=0)  # 3 LODs
    b = scenes.simple_mesh(2.5, (math.inf,), base=7000)  # 1 LOD
    c = scenes.simple_mesh(0.5, (0.0, 20.0), base=9000)  # 2 LODs
    ids = np.arange(100, dtype=np.uint32)
    m0 = scenes.MaterialDraws([scenes.DrawCommand(a, ids[0:10]), scenes.DrawCommand(b, ids[10:14])])
    m1 = scenes.MaterialDraws([scenes.DrawCommand(c, ids[14:20]), scenes.DrawCommand(a, ids[20:25]), scenes.DrawCommand(b, ids[25:25])])
    return [m0, m1], (a, b, c)


def test_static_tables_hand_checked():
    mats, (a, b, c) = _mats()
    t = host_tables.build_static_tables(mats)
    # numCubos counts LODs (:1238), not commands: (3+1) + (2+3+1) = 10; only 5 entries are written
    assert t.num_cubos == 10 and t.total_entities == 25
    assert (t.cubos["numObjects"][5:] == 0).all()
    cb = t.cubos
    # material 0: offsets restart at 0; cullingBufferOffset advances by lods*entities (:1358-1359)
    assert cb["indirectBufferOffset"][:2].tolist() == [0, 3] and cb["cullingBufferOffset"][:2].tolist() == [0, 30]
    assert cb["numLODs"][:2].tolist() == [3, 1] and cb["numObjects"][:2].tolist() == [10, 4]
    assert cb["radius"][:2].tolist() == [1.5, 2.5]
    # material 1
    assert cb["indirectBufferOffset"][2:5].tolist() == [0, 2, 5] and cb["cullingBufferOffset"][2:5].tolist() == [0, 12, 27]
    assert cb["idOutputBufferBindlessHandle"][0] == cb["idOutputBufferBindlessHandle"][1] != cb["idOutputBufferBindlessHandle"][2]
    # culling buffer = numEntities * numLODs slots per material (:1288): 14*4, 11*6
    assert t.culling_slots == [56, 66]
    # indirect init (:1296-1320): baseInstance runs over (command, LOD) pairs in steps of the command's entity count
    i0 = t.indirect_init[0]
    assert i0[:, 4].tolist() == [0, 10, 20, 30] and (i0[:, 1] == 0).all()
    assert i0[:, 0].tolist() == [l.index_count for l in a.lods] + [b.lods[0].index_count]
    assert i0[:, 2].tolist() == [l.index_start for l in a.lods] + [b.lods[0].index_start]
    assert t.indirect_init[1][:, 4].tolist() == [0, 6, 12, 17, 22, 27]


def test_static_tables_reference_slot_quirk():
    """(meshID + lodID) slot arithmetic of :1314 collides when a material holds several multi-LOD collections."""
    mats, (a, b, c) = _mats()
    q = host_tables.build_static_tables(mats, quirk_slot=True)
    # material 0: command 0 writes slots 0,1,2; command 1 (meshID 1) writes slot 1 -> overwrites LOD 1 of command 0
    i0 = q.indirect_init[0]
    assert i0[1, 0] == b.lods[0].index_count and i0[1, 4] == 30
    assert (i0[3] == 0).all()  # the slot the shader reads for command 1 is never initialised
    # a material with one collection is unaffected
    one = [scenes.MaterialDraws([scenes.DrawCommand(a, np.arange(5, dtype=np.uint32))])]
    np.testing.assert_array_equal(host_tables.build_static_tables(one, True).indirect_init[0], host_tables.build_static_tables(one).indirect_init[0])


def test_skinned_tables_hand_checked():
    mesh_a = SkinnedMeshDesc(num_verts=100, num_indices=300, index_start=5000, vertex_start=40, radius=1.0)
    mesh_b = SkinnedMeshDesc(num_verts=30, num_indices=60, index_start=9000, vertex_start=400, radius=2.0)
    m0 = [SkinnedCommandDesc(mesh_a, 8, np.array([3, 4, 5], np.uint32)), SkinnedCommandDesc(mesh_b, 4, np.array([9, 10], np.uint32))]
    m1 = [SkinnedCommandDesc(mesh_b, 4, np.arange(20, 90, dtype=np.uint32))]
    t = host_tables.build_skinned_tables([m0, m1])
    # :1081-1123 — every command restarts at slot 0 (meshID is loop-local), baseVertex keeps running, also across materials
    r0 = t.indirect_reset[0]
    assert r0.shape == (5, 5)
    assert r0[:2].tolist() == [[60, 0, 9000, 300, 0], [60, 0, 9000, 330, 1]]  # command 1 overwrote slots 0,1
    assert r0[2].tolist() == [300, 0, 5000, 200, 2]  # command 0's third entity survives
    assert (r0[3:] == 0).all()
    assert t.indirect_reset[1][0].tolist() == [60, 0, 9000, 360, 0] and t.indirect_reset[1][69, 3] == 360 + 69 * 30
    # :1136-1187 — one CUBO per command, numLODs 1, offsets advance by 1 and by the entity count
    c0 = t.cubos[0]
    assert c0["indirectBufferOffset"].tolist() == [0, 1] and c0["cullingBufferOffset"].tolist() == [0, 3] and c0["numLODs"].tolist() == [1, 1]
    # :1214 — dispatch sized by the LAST command only: ceil(2/64) = 1 group even though 5 entities are listed
    assert t.dispatch_groups == [1, 2]
    # :552-570 — SkinningPrepareUBO per command, offsets accumulate within a material, restart per material
    assert t.prepare_ubos[0] == [(5000, 0, 300, 100, 3, 0, 0), (9000, 100, 60, 30, 2, 3, 3)]
    assert t.prepare_ubos[1] == [(9000, 0, 60, 30, 70, 0, 0)]
    # :597-625 — SkinningUBO offsets run over all commands of all materials
    assert t.skinning_ubos == [(3, 100, 8, 0, 0, 40), (2, 30, 4, 24, 300, 400), (70, 30, 4, 32, 360, 400)]


def test_oracle_serial_threads_taskflow_agree():
    rng = np.random.default_rng(4)
    n = 30_000
    model = scenes.random_transforms(rng, n, 300.0, 40.0)
    layers = np.full(n, 0xFFFFFFFF, np.uint32)
    flags = np.full(n, 0xFFFF, np.uint16)
    ids = rng.permutation(n).astype(np.uint32)
    mats = [scenes.MaterialDraws([scenes.DrawCommand(scenes.simple_mesh(1.7, (0.0, 40.0, 120.0)), ids[: n // 2])]),
            scenes.MaterialDraws([scenes.DrawCommand(scenes.simple_mesh(1.2), ids[n // 2:])])]
    cam = scenes.default_camera(z_far=1000.0)
    depth = scenes.synth_depth(320, 180, cam, n_occluders=20)
    pyr, dim, levels = oracle.pyramid_build(depth, 180)
    t = host_tables.build_static_tables(mats)
    ubo = make_ubo(cam.view_proj_cm, cam.pos, t.num_cubos)
    runs = [oracle_cull(t, ubo, model, layers, flags, pyr, dim, levels, nthreads=nt, taskflow=tf) for nt, tf in ((1, False), (6, False), (6, True))]
    ref = runs[0]
    for other in runs[1:]:
        ci = 0
        for m in mats:
            cb = t.cubos[ci]
            h_c, h_i = int(cb["idOutputBufferBindlessHandle"]), int(cb["indirectOutputBufferBindlessHandle"])
            np.testing.assert_array_equal(ref[h_i], other[h_i])
            for x, y in zip(host_tables.sorted_slices(ref[h_i], ref[h_c], t.cubos, range(ci, ci + 1)),
                            host_tables.sorted_slices(other[h_i], other[h_c], t.cubos, range(ci, ci + 1))):
                np.testing.assert_array_equal(x, y)
            ci += len(m.commands)
    vis = sum(int(b[:, 1].sum()) for b in ref if b is not None and b.ndim == 2)
    assert 0 < vis < n


def test_gen_instances_is_pure_function_of_index():
    """orc_gen_instances(seed, global index): any shard of the C5 scene can be generated independently."""
    from ravengine_b200 import SceneGenParams

    p = SceneGenParams(77, 500.0, 50.0, 0.5, 2.0, 0.05, 0.05, 0)
    def gen(first, count):
        m, l = np.zeros((count, 16), np.float32), np.zeros(count, np.uint32)
        f, i = np.zeros(count, np.uint16), np.zeros(count, np.uint32)
        oracle.gen_instances(p, first, count, m, l, f, i)
        return m, l, f
    whole = gen(1000, 600)
    a, b = gen(1000, 250), gen(1250, 350)
    for w, x, y in zip(whole, a, b):
        np.testing.assert_array_equal(w, np.concatenate([x, y]))
    m = whole[0]
    assert np.all(m[:, 15] == 1.0) and np.all(np.abs(m[:, 12]) <= 500.0) and np.all(np.abs(m[:, 13]) <= 50.0)
    s = np.linalg.norm(m[:, 0:3], axis=1)
    assert s.min() >= 0.49 and s.max() <= 2.01


def test_restated_sparse_set_known_answers():
    """SparseSetRef (the checker of tests/test_gpu_sparse_set.py) on a hand-worked sequence: swap-remove moves the last
    dense element into the hole, reverse map and sparse map follow, flags mark the touched dense slots
    (include/RavEngine/SparseSet.hpp:22-47, include/RavEngine/BufferedVRAMVector.hpp:153-163)."""
    from oracle.host_tables import SparseSetRef

    s = SparseSetRef()
    for i in (5, 2, 9, 7):
        s.emplace(i, i * 10)
    assert s.dense == [50, 20, 90, 70] and s.reverse == [5, 2, 9, 7]
    assert len(s.sparse) == 10 and s.sparse[5] == 0 and s.sparse[7] == 3 and s.sparse[0] == SparseSetRef.INVALID
    s.flags = [False] * 4
    s.erase(2)  # dense slot 1 <- last element (70, owner 7)
    assert s.dense == [50, 70, 90] and s.reverse == [5, 7, 9] and s.sparse[7] == 1 and not s.has(2)
    assert s.flags == [False, True, False, False]
    s.erase(9)  # the last element itself: no move
    assert s.dense == [50, 70] and s.reverse == [5, 7] and s.flags[2]
    s.emplace(5, 123)  # already present: no-op
    assert s.dense == [50, 70]
    s.emplace(2, 21)
    assert s.dense == [50, 70, 21] and s.sparse[2] == 2 and s.flags[2]
