"""SURVEY 8f-4 on the device: the arena's lists and buffer contents against oracle/mesh_arena.py over random
allocate / deallocate / growth sequences, and a skinned `.rvem` asset driving the skinning path end to end."""
import numpy as np
import pytest

import oracle
from oracle import host_tables
from oracle import mesh_arena as oma
from oracle.host_tables import SkinnedCommandDesc, SkinnedMeshDesc
from ravengine_b200 import SkinningUBO, scenes
from ravengine_b200.mesh_arena import (INDEX_ALLOCATED, INDEX_FREE, VERTEX_ALLOCATED, VERTEX_FREE, WEIGHT, MeshArena, write_rvem)
from ravengine_b200.world import World

pytestmark = pytest.mark.gpu


def _mesh(rng, nv, ni, lightmap=False):
    s = [rng.random((nv, w), dtype=np.float32) for w in (3, 3, 3, 3, 2)] + [rng.random((nv, 2), dtype=np.float32) if lightmap else None]
    return s, rng.integers(0, max(nv, 1), ni).astype(np.uint32)


def _lists_equal(arena, ref):
    for which, fam, kind in ((VERTEX_FREE, 0, "free"), (INDEX_FREE, 1, "free"), (VERTEX_ALLOCATED, 0, "allocated"), (INDEX_ALLOCATED, 1, "allocated")):
        want = [tuple(n) for n in getattr(ref, kind)[fam]]
        assert arena.list(which) == want, (which, arena.list(which), want)


@pytest.mark.parametrize("seed,init", [(0, (0, 0)), (1, (64, 96)), (2, (8, 8))])
def test_arena_matches_oracle_over_random_sequences(ctx, seed, init):
    rng = np.random.default_rng(100 + seed)
    arena = MeshArena(ctx, *init)
    ref = oma.Arena(init[0] or 1024, init[1] or 1536)
    live = []  # (device range handle, oracle vertex node, oracle index node)
    for step in range(60):
        if live and rng.random() < 0.3:
            h, vn, inode = live.pop(int(rng.integers(len(live))))
            # the oracle erases the first node with start >= and count >=, which may belong to another mesh: mirror
            # which handles die by looking at node identity afterwards
            arena.deallocate(h)
            ref.deallocate(vn, inode)
            live = [(hh, v, i) for hh, v, i in live if any(v is n for n in ref.allocated[0]) and any(i is n for n in ref.allocated[1])]
        else:
            nv, ni = int(rng.integers(1, 200)), int(rng.integers(0, 300))
            s, idx = _mesh(rng, nv, ni, lightmap=bool(step % 3 == 0))
            try:
                vn, inode = ref.allocate(s, idx)
            except RuntimeError:
                with pytest.raises(Exception):
                    arena.allocate(*s, indices=idx)
                break
            h = arena.allocate(*s, indices=idx)
            live.append((h, vn, inode))
        _lists_equal(arena, ref)
        for h, vn, inode in live:  # MeshRange follows re-packs
            assert arena.range(h) == (vn[0], vn[1], inode[0], inode[1])
    bufs, vcap, icap = arena.buffers()
    assert (vcap, icap) == tuple(ref.size)
    ctx.synchronize()
    for s in range(7):
        got = bufs[s].download(np.uint32)
        # compare what any allocation ever wrote or a re-pack moved; untouched elements of a fresh buffer are undefined
        # on the device (cudaMalloc) and zero in the oracle, so restrict to the allocated ranges
        fam = 0 if s < 6 else 1
        w = oma.STRIDE_WORDS[s]
        for start, count in ref.allocated[fam]:
            assert np.array_equal(got[start * w:(start + count) * w], ref.streams[s][start * w:(start + count) * w]), (s, start, count)
    arena.destroy()


def test_skinned_rvem_asset_drives_the_skinning_path(ctx):
    """write .rvem (skinned) -> rvs_arena_load_rvem -> MeshRange offsets -> shim skinning == oracle skinning on the
    arrays the file was made from. A static mesh loaded first shifts the skinned one inside the arena."""
    rng = np.random.default_rng(7)
    arena = MeshArena(ctx, 256, 512)
    filler, fidx = _mesh(rng, 77, 120)
    arena.load_rvem(write_rvem(*filler[:5], fidx))
    nv, ni, nb, n_obj = 150, 420, 6, 9
    s, idx = _mesh(rng, nv, ni)
    wts = scenes.skinning_inputs(rng, nv, 1, nb)["weights"]
    blob = write_rvem(*s[:5], idx, weights=wts.view(WEIGHT).reshape(nv, 4) if wts.dtype != WEIGHT else wts)
    h, wbuf = arena.load_rvem(blob, skinned=True)
    v0, vc, i0, ic = arena.range(h)
    assert (v0, vc, ic) == (77, nv, ni) and i0 == 120
    assert np.array_equal(wbuf.download(np.uint8), np.ascontiguousarray(wts).view(np.uint8).reshape(-1))
    bufs, vcap, _ = arena.buffers()
    assert vcap == 256  # 77 + 150 fit the initial arena
    n_ent = 40
    model = scenes.random_transforms(rng, n_ent, 30.0, 10.0)
    ids = rng.permutation(n_ent).astype(np.uint32)[:n_obj]
    pose = scenes.skinning_inputs(rng, 1, n_obj, nb)["pose"]
    mesh = SkinnedMeshDesc(nv, ni, i0, v0, 1.0)
    st = host_tables.build_skinned_tables([[SkinnedCommandDesc(mesh, nb, ids)]])
    # oracle: skin out of a host copy of the arena streams
    host_arena = [np.zeros((vcap, k), np.float32) for k in (3, 3, 3, 3, 2)]
    for k in range(5):
        host_arena[k][0:77] = filler[k]
        host_arena[k][v0:v0 + nv] = s[k]
    outs_ref = [np.zeros((nv * n_obj, k), np.float32) for k in (3, 3, 3, 3, 2)]
    oracle.skin(SkinningUBO(*st.skinning_ubos[0]), outs_ref, host_arena, pose, wts)
    w = World(ctx)
    keep = [ctx.buffer_from_numpy(model), ctx.buffer_from_numpy(np.full(n_ent, 0xFFFFFFFF, np.uint32)), ctx.buffer_from_numpy(np.full(n_ent, 0xFFFF, np.uint16))]
    w.set_entity_buffers(*keep)
    w.set_mesh_arena(bufs[:5])
    mh = w.skinned_mesh(nv, ni, i0, v0, 1.0, wbuf)
    mat = w.skinned_material()
    idb = ctx.buffer_from_numpy(ids)
    w.add_skinned_command(mat, mh, nb, idb, n_obj, pose)
    assert w.prepare_skeletal_mesh_buffers()
    w.pose_skeletal_meshes()
    ctx.synchronize()
    outs, _ = w.skinned_outputs()
    for b, r in zip(outs, outs_ref):
        np.testing.assert_array_equal(b.download(np.float32)[: r.size].reshape(r.shape), r)
    w.destroy()
    arena.destroy()
