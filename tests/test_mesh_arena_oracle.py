"""SURVEY 8f-4 on the CPU: hand-worked allocation sequences pinning oracle/mesh_arena.py to the list semantics of
src/RenderEngine_Allocate.cpp, the `.rvem` byte image (writer = product, reader = oracle, independent code), and the
host-only C-ABI parser."""
import numpy as np
import pytest

from oracle import mesh_arena as oma
from ravengine_b200 import RveError
from ravengine_b200.mesh_arena import (HAS_LIGHTMAP_UV, HAS_POSITIONS, SKINNED, WEIGHT, parse_rvem, write_rvem)


def _mesh(rng, nv, ni):
    return [rng.random((nv, w), dtype=np.float32) for w in (3, 3, 3, 3, 2)] + [None], rng.integers(0, max(nv, 1), ni).astype(np.uint32)


def test_power_of_1_5_growth_sizes():
    # 1.5^17 = 985.3, 1.5^18 = 1477.9, 1.5^19 = 2216.8, 1.5^20 = 3325.3: truncated to u32 as the reference's conversion does
    assert oma.closest_power_of_1_5(1034) == 1477
    assert oma.closest_power_of_1_5(1478) == 2216
    assert oma.closest_power_of_1_5(2217) == 3325
    assert oma.closest_power_of_1_5(3) == 3  # 1.5^3 = 3.375


def test_bump_allocation_exact_fit_and_growth():
    rng = np.random.default_rng(0)
    a = oma.Arena()
    s, i = _mesh(rng, 100, 300)
    v0, i0 = a.allocate(s, i)
    assert (v0, i0) == ([0, 100], [0, 300])
    assert a.free == [[[100, 924]], [[300, 1236]]]
    s, i = _mesh(rng, 924, 6)  # exact fit: the free range is erased, the list is empty
    v1, _ = a.allocate(s, i)
    assert v1 == [100, 924] and a.free[0] == []
    s, i = _mesh(rng, 10, 6)  # empty list -> growth to closest_power_of(1024 + 10) = 1477, new range appended
    v2, _ = a.allocate(s, i)
    assert a.size[0] == 1477 and v2 == [1024, 10] and a.free[0] == [[1034, 443]] and a.growths == 1
    assert a.size[1] == 1536  # the index family did not grow


def test_first_free_range_decides_and_holes_go_to_the_back():
    rng = np.random.default_rng(1)
    a = oma.Arena(64, 64)
    nodes = [a.allocate(*_mesh(rng, 10, 10)) for _ in range(3)]  # [0,10] [10,10] [20,10]; free [[30,34]]
    a.deallocate(*nodes[1])  # borders nothing that is free: appended behind the tail range
    assert a.free[0] == [[30, 34], [10, 10]]
    v, _ = a.allocate(*_mesh(rng, 10, 10))  # a perfect hole exists, but placement only ever looks at the first range
    assert v == [30, 10] and a.free[0] == [[40, 24], [10, 10]]
    a.deallocate(*nodes[2])  # [20,10] borders the hole [10,10] on its right side AND the tail? no: only the hole
    assert a.free[0] == [[40, 24], [10, 20]]
    a.deallocate(v, [30, 10])  # [30,10] borders the hole's end (10+20) and the tail's start (40): BOTH are widened (sic)
    assert a.free[0] == [[30, 34], [10, 30]]


def test_growth_repacks_allocated_ranges_and_keeps_stale_holes():
    rng = np.random.default_rng(2)
    a = oma.Arena(32, 32)
    m = [_mesh(rng, 8, 8) for _ in range(4)]
    nodes = [a.allocate(*x) for x in m]  # 4 x 8 = 32: exact fit, free list empty
    a.deallocate(*nodes[0])  # hole [0,8] is now the only free range
    big = _mesh(rng, 20, 4)
    v, _ = a.allocate(*big)  # does not fit the hole -> growth to cpo(32 + 20) = 57; allocated ranges re-packed from 0
    assert a.size[0] == 57
    assert [n for n in a.allocated[0]] == [[0, 8], [8, 8], [16, 8], [0, 20]]  # the new mesh lands on the stale hole, widened to the end
    assert a.free[0] == [[20, 37]]
    # the re-pack moved mesh 1's data to element 0, then the new mesh overwrote it: the reference's lists do exactly this
    w = 3
    assert np.array_equal(a.streams[0][0:20 * w], big[0][0].reshape(-1).view(np.uint32))


def test_rvem_image_round_trip_and_c_abi_parser():
    rng = np.random.default_rng(3)
    nv, ni = 37, 90
    s, idx = _mesh(rng, nv, ni)
    lm = rng.random((nv, 2), dtype=np.float32)
    wts = np.zeros((nv, 4), WEIGHT)
    wts["joint"] = rng.integers(0, 12, (nv, 4))
    wts["influence"] = rng.random((nv, 4), dtype=np.float32)
    for lightmap, skinned in ((None, None), (lm, None), (None, wts), (lm, wts)):
        blob = write_rvem(*s[:5], idx, lightmap_uvs=lightmap, weights=skinned)
        ref = oma.parse_rvem(blob)
        assert ref["num_vertices"] == nv and ref["num_indices"] == ni
        assert bool(ref["attributes"] & HAS_LIGHTMAP_UV) == (lightmap is not None) and bool(ref["attributes"] & SKINNED) == (skinned is not None)
        for k, name in enumerate(oma.ORDER[:5]):
            assert np.array_equal(ref[name], s[k])
        assert np.array_equal(ref["indices"], idx)
        info = parse_rvem(blob)  # rvs_rvem_parse through the C-ABI (host only)
        assert (info.num_vertices, info.num_indices, info.attributes) == (nv, ni, ref["attributes"])
        assert info.off_positions == 16 and info.off_normals == 16 + 12 * nv and info.off_uv0 == 16 + 48 * nv
        assert info.mesh_bytes == ref["mesh_bytes"] and info.off_indices == info.mesh_bytes - 4 * ni
        if lightmap is None:
            assert info.off_lightmap_uv == 2 ** 64 - 1 and ref["lightmap_uv"] is None
        else:
            assert np.array_equal(ref["lightmap_uv"], lm) and info.off_lightmap_uv == 16 + 56 * nv
        if skinned is None:
            assert ref["weights"] is None and info.off_weights == 2 ** 64 - 1
        else:
            assert np.array_equal(ref["weights"], wts) and info.off_weights == info.mesh_bytes and info.weights_bytes == 32 * nv


def test_rvem_rejects_what_the_reference_rejects():
    rng = np.random.default_rng(4)
    s, idx = _mesh(rng, 5, 6)
    blob = write_rvem(*s[:5], idx)
    with pytest.raises(RveError, match="not a mesh"):
        parse_rvem(b"rvex" + blob[4:])
    with pytest.raises(RveError, match="truncated"):
        parse_rvem(blob[:-4])
    with pytest.raises(RveError):
        parse_rvem(blob[:10])
    with pytest.raises(ValueError):
        oma.parse_rvem(b"mver" + blob[4:])
    # a header that announces only positions: the other streams are absent, the image is shorter
    head = bytearray(blob[:16])
    head[12] = HAS_POSITIONS
    short = bytes(head) + s[0].tobytes() + idx.tobytes()
    info = parse_rvem(short)
    assert info.off_normals == 2 ** 64 - 1 and info.off_indices == 16 + 60 and oma.parse_rvem(short)["normals"] is None


@pytest.mark.parametrize("name", ["static", "lightmap", "skinned"])
def test_rvem_bytes_pinned_by_the_reference_writer(name):
    """PINNED: tests/golden/rvem.npz holds the bytes the reference mesh compiler's own SerializeMeshPart
    (tools/rvemc/rvemc.cpp:233-292, compiled unchanged into oracle/_ref/libref_rvem.so) wrote for three assets. The Python
    writer must produce the same bytes, and rvs_rvem_parse (C-ABI) / the oracle parser must read the reference's file."""
    from pathlib import Path

    g = np.load(Path(__file__).parent / "golden" / "rvem.npz")
    blob = g[f"{name}_bytes"].tobytes()
    streams = [g[f"{name}_{k}"] for k in ("positions", "normals", "tangents", "bitangents", "uv0")]
    idx = g[f"{name}_indices"]
    lm = g[f"{name}_lightmap"] if f"{name}_lightmap" in g else None
    wts = g[f"{name}_weights"].view(WEIGHT).reshape(-1, 4) if f"{name}_weights" in g else None
    assert write_rvem(*streams, idx, lightmap_uvs=lm, weights=wts) == blob
    ref = oma.parse_rvem(blob)
    info = parse_rvem(blob)
    nv, ni = len(streams[0]), len(idx)
    assert (info.num_vertices, info.num_indices) == (nv, ni) == (ref["num_vertices"], ref["num_indices"])
    raw = np.frombuffer(blob, np.uint8)
    for k, (off, arr) in enumerate(zip((info.off_positions, info.off_normals, info.off_tangents, info.off_bitangents, info.off_uv0), streams)):
        assert np.array_equal(raw[off:off + arr.nbytes].view(np.float32).reshape(arr.shape), arr)
        assert np.array_equal(ref[oma.ORDER[k]], arr)
    assert np.array_equal(raw[info.off_indices:info.off_indices + 4 * ni].view(np.uint32), idx)
    assert info.mesh_bytes == info.off_indices + 4 * ni
    if lm is not None:
        assert np.array_equal(raw[info.off_lightmap_uv:info.off_lightmap_uv + lm.nbytes].view(np.float32).reshape(lm.shape), lm)
    else:
        assert info.off_lightmap_uv == 2 ** 64 - 1
    if wts is not None:
        assert info.off_weights == info.mesh_bytes and info.weights_bytes == 32 * nv == len(blob) - info.mesh_bytes
        assert np.array_equal(raw[info.off_weights:].view(np.uint32), wts.view(np.uint32).reshape(-1))
    else:
        assert info.off_weights == 2 ** 64 - 1 and len(blob) == info.mesh_bytes


def test_rvem_live_reference_writer_when_built():
    """where oracle/_ref/libref_rvem.so exists (this container), the committed fixture is what it writes today"""
    from pathlib import Path

    from tests.golden import make_golden_rvem as mg

    if not (Path(mg.ROOT) / "oracle" / "_ref" / "libref_rvem.so").exists():
        pytest.skip("reference writer not built here")
    g = np.load(Path(__file__).parent / "golden" / "rvem.npz")
    for name, (seed, nv, ni, lm, sk) in mg.CASES.items():
        assert mg.run_ref(mg.make_inputs(seed, nv, ni, lm, sk)) == g[f"{name}_bytes"].tobytes()
