"""Golden fixtures (tests/golden/*.npz, written by tests/golden/make_golden.py): the CPU tests pin the oracle to
them, the GPU tests pin the CUDA path to them. Bit-exact everywhere (id slices after sorting)."""
from pathlib import Path

import numpy as np
import pytest

import oracle
from oracle import host_tables
from ravengine_b200 import GridAssignUBO, GridBuildUBO, SkinningUBO, scenes
from tests.golden.make_golden import cull_materials
from tests.helpers import device_cull, make_ubo, oracle_cull

G = Path(__file__).resolve().parent / "golden"


def _load_cull():
    z = np.load(G / "cull_hiz_lod.npz")
    b = z["bounds"]
    ids_per_cmd = [z["ids"][b[i]:b[i + 1]] for i in range(len(b) - 1)]
    mats = cull_materials(ids_per_cmd, z["lods"], z["radii"], z["split"])
    t = host_tables.build_static_tables(mats)
    ubo = make_ubo(z["view_proj"], z["cam_pos"], t.num_cubos)
    return z, mats, t, ubo


def _check_cull(z, mats, t, out):
    indirect, slices, ci = [], [], 0
    for m in mats:
        cb = t.cubos[ci]
        h_c, h_i = int(cb["idOutputBufferBindlessHandle"]), int(cb["indirectOutputBufferBindlessHandle"])
        indirect.append(out[h_i])
        slices += host_tables.sorted_slices(out[h_i], out[h_c], t.cubos, range(ci, ci + len(m.commands)))
        ci += len(m.commands)
    np.testing.assert_array_equal(np.concatenate(indirect), z["indirect"])
    np.testing.assert_array_equal(np.array([len(s) for s in slices]), z["slice_sizes"])
    np.testing.assert_array_equal(np.concatenate(slices), z["slices"])


def test_oracle_cull_matches_golden():
    z, mats, t, ubo = _load_cull()
    pyr, dim, levels = oracle.pyramid_build(z["depth"], 144)
    np.testing.assert_array_equal(pyr, z["pyramid"])
    assert (dim, levels) == (int(z["pyr_dim"]), int(z["pyr_levels"]))
    for nthreads, tf in ((1, False), (4, False), (3, True)):
        _check_cull(z, mats, t, oracle_cull(t, ubo, z["model"], z["layers"], z["flags"], pyr, dim, levels, nthreads=nthreads, taskflow=tf))


def _skin_io():
    z = np.load(G / "skinning.npz")
    ins = [z[k] for k in ("pos", "nrm", "tan", "bit", "uv")]
    want = [z[k] for k in ("out_pos", "out_nrm", "out_tan", "out_bit", "out_uv")]
    nv, nb = ins[0].shape[0], 8
    return z, ins, want, SkinningUBO(want[0].shape[0] // nv, nv, nb, 0, 0, 0)


def test_oracle_skinning_matches_golden():
    z, ins, want, ubo = _skin_io()
    outs = [np.zeros_like(w) for w in want]
    oracle.skin(ubo, outs, ins, z["pose"], z["weights"], nthreads=2)
    for o, w in zip(outs, want):
        np.testing.assert_array_equal(o, w)
    # north_star tolerance: 1e-5 relative against a float64 evaluation
    f64 = [np.zeros(w.shape, np.float64) for w in want[:4]]
    oracle.skin_f64(ubo, f64, ins, z["pose"], z["weights"])
    for r, w in zip(f64, want[:4]):
        assert np.abs(w - r).max() / np.abs(r).max() <= 1e-5


def _cluster_io():
    z = np.load(G / "clusters.npz")
    u = GridBuildUBO()
    u.invProj[:] = [float(x) for x in z["inv_proj"]]
    u.gridSize[:] = [int(x) for x in z["grid"]]
    u.zNear, u.zFar = float(z["z_near"]), float(z["z_far"])
    u.screenDim[:] = [int(x) for x in z["screen"]]
    a = GridAssignUBO()
    a.viewMat[:] = [float(x) for x in z["view"]]
    a.pointLightCount, a.spotLightCount = len(z["point_pos"]), len(z["spot_pos"])
    points = scenes.point_lights_bytes(z["point_pos"], z["point_int"])
    spots = scenes.spot_lights_bytes(z["spot_pos"], z["spot_int"])
    return z, u, a, points, spots


def test_oracle_clusters_match_golden():
    z, u, a, points, spots = _cluster_io()
    cl = np.zeros_like(z["clusters"])
    oracle.cluster_build(u, cl)
    oracle.cluster_assign(a, cl, points, spots, cl.shape[0], nthreads=2)
    np.testing.assert_array_equal(cl, z["clusters"])


@pytest.mark.gpu
def test_cuda_cull_matches_golden(ctx):
    z, mats, t, ubo = _load_cull()
    got, pyr_dev = device_cull(ctx, t, ubo, z["model"], z["layers"], z["flags"], depth=z["depth"], pyr_width=144)
    np.testing.assert_array_equal(pyr_dev, z["pyramid"])
    _check_cull(z, mats, t, got)


@pytest.mark.gpu
def test_cuda_skinning_matches_golden(ctx):
    z, ins, want, ubo = _skin_io()
    d_outs = [ctx.create_buffer(w.nbytes) for w in want]
    ctx.skin(ubo, d_outs, [ctx.buffer_from_numpy(a) for a in ins], ctx.buffer_from_numpy(z["pose"]), ctx.buffer_from_numpy(z["weights"]))
    ctx.synchronize()
    for b, w in zip(d_outs, want):
        np.testing.assert_array_equal(b.download(np.float32).reshape(w.shape), w)


@pytest.mark.gpu
def test_cuda_clusters_match_golden(ctx):
    z, u, a, points, spots = _cluster_io()
    d = ctx.buffer_from_numpy(np.zeros_like(z["clusters"]))
    ctx.cluster_build(u, d)
    ctx.cluster_assign(a, d, ctx.buffer_from_numpy(points), ctx.buffer_from_numpy(spots), z["clusters"].shape[0])
    ctx.synchronize()
    np.testing.assert_array_equal(d.download(np.uint32).reshape(z["clusters"].shape), z["clusters"])
