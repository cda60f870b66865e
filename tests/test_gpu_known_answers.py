"""The hand-worked known-answer cases of tests/cases.py through the CUDA path (C-ABI), checked against the
expected answers directly AND against the oracle, plus the small stage known answers."""
import numpy as np
import pytest

import oracle
from oracle import host_tables
from ravengine_b200 import GridAssignUBO, SkinningUBO, scenes
from tests import cases
from tests.helpers import compare_cull_outputs, device_cull, make_ubo, oracle_cull

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("name", sorted(cases.CULL_CASES))
def test_cull_known_answer_gpu(ctx, name):
    c = cases.CULL_CASES[name]()
    t = host_tables.build_static_tables(c.materials)
    ubo = make_ubo(c.view_proj, c.cam_pos, t.num_cubos, layers=c.cam_layers, shadow=c.shadow)
    ref = oracle_cull(t, ubo, c.model, c.layers, c.flags, c.pyramid, c.pyr_dim, c.pyr_levels)
    got, _ = device_cull(ctx, t, ubo, c.model, c.layers, c.flags, pyr_width=c.pyr_dim, pyramid_texels=c.pyramid)
    h_ind = int(t.cubos[0]["indirectOutputBufferBindlessHandle"])
    h_ids = int(t.cubos[0]["idOutputBufferBindlessHandle"])
    assert got[h_ind][:, 1].tolist() == c.expect_counts, name
    if c.expect_ids is not None:
        for start, ids in c.expect_ids:
            assert sorted(got[h_ids][start:start + len(ids)].tolist()) == sorted(ids), name
    compare_cull_outputs(t, c.materials, ref, got)


def test_cluster_tangent_and_cap_gpu(ctx):
    ident = np.eye(4, dtype=np.float32).reshape(16)
    cl = np.zeros((1, 40), np.uint32)
    cl[0, 0:3] = np.float32([0, 0, -2]).view(np.uint32)
    cl[0, 3:6] = np.float32([1, 1, -1]).view(np.uint32)
    cl[0, 6:] = 0xDEADBEEF
    i4 = np.float32(0.01) * np.float32(4.0)
    pos = [(3.0, 0.5, -1.5), (3.0 + 2.0 ** -20, 0.5, -1.5)] + [(0.5, 0.5, -1.5)] * 18
    pts = scenes.point_lights_bytes(np.asarray(pos, np.float32), np.full(20, i4, np.float32))
    spots = scenes.spot_lights_bytes(np.tile(np.float32([0.5, 0.5, -1.5]), (20, 1)), np.full(20, 1.0, np.float32))
    a = GridAssignUBO()
    a.viewMat[:] = [float(x) for x in ident]
    a.pointLightCount, a.spotLightCount = 20, 20
    ref = cl.copy()
    oracle.cluster_assign(a, ref, pts, spots, 1)
    d = ctx.buffer_from_numpy(cl)
    ctx.cluster_assign(a, d, ctx.buffer_from_numpy(pts), ctx.buffer_from_numpy(spots), 1)
    ctx.synchronize()
    got = d.download(np.uint32).reshape(1, 40)
    assert got[0, 6] == 16 and got[0, 7] == 16
    assert got[0, 8:24].tolist() == [0] + list(range(2, 17))
    assert got[0, 24:40].tolist() == list(range(16))
    np.testing.assert_array_equal(ref, got)


def test_skinning_dyadic_identity_gpu(ctx):
    rng = np.random.default_rng(1)
    d = scenes.skinning_inputs(rng, 100, 3, 8)
    d["pose"][:] = np.eye(4, dtype=np.float32).reshape(16)
    d["weights"][:, 1::2] = np.array([0.5, 0.25, 0.125, 0.125], np.float32).view(np.uint32)
    ins = [d["pos"], d["nrm"], d["tan"], d["bit"], d["uv"]]
    d_outs = [ctx.create_buffer(300 * k * 4) for k in (3, 3, 3, 3, 2)]
    d_ins = [ctx.buffer_from_numpy(a) for a in ins]
    ctx.skin(SkinningUBO(3, 100, 8, 0, 0, 0), d_outs, d_ins, ctx.buffer_from_numpy(d["pose"]), ctx.buffer_from_numpy(d["weights"]))
    ctx.synchronize()
    for b, i, k in zip(d_outs, ins, (3, 3, 3, 3, 2)):
        np.testing.assert_array_equal(b.download(np.float32).reshape(300, k), np.tile(i, (3, 1)))
