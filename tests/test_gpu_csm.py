"""rve_csm_cascades against the glm-produced golden vectors and the numpy oracle: bit-exact."""
from pathlib import Path

import numpy as np
import pytest

from oracle import csm as ocsm
from ravengine_b200.api import CSM_CAMERA, CSM_LIGHT, csm_camera_params, csm_cascades
from tests.golden import make_golden_csm as mg
from tests.test_csm_oracle import cams_from_ref, lights_from_ref

pytestmark = pytest.mark.gpu
GOLD = Path(__file__).parent / "golden" / "csm.npz"


def _device_inputs(cams_ref, lights_ref):
    cams = np.zeros(len(cams_ref), CSM_CAMERA)
    for i, c in enumerate(cams_ref):
        cams[i]["viewOnly"], cams[i]["zNear"], cams[i]["zFar"] = c["viewOnly"], c["zNear"], c["zFar"]
        cams[i]["tanHalfFovy"], cams[i]["aspect"] = csm_camera_params(float(c["fov"]), int(c["w"]), int(c["h"]))
    lights = np.zeros(len(lights_ref), CSM_LIGHT)
    for name in ("direction", "castsShadows", "numCascades", "shadowCascades"):
        lights[name] = lights_ref[name]
    lights["maxShadowDistanceFactor"] = lights_ref["factor"]
    return cams, lights


def _run(ctx, cams, lights, with_aux=True):
    slots = len(cams) * len(lights)
    bc, bl = ctx.buffer_from_numpy(cams.view(np.uint8)), ctx.buffer_from_numpy(lights.view(np.uint8))
    outs = [ctx.create_buffer(slots * 4 * 64) for _ in range(3)]
    bd = ctx.create_buffer(len(lights) * 16)
    for b in outs + [bd]:
        b.zero()
    if with_aux:
        csm_cascades(ctx, bc, len(cams), bl, len(lights), outs[0], outs[1], outs[2], bd)
    else:
        csm_cascades(ctx, bc, len(cams), bl, len(lights), outs[0])
    ctx.synchronize()
    res = [b.download(np.float32).reshape(slots, 4, 16) for b in outs] + [bd.download(np.float32).reshape(len(lights), 4)]
    for b in outs + [bd, bc, bl]:
        b.destroy()
    return res


@pytest.mark.parametrize("name", ["a", "b", "c"])
def test_csm_matches_glm_golden(ctx, name):
    g = np.load(GOLD)
    cams, lights = _device_inputs(g[f"{name}_cams"], g[f"{name}_lights"])
    assert np.array_equal(cams["tanHalfFovy"], g[f"{name}_tanhalf"]), "host tanf differs from the one glm called"
    vp, v, p, d = _run(ctx, cams, lights)
    assert np.array_equal(v.view(np.uint32), g[f"{name}_view"].view(np.uint32))
    assert np.array_equal(p.view(np.uint32), g[f"{name}_proj"].view(np.uint32))
    assert np.array_equal(vp.view(np.uint32), g[f"{name}_viewproj"].view(np.uint32))
    assert np.array_equal(d.view(np.uint32), g[f"{name}_dist"].view(np.uint32))


def test_csm_matches_oracle_on_fresh_cases_and_optional_outputs(ctx):
    for seed in (31, 32):
        cams_ref, lights_ref = mg.make_inputs(seed, 6, 5)
        cams, lights = _device_inputs(cams_ref, lights_ref)
        want = ocsm.csm_cascades(cams_from_ref(cams_ref), lights_from_ref(lights_ref))
        got = _run(ctx, cams, lights)
        for w, g_ in zip(want, got):
            assert np.array_equal(w.view(np.uint32), g_.view(np.uint32))
        only_vp = _run(ctx, cams, lights, with_aux=False)
        assert np.array_equal(only_vp[0].view(np.uint32), want[0].view(np.uint32))
        assert not only_vp[1].any() and not only_vp[3].any()


def test_csm_rejects_small_outputs(ctx):
    from ravengine_b200 import RveError
    cams_ref, lights_ref = mg.make_inputs(5, 2, 2)
    cams, lights = _device_inputs(cams_ref, lights_ref)
    bc, bl = ctx.buffer_from_numpy(cams.view(np.uint8)), ctx.buffer_from_numpy(lights.view(np.uint8))
    small = ctx.create_buffer(64)
    with pytest.raises(RveError):
        csm_cascades(ctx, bc, 2, bl, 2, small)
