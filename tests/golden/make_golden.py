#!/usr/bin/env python
"""Generates the golden fixtures under tests/golden/ from the CPU oracle (run from the repo root:
`python tests/golden/make_golden.py`).

The reference cannot run here (its implementation of this path is GLSL behind a graphics API), so
these vectors are outputs of the oracle restatement at the commit that wrote them; they freeze the arithmetic
contract (oracle/oracle.cpp header) so that neither the oracle nor the CUDA kernels can drift silently.
Inputs are stored with the outputs: nothing depends on a random generator reproducing itself.
"""
from __future__ import annotations

import math
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))

import oracle  # noqa: E402
from oracle import host_tables  # noqa: E402
from ravengine_b200 import GridAssignUBO, GridBuildUBO, SkinningUBO, scenes  # noqa: E402
from tests.helpers import make_ubo, oracle_cull  # noqa: E402

OUT = Path(__file__).resolve().parent


def cull_materials(ids_per_cmd, lods_per_cmd, radii, mats_split):
    """Rebuilds the MaterialDraws list from plain arrays (also used by the tests to reload a fixture)."""
    mats, k = [], 0
    for n_cmds in mats_split:
        m = scenes.MaterialDraws([])
        for _ in range(n_cmds):
            lods = tuple(float(x) for x in lods_per_cmd[k] if not math.isnan(x))
            m.commands.append(scenes.DrawCommand(scenes.simple_mesh(float(radii[k]), lods, base=1000 * k), ids_per_cmd[k]))
            k += 1
        mats.append(m)
    return mats


def make_cull():
    rng = np.random.default_rng(20260101)
    n = 4000
    cam = scenes.default_camera(z_far=1000.0)
    depth = scenes.synth_depth(256, 144, cam, n_occluders=25, seed=11)
    model = scenes.random_transforms(rng, n, 300.0, 40.0)
    layers = np.full(n, 0xFFFFFFFF, np.uint32)
    layers[rng.random(n) < 0.05] = 0
    flags = np.full(n, 0xFFFF, np.uint16)
    flags[rng.random(n) < 0.05] &= 0xFFFC
    flags[rng.random(n) < 0.05] &= 0xFFFD
    ids = rng.permutation(n).astype(np.uint32)
    ids[:2000] = np.sort(ids[:2000])  # half dense-ish runs, half shuffled
    bounds = [0, 900, 1500, 2600, 2601, 4000]
    lods = [[0.0, 40.0, 120.0], [math.inf, math.nan, math.nan], [0.0, 25.0, math.nan], [0.0, 40.0, 120.0], [0.0, 60.0, 30.0]]
    radii = [1.7320508, 1.74, 0.8, 3.0, 1.2]
    ids_per_cmd = [ids[bounds[i]:bounds[i + 1]] for i in range(5)]
    split = [2, 1, 2]
    mats = cull_materials(ids_per_cmd, lods, radii, split)
    t = host_tables.build_static_tables(mats)
    ubo = make_ubo(cam.view_proj_cm, cam.pos, t.num_cubos)
    pyr, dim, levels = oracle.pyramid_build(depth, 144)
    out = oracle_cull(t, ubo, model, layers, flags, pyr, dim, levels)
    indirect, slices = [], []
    ci = 0
    for m in mats:
        cb = t.cubos[ci]
        h_c, h_i = int(cb["idOutputBufferBindlessHandle"]), int(cb["indirectOutputBufferBindlessHandle"])
        indirect.append(out[h_i].copy())
        slices += host_tables.sorted_slices(out[h_i], out[h_c], t.cubos, range(ci, ci + len(m.commands)))
        ci += len(m.commands)
    np.savez_compressed(OUT / "cull_hiz_lod.npz", model=model, layers=layers, flags=flags, depth=depth, ids=ids, bounds=np.array(bounds),
                        lods=np.array(lods, np.float64), radii=np.array(radii, np.float64), split=np.array(split),
                        view_proj=cam.view_proj_cm, cam_pos=cam.pos, pyramid=pyr, pyr_dim=dim, pyr_levels=levels,
                        indirect=np.concatenate(indirect), slice_sizes=np.array([len(s) for s in slices]),
                        slices=np.concatenate(slices) if slices else np.zeros(0, np.uint32))
    print("cull: visible per slot", np.concatenate(indirect)[:, 1].tolist())


def make_skin():
    rng = np.random.default_rng(20260102)
    nv, no, nb = 64, 3, 8
    d = scenes.skinning_inputs(rng, nv, no, nb)
    ins = [d["pos"], d["nrm"], d["tan"], d["bit"], d["uv"]]
    outs = [np.zeros((no * nv, k), np.float32) for k in (3, 3, 3, 3, 2)]
    oracle.skin(SkinningUBO(no, nv, nb, 0, 0, 0), outs, ins, d["pose"], d["weights"])
    np.savez_compressed(OUT / "skinning.npz", pos=ins[0], nrm=ins[1], tan=ins[2], bit=ins[3], uv=ins[4], pose=d["pose"], weights=d["weights"],
                        out_pos=outs[0], out_nrm=outs[1], out_tan=outs[2], out_bit=outs[3], out_uv=outs[4])


def make_clusters():
    rng = np.random.default_rng(20260103)
    cam = scenes.default_camera()
    grid = (4, 3, 5)
    u = GridBuildUBO()
    u.invProj[:] = [float(x) for x in cam.inv_proj_cm]
    u.gridSize[:] = grid
    u.zNear, u.zFar = cam.z_near, cam.z_far
    u.screenDim[:] = (1280, 720)
    n = grid[0] * grid[1] * grid[2]
    cl = np.zeros((n, 40), np.uint32)
    oracle.cluster_build(u, cl)

    def lights(k):
        z = -rng.uniform(0.2, 100.0, size=k)
        x = rng.uniform(-1, 1, size=k) * (-z)
        y = rng.uniform(-1, 1, size=k) * (-z) * 0.6
        return np.stack([x, y, z], 1).astype(np.float32), np.exp(rng.uniform(np.log(0.05), np.log(5.0), size=k)).astype(np.float32)

    pp, pi = lights(40)
    sp, si = lights(10)
    view = scenes.colmajor(scenes.look_at((1.0, 2.0, 3.0), (0.5, 1.5, -4.0)))
    a = GridAssignUBO()
    a.viewMat[:] = [float(x) for x in view]
    a.pointLightCount, a.spotLightCount = 40, 10
    points, spots = scenes.point_lights_bytes(pp, pi), scenes.spot_lights_bytes(sp, si)
    oracle.cluster_assign(a, cl, points, spots, n)
    np.savez_compressed(OUT / "clusters.npz", inv_proj=cam.inv_proj_cm, grid=np.array(grid), screen=np.array([1280, 720]),
                        z_near=cam.z_near, z_far=cam.z_far, view=view, point_pos=pp, point_int=pi, spot_pos=sp, spot_int=si, clusters=cl)
    print("clusters: point counts", cl[:, 6].tolist())


if __name__ == "__main__":
    make_cull()
    make_skin()
    make_clusters()
