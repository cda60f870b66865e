#!/usr/bin/env python
"""Launches every hot kernel of the path a few times, for one `ncu --set full` capture (profiles/rNN_*_ncu_*):
C2 step (pyramid_persistent_kernel, upload_params_kernel, cull_kernel), the same cull over churned
entity lists (gather path), skin_kernel (C3 shape, 500 characters), cluster_assign_kernel (C4, one view),
cull_views_kernel (8 views, 2 M instances), full_rows_kernel (48-B affine upload expand).

  python tools/ncu_targets.py            # plain run first, then the same under ncu (B200_PROFILING.md)
"""
from __future__ import annotations

import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))

import torch  # noqa: E402

import bench  # noqa: E402
from ravengine_b200 import GridAssignUBO, GridBuildUBO, SkinningUBO, scenes  # noqa: E402


def main():
    torch.cuda.set_device(0)
    sc = bench.DeviceScene(bench.C2["instances"], 0, 0)
    ctx = sc.ctx
    for _ in range(3):
        sc.step()
    ctx.synchronize()
    # churned entity lists: the gather path of the same kernel
    from ravengine_b200.vram_vector import SparseSet

    rng = np.random.default_rng(1)
    n, m = sc.n, bench.C2["materials"]
    host_ids = np.arange(n, dtype=np.uint32)
    for i in range(m):
        lo, hi = n * i // m, n * (i + 1) // m
        ss = SparseSet(ctx)
        ss.emplace_many(host_ids[lo:hi])
        gone = rng.choice(host_ids[lo:hi], size=(hi - lo) // 10, replace=False)
        ss.erase_many(gone)
        ss.emplace_many(rng.permutation(gone))
        host_ids[lo:hi] = ss.dense()
        ss.destroy()
    with torch.cuda.stream(sc.stream):
        sc.t_ids.copy_(torch.from_numpy(host_ids.view(np.int32)).to(sc.t_ids.device))
    for _ in range(2):
        sc.step()
    ctx.synchronize()
    # view batch: 8 views over the first 2 M instances
    cams = bench.c5_cameras(8)
    views = [(c.view_proj_cm, c.pos, 0xFFFFFFFF, False) for c in cams]
    import os

    os.environ["RVS_VIEW_BATCH_MIN"] = "1"
    for _ in range(2):
        sc.world.cull_static_views(views, [sc.pyr] * 8)
    ctx.synchronize()
    # skinning, C3 shape at 500 characters
    nv, no, nb = 20_000, 500, 64
    dd = scenes.skinning_inputs(rng, nv, 1, nb, coherent_joints=True)
    ins = [ctx.buffer_from_numpy(dd[k]) for k in ("pos", "nrm", "tan", "bit", "uv")]
    wts = ctx.buffer_from_numpy(dd["weights"])
    dev = torch.device("cuda", 0)
    with torch.cuda.stream(sc.stream):
        t_pose = torch.eye(4, device=dev).reshape(1, 16).repeat(no * nb, 1).contiguous()
        t_out = [torch.empty((no * nv, k), dtype=torch.float32, device=dev) for k in (3, 3, 3, 3, 2)]
    pose, outs = ctx.wrap_tensor(t_pose), [ctx.wrap_tensor(t) for t in t_out]
    for _ in range(2):
        ctx.skin(SkinningUBO(no, nv, nb, 0, 0, 0), outs, ins, pose, wts)
    ctx.synchronize()
    # cluster assignment, C4
    grid, nl = (16, 9, 24), 65_536
    cam = scenes.default_camera()
    z = -rng.uniform(0.2, 100.0, size=nl)
    pos = np.stack([rng.uniform(-1, 1, size=nl) * (-z) * 1.03, rng.uniform(-1, 1, size=nl) * (-z) * 0.6, z], 1).astype(np.float32)
    inten = np.exp(rng.uniform(np.log(0.05), np.log(5.0), size=nl)).astype(np.float32)
    pts = ctx.buffer_from_numpy(scenes.point_lights_bytes(pos, inten))
    ncl = grid[0] * grid[1] * grid[2]
    cl = ctx.create_buffer(ncl * 160)
    u = GridBuildUBO()
    u.invProj[:] = [float(x) for x in cam.inv_proj_cm]
    u.gridSize[:] = grid
    u.zNear, u.zFar = cam.z_near, cam.z_far
    u.screenDim[:] = (3840, 2160)
    a = GridAssignUBO()
    a.viewMat[:] = [float(x) for x in scenes.colmajor(np.eye(4))]
    a.pointLightCount, a.spotLightCount = nl, 0
    ctx.cluster_build(u, cl)
    for _ in range(2):
        ctx.cluster_assign(a, cl, pts, None, ncl)
    ctx.synchronize()
    # affine upload expand
    from ravengine_b200.vram_vector import SYNC_FULL, BufferedVRAMVector

    v = BufferedVRAMVector(ctx, 48, 2_000_000, affine_rows=True)
    v.host(np.float32)[:] = 1.0
    for _ in range(2):
        v.encode_sync(SYNC_FULL)
    ctx.synchronize()
    print("ncu targets done, launches:", ctx.launch_count)


if __name__ == "__main__":
    main()
