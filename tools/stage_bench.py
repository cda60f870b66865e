#!/usr/bin/env python
"""Per-stage timings of the render-prep path against each kernel's roofline, one JSON line per stage:
C1 (100 k cull, frustum+LOD), C2 variants (shuffled ids), pyramid (3840x2160), C3 (skinning 10 k x 20 k verts, or a
scaled-down object count with --c3-objects), skinned draw-call prep, C4 (cluster build + assign, 65 536 lights, 5 views).
`python bench.py` stays the headline (C2); this tool fills profiles/ with the other rows.

  python tools/stage_bench.py [--reps 20] [--c3-objects 10000] > profiles/rNN_stages.jsonl
"""
from __future__ import annotations

import argparse
import json
import math
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))

import torch  # noqa: E402

from ravengine_b200 import (Context, GridAssignUBO, GridBuildUBO, SceneGenParams, SkinningPrepareUBO, SkinningUBO, scenes)  # noqa: E402
from ravengine_b200.world import World  # noqa: E402

PEAK = json.loads((ROOT / "MEASURED_PEAKS.json").read_text())["hbm_gbs"] if (ROOT / "MEASURED_PEAKS.json").exists() else 6650.0


def timed(ctx, fn, reps, flush=None):
    """median / best device time of fn() over reps, CUDA events on the context's stream"""
    st = torch.cuda.ExternalStream(ctx.stream)
    for _ in range(3):
        fn()
    ctx.synchronize()
    ts = []
    for _ in range(reps):
        if flush is not None:
            with torch.cuda.stream(st):
                flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(st)
        fn()
        b.record(st)
        b.synchronize()
        ts.append(a.elapsed_time(b))
    return float(np.median(ts)), float(min(ts))


def line(stage, workload, ms_med, ms_best, algo_bytes, bound, extra=None):
    d = {"stage": stage, "workload": workload, "ms_median": ms_med, "ms_best": ms_best, "algorithmic_bytes": int(algo_bytes),
         "achieved_gbs": algo_bytes / (ms_med * 1e-3) / 1e9, "peak_gbs": PEAK, "frac_of_measured_peak": algo_bytes / (ms_med * 1e-3) / 1e9 / PEAK,
         "bound": bound}
    d.update(extra or {})
    print(json.dumps(d), flush=True)


def cull_case(ctx, n, lods, occlusion, shuffle, depth_wh, reps, name, flush):
    dev = torch.device("cuda", ctx.device)
    st = torch.cuda.ExternalStream(ctx.stream)
    with torch.cuda.stream(st):
        t_model = torch.empty((n, 16), dtype=torch.float32, device=dev)
        t_layers, t_flags = torch.empty(n, dtype=torch.int32, device=dev), torch.empty(n, dtype=torch.int16, device=dev)
        t_ids = torch.empty(n, dtype=torch.int32, device=dev)
    model, layers, flags, ids = (ctx.wrap_tensor(t) for t in (t_model, t_layers, t_flags, t_ids))
    far = 1000.0 if occlusion else 100.0
    ext = (500.0, 50.0) if occlusion else (200.0, 200.0)
    ctx.gen_instances(SceneGenParams(20240, ext[0], ext[1], 0.5, 2.0, 0.05 if occlusion else 0.0, 0.05 if occlusion else 0.0, 0), 0, n,
                      model, layers, flags, ids)
    ctx.synchronize()
    if not occlusion:
        t_flags.fill_(-3)  # 0xFFFD: occlusion bit cleared (C1 is frustum + LOD only)
    cam = scenes.default_camera(z_near=0.1, z_far=far)
    w = World(ctx)
    w.set_entity_buffers(model, layers, flags)
    n_mats = 4 if occlusion else 2
    bounds = [n * i // n_mats for i in range(n_mats + 1)]
    keep = []
    if shuffle:
        perm = torch.randperm(n, device=dev, generator=torch.Generator(device=dev).manual_seed(1)).to(torch.int32)
        t_ids.copy_(perm)
        torch.cuda.synchronize()
    for i in range(n_mats):
        sl = ctx.wrap_tensor(t_ids[bounds[i]:bounds[i + 1]])
        keep.append(sl)
        k = len(lods)
        mesh = w.mesh_collection(1.7320508 if i % 2 == 0 else 1.74, lods, [36, 24, 12][:k], [0, 36, 60][:k], [0, 24, 40][:k])
        m = w.material()
        w.add_command(m, mesh, sl, bounds[i + 1] - bounds[i])
        keep.append(m)
    pyr = ctx.create_pyramid(min(depth_wh) if depth_wh else 64)
    if depth_wh:
        depth = scenes.synth_depth(depth_wh[0], depth_wh[1], cam)
        d = ctx.buffer_from_numpy(depth)
        ctx.build_pyramid(pyr, d, depth_wh[0], depth_wh[1])
    ctx.set_profiling(True)
    for _ in range(reps + 3):
        if flush is not None:
            with torch.cuda.stream(st):
                flush.zero_()
        w.cull_static(cam.view_proj_cm, cam.pos, pyr)
    ms, cnt = ctx.profile_read(0)
    ctx.set_profiling(False)
    vis = sum(int(w.material_outputs(m)[1].download(np.uint32).reshape(-1, 5)[:len(lods), 1].sum()) for m in keep if isinstance(m, int))
    algo = n * 74 + vis * 4 + 36 * n_mats + 20 * n_mats * len(lods) + (4 * (4 * pyr.dim ** 2 - 1) // 3 if depth_wh else 0)
    k_ms = ms / cnt
    line("cull", name, k_ms, k_ms, algo, "hbm", {"instances": n, "visible": vis, "ginst_per_s": n / (k_ms * 1e-3) / 1e9,
                                                 "l2": "flushed between launches" if flush is not None else "inputs exceed L2"})
    w.destroy()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--reps", type=int, default=20)
    ap.add_argument("--c3-objects", type=int, default=10_000)
    ap.add_argument("--only", default="")
    args = ap.parse_args()
    only = set(args.only.split(",")) if args.only else None
    want = lambda s: only is None or s in only
    torch.cuda.set_device(0)
    stream = torch.cuda.Stream()
    ctx = Context(0, stream.cuda_stream)
    dev = torch.device("cuda", 0)
    with torch.cuda.stream(stream):
        flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)  # > 126 MB L2

    if want("c1"):
        cull_case(ctx, 100_000, (math.inf,), False, False, None, args.reps, "C1: 100k instances, 2 materials, 1 LOD, frustum+LOD only", flush)
    if want("c2"):
        cull_case(ctx, 10_000_000, (0.0, 40.0, 120.0), True, False, (3840, 2160), args.reps, "C2: 10M instances, dense ids (TMA path)", None)
        cull_case(ctx, 10_000_000, (0.0, 40.0, 120.0), True, True, (3840, 2160), args.reps, "C2 variant: 10M instances, shuffled ids (cp.async gather path)", None)

    if want("pyramid"):
        cam = scenes.default_camera(z_far=1000.0)
        depth = scenes.synth_depth(3840, 2160, cam)
        d = ctx.buffer_from_numpy(depth)
        pyr = ctx.create_pyramid(2160)
        med, best = timed(ctx, lambda: ctx.build_pyramid(pyr, d, 3840, 2160), args.reps, flush)
        algo = min(4 * 3840 * 2160, 16 * pyr.dim ** 2) + 4 * (4 * pyr.dim ** 2 - 1) // 3
        line("pyramid", "3840x2160 depth -> 1024^2, 11 levels, one launch", med, best, algo, "hbm", {"dim": pyr.dim, "levels": pyr.levels})

    if want("c3"):
        nv, no, nb = 20_000, args.c3_objects, 64
        rng = np.random.default_rng(3)
        dd = scenes.skinning_inputs(rng, nv, 1, nb, coherent_joints=True)
        ins = [ctx.buffer_from_numpy(dd[k]) for k in ("pos", "nrm", "tan", "bit", "uv")]
        wts = ctx.buffer_from_numpy(dd["weights"])
        with torch.cuda.stream(stream):
            t_pose = torch.eye(4, device=dev).reshape(1, 16).repeat(no * nb, 1).contiguous()
            t_pose += 0.01 * torch.randn(t_pose.shape, device=dev, generator=torch.Generator(device=dev).manual_seed(2))
            t_out = [torch.empty((no * nv, k), dtype=torch.float32, device=dev) for k in (3, 3, 3, 3, 2)]
        pose = ctx.wrap_tensor(t_pose)
        outs = [ctx.wrap_tensor(t) for t in t_out]
        ubo = SkinningUBO(no, nv, nb, 0, 0, 0)
        med, best = timed(ctx, lambda: ctx.skin(ubo, outs, ins, pose, wts), max(args.reps // 2, 5))
        algo = no * nv * 56 + no * nb * 64 + nv * 88
        line("skinning", f"C3: {no} objects x {nb} joints x {nv} verts (output {no * nv * 56 / 1e9:.2f} GB)", med, best, algo, "hbm",
             {"gverts_per_s": no * nv / (med * 1e-3) / 1e9})
        del t_out, outs
        pu = SkinningPrepareUBO(0, 0, 60_000, nv, no, 0, 0)
        ind = ctx.create_buffer(no * 20)
        med, best = timed(ctx, lambda: ctx.skinned_drawcall_prepare(pu, ind), args.reps)
        line("skinned_drawcall_prepare", f"{no} objects", med, best, no * 20, "launch latency")

    if want("c4"):
        for grid in ((16, 9, 24), (12, 12, 24)):
            cam = scenes.default_camera()
            rng = np.random.default_rng(4)
            nl = 65_536
            z = -rng.uniform(0.2, 100.0, size=nl)
            pos = np.stack([rng.uniform(-1, 1, size=nl) * (-z) * 1.03, rng.uniform(-1, 1, size=nl) * (-z) * 0.6, z], 1).astype(np.float32)
            inten = np.exp(rng.uniform(np.log(0.05), np.log(5.0), size=nl)).astype(np.float32)
            pts = ctx.buffer_from_numpy(scenes.point_lights_bytes(pos, inten))
            n = grid[0] * grid[1] * grid[2]
            cl = ctx.create_buffer(n * 160)
            u = GridBuildUBO()
            u.invProj[:] = [float(x) for x in cam.inv_proj_cm]
            u.gridSize[:] = grid
            u.zNear, u.zFar = cam.z_near, cam.z_far
            u.screenDim[:] = (3840, 2160)
            a = GridAssignUBO()
            a.pointLightCount, a.spotLightCount = nl, 0
            views = [np.eye(4)] + [scenes.look_at((10.0 * i, 5.0, 3.0), (0.0, 0.0, -30.0)) for i in range(1, 5)]
            med, best = timed(ctx, lambda: ctx.cluster_build(u, cl), args.reps)
            line("cluster_build", f"C4 grid {grid}", med, best, n * 24, "launch latency")

            def five_views():
                for v in views:
                    a.viewMat[:] = [float(x) for x in scenes.colmajor(v)]
                    ctx.cluster_assign(a, cl, pts, None, n)
            med, best = timed(ctx, five_views, args.reps)
            tests = 5 * n * nl
            line("cluster_assign", f"C4 grid {grid}, {nl} point lights, 5 views (prepare + assign per view)", med, best,
                 5 * (nl * 64 + n * 320), "fp32/lsu issue (not HBM)", {"sphere_aabb_tests": tests, "gtests_per_s": tests / (med * 1e-3) / 1e9, "ms_per_view": med / 5})
    ctx.close()


if __name__ == "__main__":
    main()
