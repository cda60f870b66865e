#!/usr/bin/env python
"""Per-stage timings of the render-prep path against each kernel's roofline, one JSON line per stage:
C1 (100 k cull, frustum+LOD), C2 variants (shuffled ids), pyramid (3840x2160), C3 (skinning 10 k x 20 k verts, or a
scaled-down object count with --c3-objects), skinned draw-call prep, C4 (cluster build + assign, 65 536 lights, 5 views).
Every stage line carries the CPU number of the same stage (the oracle on the host cores, taskflow executor, core count
stated; bounded sample where the full stage would take too long) - `cpu: false` skips those legs.

`python bench.py` imports run_stages() for the `stages` block of its N=1 line; run alone this tool fills profiles/:

  python tools/stage_bench.py [--reps 20] [--c3-objects 10000] [--no-cpu] > profiles/rNN_stages.jsonl
"""
from __future__ import annotations

import argparse
import json
import math
import sys
import time
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
    return d


def _cpu_cull(n, lods, occlusion, shuffle, depth, cam, n_mats, reps=3, ids_override=None):
    """the same cull on the host cores: oracle (CPU restatement, TEST INFRASTRUCTURE) with the engine's worker count"""
    import oracle
    from oracle import host_tables
    from tests.helpers import make_ubo

    tf = oracle.LIB_TF.exists()
    workers = oracle.default_workers()
    model = np.zeros((n, 16), np.float32)
    layers, flags, ids = np.zeros(n, np.uint32), np.zeros(n, np.uint16), np.zeros(n, np.uint32)
    ext = (500.0, 50.0) if occlusion else (200.0, 200.0)
    oracle.gen_instances(SceneGenParams(20240, ext[0], ext[1], 0.5, 2.0, 0.05 if occlusion else 0.0, 0.05 if occlusion else 0.0, 0), 0, n,
                         model, layers, flags, ids)
    if not occlusion:
        flags[:] = 0xFFFD
    if ids_override is not None:
        ids = ids_override
    elif shuffle:
        ids = np.random.default_rng(1).permutation(n).astype(np.uint32)
    bounds = [n * i // n_mats for i in range(n_mats + 1)]
    mats = []
    for i in range(n_mats):
        mesh = scenes.simple_mesh(1.7320508 if i % 2 == 0 else 1.74, lods)
        mats.append(scenes.MaterialDraws([scenes.DrawCommand(mesh, ids[bounds[i]:bounds[i + 1]])]))
    t = host_tables.build_static_tables(mats)
    ubo = make_ubo(cam.view_proj_cm, cam.pos, t.num_cubos)
    if depth is not None:
        pyr, dim, levels = oracle.pyramid_build(depth, min(depth.shape))
    else:
        pyr, dim, levels = np.zeros(int(oracle.lib().orc_pyramid_texel_count(64)), np.float32), 64, 7
    groups = (t.total_entities + 63) // 64
    ts = []
    for _ in range(reps):
        for b in t.bindless:
            if b is not None and b.ndim == 2:
                b[:, 1] = 0
        t0 = time.perf_counter()
        oracle.cull(ubo, t.cubos, model, layers, flags, pyr, dim, levels, t.bindless, groups, workers, tf)
        ts.append(time.perf_counter() - t0)
    ms = float(np.median(ts)) * 1e3
    return {"ms": ms, "ginst_per_s": n / ms / 1e6, "cores": workers, "kind": "port", "executor": "taskflow" if tf else "std::thread",
            "sample": f"{n} instances (the full stage), median of {reps}"}


def cull_case(ctx, n, lods, occlusion, shuffle, depth_wh, reps, name, flush, cpu=True, world_flags=0):
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
        with torch.cuda.stream(st):
            t_flags.fill_(-3)  # 0xFFFD: occlusion bit cleared (C1 is frustum + LOD only)
    cam = scenes.default_camera(z_near=0.1, z_far=far)
    w = World(ctx, world_flags)
    w.set_entity_buffers(model, layers, flags)
    n_mats = 4 if occlusion else 2
    bounds = [n * i // n_mats for i in range(n_mats + 1)]
    keep = []
    host_ids = None
    if shuffle == "churn":
        # every material's list goes through the engine's own mutation: 10 % of its entities erased at random (swap-remove)
        # and inserted again - BufferedVRAMSparseSet semantics through the shim (csrc/sync.cu)
        from ravengine_b200.vram_vector import SparseSet

        rng = np.random.default_rng(1)
        host_ids = np.arange(n, dtype=np.uint32)
        for i in range(n_mats):
            ss = SparseSet(ctx)
            ss.emplace_many(host_ids[bounds[i]:bounds[i + 1]])
            gone = rng.choice(host_ids[bounds[i]:bounds[i + 1]], size=(bounds[i + 1] - bounds[i]) // 10, replace=False)
            ss.erase_many(gone)
            ss.emplace_many(rng.permutation(gone))
            host_ids[bounds[i]:bounds[i + 1]] = ss.dense()
            ss.destroy()
    elif shuffle:
        host_ids = np.random.default_rng(1).permutation(n).astype(np.uint32)
    if host_ids is not None:
        perm = torch.from_numpy(host_ids.view(np.int32)).to(dev)
        with torch.cuda.stream(st):
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
    depth = None
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
    step_ms = None
    if n <= 1_000_000:  # small scenes are launch-bound: the whole step (upload + cull), enqueued call by call and as one captured graph
        def one():
            w.cull_static(cam.view_proj_cm, cam.pos, pyr)
        direct, _ = timed(ctx, one, reps, flush)
        graph = ctx.capture(one)
        replay, _ = timed(ctx, graph.launch, reps, flush)
        graph.destroy()
        step_ms = {"step_ms_call_by_call": direct, "step_ms_captured_graph": replay}
    counts = [w.material_outputs(m)[1].download(np.uint32).reshape(-1, 5)[:len(lods), 1] for m in keep if isinstance(m, int)]
    vis = int(sum(int(c.sum()) for c in counts))
    algo = n * 74 + vis * 4 + 36 * n_mats + 20 * n_mats * len(lods) + (4 * (4 * pyr.dim ** 2 - 1) // 3 if depth_wh else 0)
    k_ms = ms / cnt
    extra = {"instances": n, "visible": vis, "ginst_per_s": n / (k_ms * 1e-3) / 1e9,
             "l2": "flushed between launches" if flush is not None else "inputs exceed L2"}
    if step_ms:
        extra.update(step_ms)
    if host_ids is not None:
        runs = np.flatnonzero(np.diff(host_ids.astype(np.int64)) != 1).size + 1
        extra["id_runs"] = int(runs)  # maximal runs of consecutive ids in the lists (1 per material when untouched)
    if cpu:
        extra["cpu"] = _cpu_cull(n, lods, occlusion, shuffle, depth, cam, n_mats, ids_override=host_ids)
    w.destroy()
    return line("cull", name, k_ms, k_ms, algo, "hbm", extra)


def live_churn_case(ctx, n, per_frame, frames, sort_lists):
    """C2 with every material's list held in a BufferedVRAMSparseSet stand-in that CHANGES every frame: after an initial 10 % churn
    each frame erases `per_frame` random members of every list and inserts as many other entities (host calls, not timed), then
    enqueues the cull (set sync = packed scatter of the touched slots, sorted copies kept by deltas when `sort_lists`, cull).
    Returns the median GPU time of a frame's enqueued work."""
    from ravengine_b200.vram_vector import SparseSet
    from ravengine_b200.world import SORT_ENTITY_LISTS

    dev = torch.device("cuda", ctx.device)
    st = torch.cuda.ExternalStream(ctx.stream)
    with torch.cuda.stream(st):
        t_model = torch.empty((n, 16), dtype=torch.float32, device=dev)
        t_layers, t_flags = torch.empty(n, dtype=torch.int32, device=dev), torch.empty(n, dtype=torch.int16, device=dev)
        t_ids = torch.empty(n, dtype=torch.int32, device=dev)
    model, layers, flags, ids = (ctx.wrap_tensor(t) for t in (t_model, t_layers, t_flags, t_ids))
    ctx.gen_instances(SceneGenParams(20240, 500.0, 50.0, 0.5, 2.0, 0.05, 0.05, 0), 0, n, model, layers, flags, ids)
    ctx.synchronize()
    cam = scenes.default_camera(z_near=0.1, z_far=1000.0)
    w = World(ctx, SORT_ENTITY_LISTS if sort_lists else 0)
    w.set_entity_buffers(model, layers, flags)
    rng = np.random.default_rng(1)
    n_mats = 4
    per = n // n_mats
    sets, spare = [], []
    for i in range(n_mats):
        base = np.arange(i * per, (i + 1) * per, dtype=np.uint32)
        ss = SparseSet(ctx)
        live = base[: per - per // 50]  # 2 % of the range stays free for insertions
        ss.emplace_many(live)
        gone = rng.choice(live, size=per // 10, replace=False)
        ss.erase_many(gone)
        ss.emplace_many(rng.permutation(gone))
        sets.append(ss)
        spare.append(list(base[per - per // 50:]))
        mesh = w.mesh_collection(1.7320508 if i % 2 == 0 else 1.74, (0.0, 40.0, 120.0), [36, 24, 12], [0, 36, 60], [0, 24, 40])
        w.add_command_set(w.material(), mesh, ss)
    pyr = ctx.create_pyramid(2160)
    d = ctx.buffer_from_numpy(scenes.synth_depth(3840, 2160, cam))
    ctx.build_pyramid(pyr, d, 3840, 2160)
    for _ in range(2):
        w.cull_static(cam.view_proj_cm, cam.pos, pyr)
    ctx.synchronize()
    ts, host = [], []
    for f in range(frames):
        for i, ss in enumerate(sets):
            members = ss.dense()
            out = members[rng.integers(0, len(members), size=per_frame)]
            out = np.unique(out)
            ss.erase_many(out)
            take = [spare[i].pop() for _ in range(len(out))]
            ss.emplace_many(np.asarray(take, dtype=np.uint32))
            spare[i] = list(out) + spare[i]
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        with torch.cuda.stream(st):
            torch.cuda._sleep(12_000_000)  # ~6 ms: the stream is busy while the host scans the sets' flags and enqueues the frame,
        a.record(st)                       # so the events bracket the GPU's execution of that work, not the host's preparation
        t0 = time.perf_counter()
        w.cull_static(cam.view_proj_cm, cam.pos, pyr)
        host.append((time.perf_counter() - t0) * 1e3)
        b.record(st)
        b.synchronize()
        ts.append(a.elapsed_time(b))
    stats = w.sorted_list_stats()
    vis = int(sum(int(w.material_outputs(m)[1].download(np.uint32).reshape(-1, 5)[:3, 1].sum()) for m in range(n_mats)))
    w.destroy()
    for ss in sets:
        ss.destroy()
    return float(np.median(ts)), float(min(ts)), stats, vis, float(np.median(host))


def run_stages(ctx, stream, reps=20, c3_objects=10_000, only=None, cpu=True):
    """Generator of stage dicts (see module docstring)."""
    want = lambda s: only is None or s in only  # noqa: E731
    dev = torch.device("cuda", ctx.device)
    with torch.cuda.stream(stream):
        flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)  # > 126 MB L2

    if want("c1"):
        yield cull_case(ctx, 100_000, (math.inf,), False, False, None, reps, "C1: 100k instances, 2 materials, 1 LOD, frustum+LOD only", flush, cpu)
    if want("c2"):
        yield cull_case(ctx, 10_000_000, (0.0, 40.0, 120.0), True, False, (3840, 2160), reps, "C2: 10M instances, dense ids (TMA path)", None, False)
    if want("c2s"):
        yield cull_case(ctx, 10_000_000, (0.0, 40.0, 120.0), True, True, (3840, 2160), reps,
                        "C2 variant: 10M instances, shuffled ids (cp.async gather path)", None, cpu)

    if want("c2churn"):
        yield cull_case(ctx, 10_000_000, (0.0, 40.0, 120.0), True, "churn", (3840, 2160), reps,
                        "C2 variant: 10M instances after 10 % erase / re-insert churn of every entity list (swap-remove)", None, cpu)

    if want("c2sorted"):
        from ravengine_b200.world import SORT_ENTITY_LISTS

        yield cull_case(ctx, 10_000_000, (0.0, 40.0, 120.0), True, "churn", (3840, 2160), reps,
                        "C2 variant: the churned lists, culled through their ascending device copies (RVS_SORT_ENTITY_LISTS; "
                        "the copy is rebuilt only when a list changes - not in these launches)", None, False, SORT_ENTITY_LISTS)
        yield cull_case(ctx, 10_000_000, (0.0, 40.0, 120.0), True, True, (3840, 2160), reps,
                        "C2 variant: fully shuffled lists through their ascending device copies", None, False, SORT_ENTITY_LISTS)

    if want("c2sorted"):
        # what rebuilding such copies costs (rve_sort_entity_ids / _batch; the shim does it when lists changed): the four lists of
        # C2 - contiguous 2.5 M-entity ranges of a 10 M-entity id space after 10 % swap-remove churn each
        from ravengine_b200.vram_vector import SparseSet

        n_ids, id_space = 2_500_000, 10_000_000
        rng = np.random.default_rng(4)
        lists = []
        for i in range(4):
            ss = SparseSet(ctx)
            base = np.arange(i * n_ids, (i + 1) * n_ids, dtype=np.uint32)
            ss.emplace_many(base)
            gone = rng.choice(base, size=n_ids // 10, replace=False)
            ss.erase_many(gone)
            ss.emplace_many(rng.permutation(gone))
            lists.append(ss.dense().copy())
            ss.destroy()
        srcs = [ctx.buffer_from_numpy(l) for l in lists]
        dsts = [ctx.create_buffer(4 * n_ids) for _ in lists]
        words = (id_space + 31) // 32
        med, best = timed(ctx, lambda: ctx.sort_entity_ids(srcs[1], n_ids, id_space, dsts[1]), reps, flush)
        assert np.array_equal(dsts[1].download(np.uint32)[:n_ids], np.sort(lists[1]))
        yield line("sort_entity_ids", "one churned 2.5 M-entity list of the C2 scene (10 M-entity id space): set bits, then counts + look-back scan + expand",
                   med, best, 8 * n_ids + 3 * 4 * words, "scattered atomicOr (one per run of neighbouring ids) + two launches",
                   {"ids": n_ids, "id_space": id_space, "gids_per_s": n_ids / med / 1e6})
        med, best = timed(ctx, lambda: ctx.sort_entity_ids_batch(srcs, [n_ids] * 4, id_space, dsts), reps, flush)
        yield line("sort_entity_ids_batch", "the four churned 2.5 M-entity lists of the C2 scene in one batch (two launches)", med, best,
                   4 * (8 * n_ids + 3 * 4 * words), "scattered atomicOr + look-back scan",
                   {"ids": 4 * n_ids, "id_space": id_space, "gids_per_s": 4 * n_ids / med / 1e6,
                    "note": "a cull of these lists gains ~71 us from reading the sorted copies (0.216 -> 0.145 ms)"})
        for b in srcs + dsts:
            b.destroy()

    if want("c2live"):
        n_live, per_frame = 10_000_000, 500
        for sort_lists in (False, True):
            med, best, stats, vis, host_ms = live_churn_case(ctx, n_live, per_frame, max(reps, 8), sort_lists)
            yield line("frame_with_live_lists", f"C2 with sparse-set-backed lists that change every frame ({per_frame} erases + {per_frame} inserts per list and "
                       f"frame, after 10 % churn): set sync + " + ("sorted copies kept by deltas + cull of the copies" if sort_lists else "cull of the lists as they are")
                       + " - GPU time of the frame's enqueued work", med, best, n_live * 74 + vis * 4, "hbm",
                       {"instances": n_live, "visible": vis, "sorted_copy_full_rebuilds": stats[0], "sorted_copy_delta_updates": stats[1],
                        "host_ms_per_frame_encode": host_ms,  # wall time of the rvs_cull_static call (set sync preparation, tables, enqueues)
                        "l2": "inputs exceed L2"})

    if want("pyramid"):
        cam = scenes.default_camera(z_far=1000.0)
        depth = scenes.synth_depth(3840, 2160, cam)
        d = ctx.buffer_from_numpy(depth)
        pyr = ctx.create_pyramid(2160)
        med, best = timed(ctx, lambda: ctx.build_pyramid(pyr, d, 3840, 2160), reps, flush)
        algo = min(4 * 3840 * 2160, 16 * pyr.dim ** 2) + 4 * (4 * pyr.dim ** 2 - 1) // 3
        extra = {"dim": pyr.dim, "levels": pyr.levels, "l2": "flushed between launches",
                 "dram_bytes_expected": 4 * 3840 * 2160 + 2 * 4 * (4 * pyr.dim ** 2 - 1) // 3,
                 "dram_note": "the 2x2 footprints of the 1024^2 samples touch nearly every 32-B sector of the 33 MB source, and the "
                              "footprint-min chain (5.6 MB) is written beside the texels"}
        if cpu:
            import oracle

            workers = oracle.default_workers()
            ts = []
            for _ in range(3):
                t0 = time.perf_counter()
                oracle.pyramid_build(depth, 2160, nthreads=workers, taskflow=oracle.LIB_TF.exists())
                ts.append(time.perf_counter() - t0)
            extra["cpu"] = {"ms": float(np.median(ts)) * 1e3, "cores": workers, "kind": "port", "sample": "the full stage, median of 3"}
        yield line("pyramid", "3840x2160 depth -> 1024^2, 11 levels", med, best, algo, "hbm", extra)
        # the build as the fused encode issues it: tile levels only (persistent kernel), the rest rides in the cull that follows
        med, best = timed(ctx, lambda: ctx.build_pyramid(pyr, d, 3840, 2160, deferred=True), reps, flush)
        n0 = pyr.dim
        tile_written = 4 * sum((n0 >> l) ** 2 for l in range(6)) + 4 * (n0 + 1) ** 2  # mips 0..5 + footprint-min level 0
        yield line("pyramid_tile_levels", "3840x2160 depth -> mips 0..5 of 1024^2 + footprint-min level 0 (rve_depth_pyramid_build_deferred: what the "
                   "fused encode launches; mips 6..10 and footprint-min levels >= 1 are derived in the start-up of the cull that follows)",
                   med, best, min(4 * 3840 * 2160, 16 * n0 ** 2) + tile_written, "hbm + launch latency",
                   {"dim": pyr.dim, "l2": "flushed between launches", "dram_bytes_expected": 4 * 3840 * 2160 + tile_written,
                    "note": "an empty launch of this kernel's shape measures 8.6 us between the same events (profiles/r02f_pyramid_variants.txt)"})
        ctx.build_pyramid(pyr, d, 3840, 2160)  # leave the pyramid complete

    if want("c3"):
        nv, no, nb = 20_000, c3_objects, 64
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
        med, best = timed(ctx, lambda: ctx.skin(ubo, outs, ins, pose, wts), max(reps // 2, 5))
        algo = no * nv * 56 + no * nb * 64 + nv * 88
        extra = {"gverts_per_s": no * nv / (med * 1e-3) / 1e9}
        if cpu:
            import oracle

            workers = oracle.default_workers()
            so = min(no, 256)  # bounded sample: 256 characters = 5.1 M vertices
            h_pose = t_pose[: so * nb].cpu().numpy()
            h_outs = [np.zeros((so * nv, k), np.float32) for k in (3, 3, 3, 3, 2)]
            h_ins = [dd[k] for k in ("pos", "nrm", "tan", "bit", "uv")]
            ts = []
            for _ in range(3):
                t0 = time.perf_counter()
                oracle.skin(SkinningUBO(so, nv, nb, 0, 0, 0), h_outs, h_ins, h_pose, dd["weights"], nthreads=workers)
                ts.append(time.perf_counter() - t0)
            ms = float(np.median(ts)) * 1e3
            extra["cpu"] = {"ms": ms, "gverts_per_s": so * nv / ms / 1e6, "cores": workers, "kind": "port",
                            "sample": f"{so} of the {no} characters ({so * nv} vertices), median of 3"}
        yield line("skinning", f"C3: {no} objects x {nb} joints x {nv} verts (output {no * nv * 56 / 1e9:.2f} GB)", med, best, algo, "hbm", extra)
        del t_out, outs
        pu = SkinningPrepareUBO(0, 0, 60_000, nv, no, 0, 0)
        ind = ctx.create_buffer(no * 20)
        med, best = timed(ctx, lambda: ctx.skinned_drawcall_prepare(pu, ind), reps)
        yield line("skinned_drawcall_prepare", f"{no} objects", med, best, no * 20, "launch latency")

    if want("c4"):
        for grid in ((16, 9, 24), (12, 12, 24)):
            cam = scenes.default_camera()
            rng = np.random.default_rng(4)
            nl = 65_536
            z = -rng.uniform(0.2, 100.0, size=nl)
            pos = np.stack([rng.uniform(-1, 1, size=nl) * (-z) * 1.03, rng.uniform(-1, 1, size=nl) * (-z) * 0.6, z], 1).astype(np.float32)
            inten = np.exp(rng.uniform(np.log(0.05), np.log(5.0), size=nl)).astype(np.float32)
            h_pts = scenes.point_lights_bytes(pos, inten)
            pts = ctx.buffer_from_numpy(h_pts)
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
            med, best = timed(ctx, lambda: ctx.cluster_build(u, cl), reps)
            yield line("cluster_build", f"C4 grid {grid}", med, best, n * 24, "launch latency")

            ubos = []
            for v in views:
                a_v = GridAssignUBO()
                a_v.viewMat[:] = [float(x) for x in scenes.colmajor(v)]
                a_v.pointLightCount, a_v.spotLightCount = nl, 0
                ubos.append(a_v)

            def five_views():
                for a_v in ubos:
                    ctx.cluster_assign(a_v, cl, pts, None, n)
            med, best = timed(ctx, five_views, reps)
            ctx.set_profiling(True)  # the assign kernel alone (events inside the library), without the light prepare pass and launch gaps
            for _ in range(3):
                five_views()
            k_ms, k_n = ctx.profile_read(3)
            ctx.set_profiling(False)
            tests = 5 * n * nl
            extra = {"sphere_aabb_tests": tests, "gtests_per_s": tests / (med * 1e-3) / 1e9, "ms_per_view": med / 5,
                     "assign_kernel_ms_per_view": k_ms / max(k_n, 1),
                     "note": "ms_per_view = light prepare + assign as enqueued from Python (two launches per view; at ~20 us of GPU work per "
                             "view the loop is partly host-bound); sphere_aabb_tests counts every (cluster, light) pair, most of which the "
                             "per-CTA union-box prefilter rejects eight clusters at a time"}
            if cpu and grid == (16, 9, 24):
                import oracle

                workers = oracle.default_workers()
                h_cl = np.zeros(n * 160, np.uint8)
                oracle.cluster_build(u, h_cl)
                a.viewMat[:] = [float(x) for x in scenes.colmajor(views[1])]
                t0 = time.perf_counter()
                oracle.cluster_assign(a, h_cl, h_pts, None, n, nthreads=workers)
                ms = (time.perf_counter() - t0) * 1e3
                extra["cpu"] = {"ms_per_view": ms, "gtests_per_s": n * nl / ms / 1e6, "cores": workers, "kind": "port",
                                "sample": "one of the five views (the oracle exits a cluster early once it holds 16 lights, like the shader)"}
            yield line("cluster_assign", f"C4 grid {grid}, {nl} point lights, 5 views (prepare + assign per view)", med, best,
                       5 * (nl * 64 + n * 320), "fp32 issue + shared-memory bandwidth (not HBM)", extra)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--reps", type=int, default=20)
    ap.add_argument("--c3-objects", type=int, default=10_000)
    ap.add_argument("--only", default="")
    ap.add_argument("--no-cpu", action="store_true")
    args = ap.parse_args()
    only = set(args.only.split(",")) if args.only else None
    torch.cuda.set_device(0)
    stream = torch.cuda.Stream()
    ctx = Context(0, stream.cuda_stream)
    for d in run_stages(ctx, stream, args.reps, args.c3_objects, only, not args.no_cpu):
        print(json.dumps(d), flush=True)
    ctx.close()


if __name__ == "__main__":
    main()
