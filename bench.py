#!/usr/bin/env python
"""bench.py — culled instances/s on BASELINE.json config C2 (10 M instances, 3 LODs, frustum + Hi-Z
against a 3840x2160 depth pyramid, indirect-command compaction), one rank per GPU.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
  torchrun ... bench.py --gpus N ...        (N > 1: instance-range sharding, weak scaling)

A step = one pass of the hot path for one view: depth pyramid build + host table build/upload +
indirect re-init + cull/LOD/compaction (+ for N>1 the per-slot count exchange and offset scan).
Prints ONE JSON line (rank 0). See DESIGN.md "Measurement" for every field.
"""
from __future__ import annotations

import argparse
import json
import math
import os
import subprocess
import sys
import threading
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

C2 = dict(instances=10_000_000, materials=4, lod_distances=(0.0, 40.0, 120.0), extent_xy=500.0, extent_y=50.0,
          z_near=0.1, z_far=1000.0, depth=(3840, 2160), frac_layer0=0.05, frac_nocull=0.05, seed=20240)
BYTES_PER_INSTANCE = 74  # 4 id + 4 layer + 2 flags + 64 transform


def peaks() -> tuple[float, str]:
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        return float(json.loads(p.read_text())["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=lambda: [self.lines.append(l) for l in self.proc.stdout], daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def stop(self) -> dict:
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        self.t.join(timeout=2)
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for l in self.lines:
            f = [x.strip() for x in l.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1]))
                mx.append(float(f[2]))
            except ValueError:
                continue
            for nm, v in zip(names, f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def camera_and_depth():
    from ravengine_b200 import scenes

    cam = scenes.default_camera(z_near=C2["z_near"], z_far=C2["z_far"])
    depth = scenes.synth_depth(C2["depth"][0], C2["depth"][1], cam)
    return cam, depth


def gen_params(seed: int):
    from ravengine_b200 import SceneGenParams

    return SceneGenParams(seed, C2["extent_xy"], C2["extent_y"], 0.5, 2.0, C2["frac_layer0"], C2["frac_nocull"], 0)


# --------------------------------------------------------------------------------------
# ours: CUDA path through the C-ABI / C++ shim
# --------------------------------------------------------------------------------------
class DeviceScene:
    """C2 scene resident in HBM on one rank: per-entity arrays generated on the device from
    (seed, global index); 4 materials x 1 mesh collection x 3 LODs over contiguous id ranges."""

    def __init__(self, n: int, rank: int, device: int, first_index: int | None = None, shard_of: tuple[int, int] | None = None):
        """first_index: global index of this scene's first instance (default rank * n: every rank its own 10 M instances).
        shard_of = (total, world): this rank's instance-range shard of ONE global scene of `total` instances - material m of
        the global scene owns global indices [total*m/M, total*(m+1)/M) and every rank holds its contiguous share of each
        material's list (SURVEY 8e partition A), so an instance keeps its material whatever the world size."""
        import torch

        from ravengine_b200 import Context
        from ravengine_b200.world import World

        self.torch, self.n, self.rank = torch, n, rank
        self.stream = torch.cuda.Stream(device=device)
        self.ctx = Context(device, self.stream.cuda_stream)
        dev = torch.device("cuda", device)
        with torch.cuda.stream(self.stream):
            self.t_model = torch.empty((n, 16), dtype=torch.float32, device=dev)
            self.t_layers = torch.empty(n, dtype=torch.int32, device=dev)
            self.t_flags = torch.empty(n, dtype=torch.int16, device=dev)
            self.t_ids = torch.empty(n, dtype=torch.int32, device=dev)
        c = self.ctx
        self.model, self.layers = c.wrap_tensor(self.t_model), c.wrap_tensor(self.t_layers)
        self.flags, self.ids = c.wrap_tensor(self.t_flags), c.wrap_tensor(self.t_ids)
        m_ = C2["materials"]
        if shard_of is None:
            c.gen_instances(gen_params(C2["seed"]), rank * n if first_index is None else first_index, n, self.model, self.layers, self.flags, self.ids)
        else:
            total, world = shard_of
            assert total % (m_ * world) == 0 and n == total // world, "the C5 scene splits evenly over materials and ranks"
            per = total // m_ // world  # instances of one material on one rank
            for i in range(m_):
                lo = i * per
                ids_view = c.wrap_tensor(self.t_ids[lo:lo + per])
                p_ = gen_params(C2["seed"])
                p_.id_base = lo  # local entity id = position in this rank's arrays (the generator writes entity arrays at id_base + i)
                c.gen_instances(p_, (total // m_) * i + rank * per, per, self.model, self.layers, self.flags, ids_view)
                c.synchronize()
                ids_view.destroy()
        self.cam, depth = camera_and_depth()
        self.depth_host = depth
        self.t_depth = torch.from_numpy(depth).to(dev)
        self.depth = c.wrap_tensor(self.t_depth)
        self.pyr = c.create_pyramid(min(C2["depth"]))
        self.id_slices = []
        self.world, self.mats = self._make_world()
        self.worlds, self.mats_of = [self.world], [self.mats]
        self.vp, self.cam_pos = self.cam.view_proj_cm, self.cam.pos
        self.slots = C2["materials"] * len(C2["lod_distances"])
        c.synchronize()

    def _make_world(self):
        """a render world over the scene's entity arrays: its own materials, commands and output (indirect / culling) buffers"""
        from ravengine_b200.world import World

        c, n = self.ctx, self.n
        world = World(c)
        world.set_entity_buffers(self.model, self.layers, self.flags)
        m = C2["materials"]
        bounds = [n * i // m for i in range(m + 1)]
        mats = []
        for i in range(m):
            cnt = bounds[i + 1] - bounds[i]
            sl = c.wrap_tensor(self.t_ids[bounds[i]:bounds[i + 1]]) if cnt else c.create_buffer(16)
            radius = 1.7320508 if i % 2 == 0 else 1.74  # meshes/cube.obj, meshes/sphere.obj
            mesh = world.mesh_collection(radius, C2["lod_distances"], [36, 24, 12], [0, 36, 60], [0, 24, 40])
            mat = world.material()
            world.add_command(mat, mesh, sl, cnt)
            self.id_slices.append(sl)
            mats.append(mat)
        return world, mats

    def add_frame_slot(self) -> int:
        """Second frame in flight: another world over the same entity arrays, i.e. a second set of indirect-command and culling
        buffers, so that frame f+1's encode does not have to wait until frame f's counts have been read (N > 1: the exchange
        kernel on the side stream). Returns the slot index for step(slot=..)."""
        w, mats = self._make_world()
        self.worlds.append(w)
        self.mats_of.append(mats)
        self.ctx.synchronize()
        return len(self.worlds) - 1

    def step(self, before_cull=None, slot: int = 0, fused: bool = False):
        """one pass of the hot path for one view: pyramid build + table build/upload + indirect re-init + cull, as one encode
        (rvs_cull_static_with_pyramid); RVE_BENCH_SPLIT_STEP=1 issues the two reference-shaped calls instead. slot: which
        frame slot's world (output buffers) the pass writes; fused=True with before_cull: the wait first, then the one encode
        (two frames in flight: the wait is for the exchange of two frames ago and does not stall)"""
        if fused and os.environ.get("RVE_BENCH_SPLIT_STEP") != "1":
            if before_cull:
                before_cull()
            self.worlds[slot].cull_static_with_pyramid(self.vp, self.cam_pos, self.pyr, self.depth, C2["depth"][0], C2["depth"][1])
            return
        if before_cull or os.environ.get("RVE_BENCH_SPLIT_STEP") == "1":
            # N > 1: the pyramid build does not touch the indirect buffers, so it is enqueued BEFORE the wait for last frame's
            # count exchange (before_cull) and runs while the slowest peer catches up (the fused order cost 15 us per step at N = 8)
            # (a complete build: here it runs in the shadow of that wait, so deferring its upper levels into the cull only
            # lengthens the cull - N=2: 165.5 us per step as it is, 167.6 us with rve_depth_pyramid_build_deferred; RVE_BENCH_DEFER=1)
            self.ctx.build_pyramid(self.pyr, self.depth, C2["depth"][0], C2["depth"][1], deferred=os.environ.get("RVE_BENCH_DEFER") == "1")
            if before_cull:
                before_cull()
            self.worlds[slot].cull_static(self.vp, self.cam_pos, self.pyr)
            return
        self.worlds[slot].cull_static_with_pyramid(self.vp, self.cam_pos, self.pyr, self.depth, C2["depth"][0], C2["depth"][1])

    def visible_counts(self, slot: int = 0) -> np.ndarray:
        out = []
        for mat in self.mats_of[slot]:
            _, ind, lods, _ = self.worlds[slot].material_outputs(mat)
            out.append(ind.download(np.uint32).reshape(-1, 5)[:lods, 1])
        return np.concatenate(out)

    def slot_plans(self, n: int | None = None):
        """sharding.SlotPlan per material for `n` instances split over the materials like this scene's"""
        from ravengine_b200.sharding import SlotPlan

        n = self.n if n is None else n
        m = C2["materials"]
        return [SlotPlan((n * (i + 1) // m - n * i // m,), (len(C2["lod_distances"]),)) for i in range(m)]

    def release(self):
        """free the scene's device memory (the C5 scene needs the room at N = 1)"""
        self.ctx.synchronize()
        for w in self.worlds:
            w.destroy()
        self.ctx.close()
        for k in ("t_model", "t_layers", "t_flags", "t_ids", "t_depth"):
            setattr(self, k, None)
        self.torch.cuda.empty_cache()


def cpu_sample_run(sample: int, reps: int, cam, depth, workers: int | None = None, warmup: int = 0):
    """The CPU path of the same step on the host cores: the oracle (CPU restatement of depthpyramidcopy.fsh / depthpyramid.csh /
    defaultcull.csh; TEST INFRASTRUCTURE, timed here as the reported baseline) through the reference's own executor (vendored
    taskflow) with the engine's worker count. One step = pyramid build + indirect re-init + cull/LOD/append over `sample`
    instances of the C2 generator. Returns (instances_per_s, cores, kind, sample_desc, per-step seconds after warm-up)."""
    import oracle
    from oracle import host_tables
    from ravengine_b200 import scenes
    from tests.helpers import make_ubo

    tf = oracle.LIB_TF.exists()
    workers = workers or oracle.default_workers()
    model = np.zeros((sample, 16), np.float32)
    layers, flags = np.zeros(sample, np.uint32), np.zeros(sample, np.uint16)
    ids = np.zeros(sample, np.uint32)
    oracle.gen_instances(gen_params(C2["seed"]), 0, sample, model, layers, flags, ids)
    m = C2["materials"]
    bounds = [sample * i // m for i in range(m + 1)]
    mats = []
    for i in range(m):
        mesh = scenes.simple_mesh(1.7320508 if i % 2 == 0 else 1.74, C2["lod_distances"])
        mats.append(scenes.MaterialDraws([scenes.DrawCommand(mesh, ids[bounds[i]:bounds[i + 1]])]))
    t = host_tables.build_static_tables(mats)
    ubo = make_ubo(cam.view_proj_cm, cam.pos, t.num_cubos)
    groups = (t.total_entities + 63) // 64
    times, t_pyr = [], []
    for _ in range(warmup + reps):
        t0 = time.perf_counter()
        pyr, dim, levels = oracle.pyramid_build(depth, min(C2["depth"]), nthreads=workers, taskflow=tf)
        t1 = time.perf_counter()
        for b in t.bindless:  # indirect re-init, like the reference's per-pass CopyBufferToBuffer
            if b is not None and b.ndim == 2:
                b[:, 1] = 0
        oracle.cull(ubo, t.cubos, model, layers, flags, pyr, dim, levels, t.bindless, groups, workers, tf)
        times.append(time.perf_counter() - t0)
        t_pyr.append(t1 - t0)
    times, t_pyr = times[warmup:], t_pyr[warmup:]
    best = float(np.median(times))
    visible = int(sum(int(b[:, 1].sum()) for b in t.bindless if b is not None and b.ndim == 2))
    desc = (f"{sample} instances of the C2 generator (seed {C2['seed']}){' = the full C2 scene' if sample == C2['instances'] else ''}; per step: "
            f"pyramid build ({float(np.median(t_pyr)) * 1e3:.1f} ms) + indirect re-init + cull/LOD/append, median of {reps}; "
            f"{'taskflow' if tf else 'std::thread'} executor, {workers} workers; visible {visible}")
    return sample / best, workers, "port", desc, times


def run_reference(args, rank: int):
    if rank != 0:
        return
    cam, depth = camera_and_depth()
    sample = args.cpu_sample or C2["instances"]
    v, cores, kind, desc, times = cpu_sample_run(sample, args.steps, cam, depth, warmup=args.warmup)
    ms = float(np.mean(times)) * 1e3
    v = sample / (ms / 1e3)
    line = {"impl": "reference", "metric": "culled instances/s", "value": v, "unit": "instances/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": "C2: 10M instances, 4 materials x 3 LODs {0,40,120}, frustum + Hi-Z (3840x2160 depth -> 1024^2 pyramid, "
                                   "rebuilt every step) + LOD + indirect compaction; CPU restatement of the shaders on host cores",
                       "instances_per_step": sample, "lods": 3, "depth": list(C2["depth"]), "same_config_as_gpu_arm": sample == C2["instances"]},
            "cpu_baseline": {"value": v, "unit": "instances/s", "cores": cores, "kind": kind, "sample": desc},
            "e2e": {"value": v, "unit": "instances/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0,
            "note": "the reference's own implementation needs a graphics API (unavailable on this box); this is the "
                    "CPU oracle port on all host worker threads the engine would use (hardware_concurrency-2), one rank "
                    "regardless of --gpus (the CPU path does not shard)"}
    print(json.dumps(line), flush=True)


def place_rank(local_rank: int, world_size: int, device_count: int, mode: str) -> int:
    """Which GPU a rank drives. "spread" (default): ranks are spread evenly over the box's GPUs (stride
    device_count / world_size) instead of packed onto the first N, so that a job smaller than the box uses every PCIe root /
    socket it can: host->device uploads are the e2e bound, and on the 8-GPU pool boxes GPUs 0-3 share one ~115 GB/s path to
    host memory while 0,2,4,6 reach 4 x 54 GB/s (profiles/r02_h2d_probe2_n8.jsonl). NVSwitch makes every pair of GPUs
    equally near for the count exchange. "packed": device = local rank."""
    if mode == "spread" and 0 < world_size < device_count and device_count % world_size == 0:
        return local_rank * (device_count // world_size)
    return local_rank


def run_ours(args, rank: int, world_size: int, local_rank: int):
    import torch

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — the render-prep path has no CPU fallback (use --impl reference)")
    device = place_rank(local_rank, world_size, torch.cuda.device_count(), args.placement)
    torch.cuda.set_device(device)
    dist = None
    if world_size > 1:
        import torch.distributed as dist

        dist.init_process_group("nccl", device_id=torch.device("cuda", device))
    n = args.instances
    sc = DeviceScene(n, rank, device)
    # host placement before any pinned allocation: this process onto the CPUs / NUMA node of its GPU's PCIe root
    affinity = sc.ctx.bind_host_to_device(cpus=not args.no_numa_bind, memory=not args.no_numa_bind)
    ex = None
    if world_size > 1 or args.force_exchange:
        from ravengine_b200.sharding import ShardedCull

        ex = ShardedCull(sc.ctx, sc.world, sc.mats, [len(C2["lod_distances"])] * len(sc.mats), world_size, sc.stream, rank=rank,
                         exchange=args.exchange)

    def barrier():
        if dist:
            dist.barrier()
        torch.cuda.synchronize()

    # N > 1: two frames in flight. Each frame slot has its own indirect / culling buffers, so a frame's encode does not wait
    # for the previous frame's count exchange (side stream) to have read them, and the step is the same fused encode as at N = 1.
    two_frames = world_size > 1 and not args.one_frame_in_flight
    if two_frames:
        ex.add_frame_slot(sc.worlds[sc.add_frame_slot()], sc.mats_of[1])
    frame = [0]

    def one_step():
        if two_frames:
            slot = frame[0] & 1
            frame[0] += 1
            dbg = os.environ.get("RVE_BENCH_EXCHANGE_DEBUG", "")  # timing experiments only: "nowait" skips begin_cull, "noexchange" the exchange
            sc.step(None if dbg else (lambda: ex.begin_cull(slot)), slot=slot, fused=True)
            if dbg != "noexchange":
                ex.exchange(slot=slot)
            return
        sc.step(ex.begin_cull if ex else None)
        if ex:
            ex.exchange()

    for _ in range(max(args.warmup, 3)):
        one_step()
    sampler = ClockSampler(device)
    if rank == 0:
        sampler.start()  # before the barrier: spawning nvidia-smi takes milliseconds and would skew rank 0's start
    barrier()
    launches0 = sc.ctx.launch_count
    ex_launches0 = ex.launch_count if ex else 0
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record(sc.stream)
    t_host0 = time.perf_counter()
    for _ in range(args.steps):
        one_step()
    if ex:
        ex.finish()  # the side stream's count exchange + offset scan belong to the timed region
    ev1.record(sc.stream)
    host_enqueue_ms = (time.perf_counter() - t_host0) * 1e3 / args.steps  # CPU time to enqueue one step (no sync inside)
    barrier()
    launches = sc.ctx.launch_count - launches0 + (ex.launch_count - ex_launches0 if ex else 0)
    ms_total = ev0.elapsed_time(ev1)
    clocks = sampler.stop() if rank == 0 else None
    by_rank = [ms_total / args.steps]
    if dist:
        t = torch.tensor([ms_total], device="cuda")
        allt = torch.zeros(world_size, device="cuda")
        dist.all_gather_into_tensor(allt, t)
        by_rank = [float(x) / args.steps for x in allt.tolist()]
        ms_total = float(allt.max().item())  # MAX over ranks
    ms_step = ms_total / args.steps
    counts = sc.visible_counts()
    visible = int(counts.sum())

    # --- per-kernel roofline pass: the cull kernel alone, bracketed by events inside the library ---
    sc.ctx.set_profiling(True)
    for _ in range(args.steps):
        sc.step()
    cull_ms, cull_n = sc.ctx.profile_read(0)
    pyr_ms, pyr_n = sc.ctx.profile_read(1)
    sc.ctx.set_profiling(False)
    cull_avg_ms = cull_ms / max(cull_n, 1)
    cull_by_rank = [cull_avg_ms]
    if dist:
        allk = torch.zeros(world_size, device="cuda")
        dist.all_gather_into_tensor(allk, torch.tensor([cull_avg_ms], device="cuda"))
        cull_by_rank = [float(x) for x in allk.tolist()]
    peak, peak_src = peaks()
    pyr_bytes = 4 * (4 * sc.pyr.dim ** 2 - 1) // 3
    algo_bytes = n * BYTES_PER_INSTANCE + visible * 4 + 36 * C2["materials"] + 20 * sc.slots + pyr_bytes
    achieved = algo_bytes / (cull_avg_ms * 1e-3) / 1e9
    traffic, traffic_src = None, None
    tf = ROOT / "profiles" / "cull_kernel_traffic.json"  # dram__bytes_read.sum + dram__bytes_write.sum of one `ncu --set full` launch
    if tf.exists() and n == C2["instances"]:
        tj = json.loads(tf.read_text())
        traffic, traffic_src = tj["traffic_bytes"], tj["source"]

    # --- e2e: host buffers through the shim, H2D + D2H inside the timed region ---
    if args.no_e2e:
        e2e = None
    else:
        e2e = run_e2e(sc, args, torch, barrier, dist, affinity)

    ex_mode = ex.mode if ex else None
    sc_slots, pyr_dim = sc.slots, sc.pyr.dim
    # --- device-side gather of the compacted ids for one consumer (rank 0), N > 1 only ---
    gather = run_gather(sc, ex, args, torch, barrier, dist, rank, world_size) if (ex and world_size > 1 and ex.mode == "p2p" and not args.no_gather) else None

    # --- C5: 512 camera views sharded by view range, 1 B instances sharded by instance range (fixed totals: strong scaling) ---
    c5 = None
    if not args.no_c5 and n == C2["instances"]:
        cam, depth_host = sc.cam, sc.depth_host
        if ex:
            ex.close()
        sc.release()
        sc = ex = None
        c5 = {"views512": run_c5_views(args, torch, barrier, dist, rank, world_size, device)}
        c5["instances_1B"] = run_c5_instances(args, torch, barrier, dist, rank, world_size, device)
    else:
        cam, depth_host = sc.cam, sc.depth_host

    if rank != 0:
        if dist:
            dist.destroy_process_group()
        return
    stages = None
    if world_size == 1 and not args.no_stages and n == C2["instances"]:
        from tools import stage_bench

        st_stream = torch.cuda.Stream(device=device)
        from ravengine_b200 import Context

        st_ctx = Context(device, st_stream.cuda_stream)
        stages = list(stage_bench.run_stages(st_ctx, st_stream, reps=10, only={"c1", "c2s", "c2churn", "c2sorted", "c2live", "pyramid", "c3", "c4"}, cpu=not args.no_cpu_baseline))
        st_ctx.close()
    cpu = None
    if not args.no_cpu_baseline:
        cpu_v, cores, kind, desc, _ = cpu_sample_run(args.cpu_sample or C2["instances"], 5, cam, depth_host, warmup=1)
        cpu = {"value": cpu_v, "unit": "instances/s", "cores": cores, "kind": kind, "sample": desc}
    total_instances = n * world_size
    slots = C2["materials"] * len(C2["lod_distances"])
    line = {
        "metric": "culled instances/s", "value": total_instances / (ms_step * 1e-3), "unit": "instances/s",
        "n_gpus": world_size, "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms_step,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": ("C2" if n == C2["instances"] else "C5 shard" if n >= 100_000_000 else "C2 distribution")
                               + f": {n / 1e6:g}M instances/GPU, 4 materials x 3 LODs {{0,40,120}}, frustum + Hi-Z (3840x2160 depth -> "
                               "1024^2 pyramid, rebuilt every step) + LOD + indirect compaction",
                   "instances_per_gpu": n, "ids": "identity order", "visible": visible,
                   "l2": f"inputs (74 B x {n / 1e6:g}M = {n * 74 / 1e6:.0f} MB) exceed the 126 MB L2; no explicit flush",
                   "device_of_rank0": device, "placement": args.placement,
                   "sharding": ("instance range per rank; per-slot counts exchanged by "
                                + ("one kernel over NVLink peer memory (peer stores + flags + scan)" if ex_mode == "p2p"
                                   else "NCCL all-gather + device offset scan")
                                + ("; two frames in flight (two sets of indirect / culling buffers: a frame's encode does not wait for "
                                   "the previous frame's exchange)" if two_frames else "; one frame in flight")) if world_size > 1 else "single GPU"},
        "roofline": {"bound": "hbm", "kernel": "cull_kernel", "achieved": achieved, "peak": peak, "unit": "GB/s",
                     "frac": achieved / peak, "traffic": traffic, "traffic_source": traffic_src, "peak_source": peak_src,
                     "algorithmic_bytes_per_launch": algo_bytes, "kernel_ms": cull_avg_ms, "slots": slots,
                     "frac_of_8000_datasheet": achieved / 8000.0, "pyramid_kernel_ms": pyr_ms / max(pyr_n, 1),
                     "note": ("fused encode: the pyramid launch stops after the tile levels (mips 0..5, footprint-min level 0) and cull_kernel "
                              "derives the rest in its start-up - kernel_ms contains that work (about +2.3 us; 144.4 us = 0.80 of peak without it: "
                              "profiles/r02f_pyramid_variants.txt), algorithmic_bytes_per_launch does not count it"
                              if (world_size == 1 or two_frames) and os.environ.get("RVE_BENCH_SPLIT_STEP") != "1" and os.environ.get("RVE_PYRAMID_DEFER") != "0"
                              else "pyramid build (complete, two launches) and cull are separate calls here - the build is enqueued before the wait "
                                   "for last frame's count exchange and runs in its shadow")},
        "e2e": e2e, "gpu_launches": int(launches), "clocks": clocks, "host_enqueue_ms_per_step": host_enqueue_ms,
        "ms_per_step_by_rank": by_rank, "cull_kernel_ms_by_rank": cull_by_rank,
        "visible_per_slot": counts.tolist(),
    }
    if cpu:
        line["cpu_baseline"] = cpu
    if gather:
        line["gather"] = gather
    if c5:
        line["c5"] = c5
    if stages:
        line["stages"] = stages
    print(json.dumps(line), flush=True)
    if dist:
        dist.destroy_process_group()


def run_gather(sc, ex, args, torch, barrier, dist, rank, world_size):
    """N > 1: the step followed by the device-side gather of every rank's compacted ids into rank 0's GLOBAL culling buffers
    (rve_gather_ids_p2p over NVLink peer memory, chained behind the count exchange on the side stream; no host sync).
    Checked in place: per slot, the ids that landed on rank 0 are exactly `total` distinct values whose sum and sum of squares
    (mod 2^64) equal the all-reduced sums of every rank's local slices shifted by its id base."""
    n = sc.n
    from ravengine_b200.sharding import SlotPlan

    local_plan = sc.slot_plans(n)
    global_plan = [SlotPlan(tuple(c * world_size for c in p.counts), p.lods) for p in local_plan]  # every rank holds the same split
    # entity-id lists of the bench are contiguous per material on every rank: rank r's material m holds local ids
    # [n*m/M, n*(m+1)/M); globally that is NOT one contiguous range, so the "global id" of a local id is rank*n + id
    ex.setup_gather(local_plan, global_plan, [0], rank * n)
    steps = max(3, min(args.steps, 10))

    def one():
        sc.step(ex.begin_cull)
        b = ex.exchange()
        ex.gather(b)
        return b

    for _ in range(2):
        one()
    ex.finish()
    barrier()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record(sc.stream)
    for _ in range(steps):
        b = one()
    ex.finish()
    ev1.record(sc.stream)
    barrier()
    ms = ev0.elapsed_time(ev1) / steps
    # the gather alone: exchange + gather on an otherwise idle GPU, events on the side stream
    g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    g0.record(ex.side)
    b = ex.exchange()
    ex.gather(b)
    g1.record(ex.side)
    ex.finish()
    barrier()
    ms_alone = g0.elapsed_time(g1)
    t = torch.tensor([ms, ms_alone], device="cuda")
    allt = torch.zeros((world_size, 2), device="cuda")
    dist.all_gather_into_tensor(allt.view(-1), t)
    ms, ms_alone = float(allt[:, 0].max().item()), float(allt[:, 1].max().item())
    # verification
    totals = ex.t_tot[b].cpu().numpy().view(np.uint32).astype(np.int64)
    offs = ex.t_off[b].cpu().numpy().view(np.uint32).astype(np.int64)
    sums = np.zeros((len(totals), 2), np.uint64)
    s = 0
    mine = 0
    for m, plan in enumerate(local_plan):
        ids = sc.world.material_outputs(sc.mats[m])[0].download(np.uint32)
        for start in plan.slot_starts():
            cnt = int((offs[rank + 1, s] if rank + 1 < world_size else totals[s]) - offs[rank, s])
            v = ids[start:start + cnt].astype(np.uint64) + np.uint64(rank * n)
            sums[s] = (v.sum(dtype=np.uint64), (v * v).sum(dtype=np.uint64))
            mine += cnt
            s += 1
    ts = torch.from_numpy(sums.view(np.int64)).cuda()
    parts = torch.zeros((world_size,) + tuple(ts.shape), dtype=torch.int64, device="cuda")
    dist.all_gather_into_tensor(parts.view(-1), ts.view(-1))
    want = parts.cpu().numpy().view(np.uint64).sum(axis=0, dtype=np.uint64)
    ok, landed = None, None
    if rank == 0:
        ok, landed, s = True, 0, 0
        for m, plan in enumerate(global_plan):
            got = ex.global_ids[m].download(np.uint32)
            for start in plan.slot_starts():
                v = got[start:start + int(totals[s])].astype(np.uint64)
                ok = ok and len(np.unique(v)) == len(v) and v.sum(dtype=np.uint64) == want[s, 0] and (v * v).sum(dtype=np.uint64) == want[s, 1]
                landed += len(v)
                s += 1
        ok = bool(ok) and not ex.timed_out()
    return {"ms_per_step_with_gather": ms, "ms_exchange_plus_gather_alone": ms_alone, "steps": steps, "consumer": 0,
            "ids_landed_on_consumer": landed, "nvlink_bytes_per_step": None if landed is None else 4 * (landed - mine),
            "verified": ok, "check": "per slot: distinct ids, sum and sum of squares (mod 2^64) equal the all-reduced local slices",
            "note": "one gather kernel per rank writes its slices straight into rank 0's global culling buffers "
                    "(cullingBufferOffset + numObjects*lod + offset[rank][slot], defaultcull.csh:305-312) over peer memory"}


def c5_cameras(num_views: int):
    """BASELINE C5's 32 x 16 camera grid over the scene, every camera looking down -Z like the C2 camera"""
    from ravengine_b200 import scenes

    grid = [(x, z) for z in np.linspace(-400, 400, 16) for x in np.linspace(-400, 400, 32)]
    return [scenes.default_camera(z_near=C2["z_near"], z_far=C2["z_far"], eye=(grid[i % 512][0], 5.0, grid[i % 512][1]),
                                  target=(grid[i % 512][0], 5.0, grid[i % 512][1] - 1.0)) for i in range(num_views)]


def run_c5_views(args, torch, barrier, dist, rank, world_size, device):
    """Partition B: `--c5-views` (512) independent camera views over ONE 10 M-instance scene (the C2 scene of rank 0, held by every
    rank), each rank culling its share of the views with the view-batched kernel (no collective; views dealt round-robin so
    the visible work balances). Fixed total work: strong scaling."""
    from ravengine_b200.sharding import ShardedViews

    sc = DeviceScene(C2["instances"], 0, device, first_index=0)
    V = args.c5_views
    cams = c5_cameras(V)
    views = [(c.view_proj_cm, c.pos, 0xFFFFFFFF, False) for c in cams]
    sv = ShardedViews(sc.world, views, [sc.pyr] * V, rank, world_size)
    sc.step()
    sv.cull_chunk(0)  # warm-up: buffers for up to 128 views per material
    barrier()
    reps = 2
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record(sc.stream)
    for _ in range(reps):
        sv.cull_all()
    ev1.record(sc.stream)
    barrier()
    ms = ev0.elapsed_time(ev1) / reps
    vis = 0
    for i in range(len(sv.chunks)):  # visible (instance, view) pairs, outside the timed region
        owned = sv.cull_chunk(i)
        sc.ctx.synchronize()
        for m in sc.mats:
            _, ind, lods, _ = sc.world.material_outputs(m)
            vis += int(ind.download(np.uint32).reshape(-1, 5)[:len(owned) * lods, 1].astype(np.int64).sum())
    by_rank = [ms]
    if dist:
        t = torch.tensor([ms, float(vis)], device="cuda", dtype=torch.float64)
        allt = torch.zeros((world_size, 2), device="cuda", dtype=torch.float64)
        dist.all_gather_into_tensor(allt.view(-1), t)
        by_rank = [float(x) for x in allt[:, 0].tolist()]
        ms, vis = float(allt[:, 0].max().item()), int(allt[:, 1].sum().item())
    out = {"value": sc.n * V / (ms * 1e-3), "unit": "instance-views/s", "views": V, "instances": sc.n, "ms_per_pass": ms,
           "ms_per_pass_by_rank": by_rank, "views_per_rank": len(sv.owned), "visible_instance_views": vis, "scaling": "strong",
           "sharding": "views dealt round-robin over the ranks (partition B), no collective; every rank holds the whole scene",
           "check": "visible_instance_views is the same at every N (tests/test_gpu_sharding.py pins the per-rank views to the one-GPU batch)"}
    sc.release()
    return out


def run_c5_instances(args, torch, barrier, dist, rank, world_size, device):
    """Partition A at C5 size: a FIXED scene of `--c5-instances` (1 B) instances split by instance range over the ranks
    (strong scaling; 74 GB of inputs at N = 1), per-slot counts exchanged by the peer-memory kernel."""
    total = args.c5_instances
    n = total // world_size
    sc = DeviceScene(n, rank, device, shard_of=(total, world_size))
    ex = None
    if world_size > 1:
        from ravengine_b200.sharding import ShardedCull

        ex = ShardedCull(sc.ctx, sc.world, sc.mats, [len(C2["lod_distances"])] * len(sc.mats), world_size, sc.stream, rank=rank, exchange=args.exchange)

    def one():
        sc.step(ex.begin_cull if ex else None)
        return ex.exchange() if ex else 0

    for _ in range(2):
        one()
    barrier()
    steps = 5
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record(sc.stream)
    for _ in range(steps):
        b = one()
    if ex:
        ex.finish()
    ev1.record(sc.stream)
    barrier()
    ms = ev0.elapsed_time(ev1) / steps
    sc.ctx.set_profiling(True)
    for _ in range(3):
        sc.step()
    k_ms, k_n = sc.ctx.profile_read(0)
    sc.ctx.set_profiling(False)
    k_ms /= max(k_n, 1)
    counts = sc.visible_counts().astype(np.int64)
    by_rank, k_by_rank = [ms], [k_ms]
    if dist:
        t = torch.tensor([ms, k_ms], device="cuda")
        allt = torch.zeros((world_size, 2), device="cuda")
        dist.all_gather_into_tensor(allt.view(-1), t)
        by_rank, k_by_rank = [float(x) for x in allt[:, 0].tolist()], [float(x) for x in allt[:, 1].tolist()]
        ms = float(allt[:, 0].max().item())
        totals = ex.t_tot[b].cpu().numpy().view(np.uint32).astype(np.int64)  # the exchange's global instanceCount per slot
        tc = torch.from_numpy(counts).cuda()
        dist.all_reduce(tc)
        assert np.array_equal(tc.cpu().numpy(), totals), "count exchange totals differ from the all-reduced local counts"
        counts = totals
    peak, _ = peaks()
    vis_local = int(sc.visible_counts().sum())
    algo = n * BYTES_PER_INSTANCE + vis_local * 4 + 4 * (4 * sc.pyr.dim ** 2 - 1) // 3
    out = {"value": total / (ms * 1e-3), "unit": "instances/s", "instances_total": total, "instances_per_rank": n, "ms_per_step": ms,
           "ms_per_step_by_rank": by_rank, "cull_kernel_ms_by_rank": k_by_rank, "steps": steps, "scaling": "strong",
           "cull_kernel_frac_of_measured_peak_rank0": algo / (k_ms * 1e-3) / 1e9 / peak,
           "visible_per_slot_global": counts.tolist(), "visible_total": int(counts.sum()),
           "sharding": "instance range per rank (partition A); per-slot counts by one peer-memory kernel" if world_size > 1 else "single GPU",
           "check": "visible_per_slot_global is the same at every N (same generator, same camera)"}
    if ex:
        ex.close()
    sc.release()
    return out


class HostScene:
    """The per-entity inputs of a DeviceScene as the engine holds them on the HOST: BufferedVRAMVector mirrors
    (pinned host array + modification flags + device-private buffer, csrc/sync.cu) of worldTransforms (48-B affine or
    64-B mat4 rows), renderLayers, perObjectAttributes and the entity-id list. Created after the thread was bound to
    the GPU's NUMA node, so the pinned pages are local to the GPU's PCIe root."""

    def __init__(self, sc: DeviceScene, affine: bool, shared_host: bool = False):
        from ravengine_b200.vram_vector import SYNC_FULL, BufferedVRAMVector

        n, ctx = sc.n, sc.ctx
        self.sc, self.affine = sc, affine
        self.model = BufferedVRAMVector(ctx, 48 if affine else 64, n, affine_rows=affine, shared_host=shared_host)
        host = self.model.host(np.float32).reshape(n, 12 if affine else 16)
        for a in range(0, n, 1 << 20):  # device -> pinned host in 1 M-row pieces (no second host copy of the scene)
            rows = sc.t_model[a:a + (1 << 20)]
            host[a:a + rows.shape[0]] = (rows.view(-1, 4, 4)[:, :, :3].reshape(-1, 12) if affine else rows).cpu().numpy()
        self.layers = BufferedVRAMVector(ctx, 4, n)
        self.layers.host(np.int32)[:] = sc.t_layers.cpu().numpy()
        self.flags = BufferedVRAMVector(ctx, 2, n)
        self.flags.host(np.int16)[:] = sc.t_flags.cpu().numpy()
        self.ids = BufferedVRAMVector(ctx, 4, n)
        self.ids.host(np.int32)[:] = sc.t_ids.cpu().numpy()
        for v in (self.model, self.layers, self.flags, self.ids):
            v.encode_sync(SYNC_FULL)  # first sync after Resize: everything goes up once, outside any timed region
        ctx.synchronize()
        # the scene's commands now read the vectors' private buffers
        sc.world.set_entity_buffers(self.model.private_buffer, self.layers.private_buffer, self.flags.private_buffer)
        m = C2["materials"]
        bounds = [n * i // m for i in range(m + 1)]
        import ctypes as C

        self._id_views = []
        idbuf = self.ids.private_buffer
        for i, mat in enumerate(sc.mats):
            cnt = bounds[i + 1] - bounds[i]
            if cnt:
                view = ctx.wrap_pointer(idbuf.ptr + 4 * bounds[i], 4 * cnt, keepalive=self.ids)
                sc.world.set_command_entities(mat, 0, view, cnt)
                self._id_views.append(view)

    def bytes_model(self) -> int:
        return self.sc.n * (48 if self.affine else 64)

    def close(self):
        sc = self.sc
        sc.world.set_entity_buffers(sc.model, sc.layers, sc.flags)
        for i, mat in enumerate(sc.mats):
            if i < len(sc.id_slices):
                m = C2["materials"]
                cnt = sc.n * (i + 1) // m - sc.n * i // m
                if cnt:
                    sc.world.set_command_entities(mat, 0, sc.id_slices[i], cnt)
        sc.ctx.synchronize()
        for view in self._id_views:
            view.destroy()
        for v in (self.model, self.layers, self.flags, self.ids):
            v.destroy()


def _timed_host_loop(step, steps, barrier, torch, dist):
    for i in range(2):
        step(i)
    barrier()
    t0 = time.perf_counter()
    for i in range(steps):
        step(i)
    barrier()
    dt = time.perf_counter() - t0
    ws = 1
    by_rank = [dt / steps * 1e3]
    if dist:
        ws = dist.get_world_size()
        t = torch.tensor([dt], device="cuda")
        allt = torch.zeros(ws, device="cuda")
        dist.all_gather_into_tensor(allt, t)
        by_rank = [float(x) / steps * 1e3 for x in allt.tolist()]
        dt = float(allt.max().item())  # MAX over ranks
    return dt, ws, by_rank


def run_e2e(sc: DeviceScene, args, torch, barrier, dist, affinity):
    """The same step through the shim with HOST inputs, H2D + D2H inside the timed region.
    Headline leg: every transform is dirty every step (the worst case of the reference's EncodeSync: all entities moved)
    and goes up as a 48-B affine row (RVS_VECTOR_AFFINE_ROWS: glm::mat4x3 on the host, the reference's 64-B mat4 in the
    private buffer); renderLayers / perObjectAttributes / entity ids live in flag-tracked vectors like the reference's
    (src/World.cpp:330-378 touches only worldTransforms for entities that moved), are synced every step and upload only
    what was written - nothing in steady state. The indirect commands (the step's result) are read back.
    by_layout repeats the leg with the reference's unmodified 64-B host rows; by_dirty_fraction with 10 % / 1 % of the rows
    flagged per step (GPU-pull sync)."""
    from ravengine_b200.vram_vector import SYNC_FULL, SYNC_FULL_PULL, SYNC_PULL

    n = sc.n
    lib, ctx = sc.ctx.lib, sc.ctx
    lods = len(C2["lod_distances"])
    steps = max(2, min(args.steps, 10))
    out = {}
    full_mode = {"copy": SYNC_FULL, "pull": SYNC_FULL_PULL}[args.e2e_upload]
    # the 64-B-row variant only at N = 1: the pool's 8-GPU VMs have 15 GB of host memory for eight ranks' pinned arrays
    legs = [("affine48", True)] + ([("mat4_64", False)] if not args.e2e_headline_only and not dist else [])
    for name, affine in legs:
        hs = HostScene(sc, affine)
        outs = [sc.world.material_outputs(m)[1] for m in sc.mats]
        res = [np.zeros(lods * 5, np.uint32) for _ in outs]
        d2h = sum(r.nbytes for r in res)

        def step(_i):
            hs.model.encode_sync(full_mode)   # 100 % of the transforms
            hs.layers.encode_sync(SYNC_PULL)  # flag-tracked: nothing written -> nothing uploaded
            hs.flags.encode_sync(SYNC_PULL)
            hs.ids.encode_sync(SYNC_PULL)
            sc.step()
            for ind, r in zip(outs, res):
                rc = lib.rve_buffer_download(ctx.h, ind.handle, 0, r.ctypes.data, r.nbytes)  # synchronises the stream
                assert rc == 0

        dt, ws, by_rank = _timed_host_loop(step, steps, barrier, torch, dist)
        counts = np.concatenate([r.reshape(-1, 5)[:, 1] for r in res])
        # the link's practical rate for this buffer: the transform upload alone, timed with events on the stream
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        best = 0.0
        for _ in range(3):
            barrier()
            e0.record(sc.stream)
            hs.model.encode_sync(full_mode)
            e1.record(sc.stream)
            e1.synchronize()
            best = max(best, hs.bytes_model() / (e0.elapsed_time(e1) * 1e-3) / 1e9)
        h2d = hs.bytes_model()
        leg = {"value": n * ws * steps / dt, "unit": "instances/s", "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
               "ms_per_step": dt / steps * 1e3, "ms_per_step_by_rank": by_rank, "steps": steps, "h2d_gbs_achieved": h2d / (dt / steps) / 1e9,
               "h2d_gbs_upload_alone": best, "visible": int(counts.sum()), "upload": args.e2e_upload}
        if dist:
            allb = torch.zeros(ws, device="cuda")
            dist.all_gather_into_tensor(allb, torch.tensor([best], device="cuda"))
            leg["h2d_gbs_upload_alone_by_rank"] = [float(x) for x in allb.tolist()]
        out[name] = leg
        if name == "affine48" and not args.e2e_headline_only:
            out["by_dirty_fraction"] = run_e2e_dirty(sc, hs, args, torch, barrier, dist)
        hs.close()
    e2e = dict(out["affine48"])
    if dist and not args.no_balance:
        bal = run_e2e_balanced(sc, args, torch, barrier, dist, out["affine48"]["h2d_gbs_upload_alone_by_rank"], out["affine48"]["visible"])
        if bal is None:
            e2e["forwarding"] = "not used: every rank's upload rate is within 15 % of every other's"
        elif bal["verified"] and bal["value"] > e2e["value"]:
            plain = {k: e2e[k] for k in ("value", "ms_per_step", "ms_per_step_by_rank", "h2d_gbs_achieved", "h2d_gbs_upload_alone_by_rank")}
            for k in ("value", "ms_per_step", "ms_per_step_by_rank", "steps"):
                e2e[k] = bal[k]
            e2e["h2d_gbs_achieved"] = e2e["h2d_bytes_per_step"] / (bal["ms_per_step"] * 1e-3) / 1e9
            e2e["forwarding"] = {"used": True, "verified": True, "pairs": bal["pairs"], "split_search_ms_per_step": bal["split_search_ms_per_step"],
                                 "upload": bal["upload"], "without_forwarding": plain,
                                 "note": "ranks behind the slower host links upload the head of their rows; their partner on a faster link reads the "
                                         "tail from the owner's (shared, pinned) host block over its own PCIe link and writes it into the owner's "
                                         "private buffer over NVLink; same bytes per rank per step"}
        else:
            e2e["forwarding"] = {"used": False, "verified": bal["verified"], "balanced_ms_per_step": bal["ms_per_step"], "pairs": bal["pairs"]}
    e2e["dirty_fraction"] = 1.0
    e2e["note"] = ("per step: 100 % of the transforms re-uploaded from pinned host memory as 48-B affine rows (expanded to the reference's "
                   "64-B mat4 layout on the device; exact h2d_bytes_per_step = 48 x instances), renderLayers / perObjectAttributes / entity "
                   "ids in flag-tracked vectors synced every step (unchanged -> 0 B, the reference's policy), indirect commands read back; "
                   "PCIe-bound: compare h2d_gbs_achieved with h2d_gbs_upload_alone")
    e2e["host_affinity"] = affinity
    if "mat4_64" in out:
        e2e["by_layout"] = {"mat4_64": out["mat4_64"]}
        e2e["by_layout_note"] = "same leg with the reference's unmodified 64-B mat4 host rows (no host-side change at all)"
    if "by_dirty_fraction" in out:
        e2e["by_dirty_fraction"] = out["by_dirty_fraction"]
        e2e["by_dirty_fraction_note"] = ("same step, only rows flagged since the last frame are uploaded (the reference's EncodeSync policy, "
                                         "GPU-pull variant: one kernel reads the flags and the flagged 48-B rows over PCIe); keys = fraction flagged per step")
    return e2e


def pair_ranks(rates: list[float], min_ratio: float = 1.15) -> dict[int, tuple[int, str]]:
    """Upload forwarding plan from the per-rank upload rates measured with every rank uploading at once: the i-th slowest
    rank is paired with the i-th fastest while the faster link is at least `min_ratio` x the slower one.
    Returns {rank: (partner, "owner" | "helper")}; ranks without an entry upload on their own."""
    order = sorted(range(len(rates)), key=lambda r: (rates[r], r))
    plan = {}
    for i in range(len(order) // 2):
        slow, fast = order[i], order[-1 - i]
        if rates[slow] > 0 and rates[fast] / rates[slow] >= min_ratio:
            plan[slow], plan[fast] = (fast, "owner"), (slow, "helper")
    return plan


def run_e2e_balanced(sc: DeviceScene, args, torch, barrier, dist, rates: list[float], want_visible: int):
    """The headline e2e leg with upload forwarding: ranks behind a slow host link (on the pool's 8-GPU
    boxes GPUs 0-3 share one path to host memory) upload only the head of their transform rows; a partner rank on a fast link
    maps the owner's host block, reads the tail over ITS PCIe link and writes it, expanded, into the owner's private buffer
    over NVLink, concurrently with its own upload. Per step and pair: owner signals "private buffer free", helper forwards
    and signals "rows landed", owner's cull waits for that flag. Bytes per step per rank are unchanged (48 x instances)."""
    from ravengine_b200 import Context
    from ravengine_b200.vram_vector import SYNC_FULL, SYNC_PULL

    rank, ws = dist.get_rank(), dist.get_world_size()
    n, ctx, lib = sc.n, sc.ctx, sc.ctx.lib
    if os.environ.get("RVE_BENCH_FAKE_RATES"):  # development: exercise the forwarding path on a box whose links are equal
        rates = [float(x) for x in os.environ["RVE_BENCH_FAKE_RATES"].split(",")][:ws]
    plan = pair_ranks(rates)
    if not plan:
        return None
    hs = HostScene(sc, True, shared_host=True)
    flags, status = ctx.create_buffer(64), ctx.create_buffer(16)
    flags.zero()
    status.zero()
    ctx.synchronize()
    pid, fd, nbytes = hs.model.share()
    priv = hs.model.private_buffer
    mine = {"pid": pid, "fd": fd, "bytes": nbytes, "flags": ctx.ipc_export(flags), "priv": ctx.ipc_export(priv), "priv_bytes": priv.nbytes}
    everyone = [None] * ws
    dist.all_gather_object(everyone, mine)
    role = plan.get(rank)
    k = n
    hctx = hstream = host_ptr = peer_priv = peer_flags = my_flags_h = hstatus = None
    if role:
        partner, what = role
        slow, fast = (rank, partner) if what == "owner" else (partner, rank)
        k0 = int(2 * n * rates[slow] / (rates[slow] + rates[fast]))  # rows the owner keeps: both links finish together
        k = max(0, min(n, k0 // 64 * 64))
        info = everyone[partner]
        if what == "owner":
            peer_flags = ctx.ipc_import(info["flags"], 64)
        else:
            hstream = torch.cuda.Stream(device=torch.device("cuda", ctx.device))
            hctx = Context(ctx.device, hstream.cuda_stream)
            host_ptr = hctx.host_import(info["pid"], info["fd"], info["bytes"])
            peer_priv = hctx.ipc_import(info["priv"], info["priv_bytes"])
            peer_flags = hctx.ipc_import(info["flags"], 64)
            my_flags_h = hctx.wrap_pointer(flags.ptr, 64)
            hstatus = hctx.wrap_pointer(status.ptr, 16)
    outs = [sc.world.material_outputs(m)[1] for m in sc.mats]
    lods = len(C2["lod_distances"])
    res = [np.zeros(lods * 5, np.uint32) for _ in outs]
    epoch = [0]

    def step(_i):
        nonlocal k
        epoch[0] += 1
        e = epoch[0]
        if role and role[1] == "helper":  # first: the partner's tail rows start crossing this rank's link right away
            hctx.flag_wait(my_flags_h, 0, e, hstatus)
            hctx.forward_rows(host_ptr, 48, True, k, n - k, peer_priv)
            hctx.flag_signal(peer_flags, 1, e)
        if role and role[1] == "owner":
            ctx.flag_signal(peer_flags, 0, e)  # behind last frame's cull on this stream: the private buffer is free
            hs.model.encode_sync_rows(SYNC_FULL, 0, k)
        else:
            hs.model.encode_sync(SYNC_FULL)
        hs.layers.encode_sync(SYNC_PULL)
        hs.flags.encode_sync(SYNC_PULL)
        hs.ids.encode_sync(SYNC_PULL)
        if role and role[1] == "owner":
            ctx.flag_wait(flags, 1, e, status)  # the forwarded rows have landed
        sc.step()
        for ind, r in zip(outs, res):
            rc = lib.rve_buffer_download(ctx.h, ind.handle, 0, r.ctypes.data, r.nbytes)  # synchronises the stream
            assert rc == 0

    steps = max(2, min(args.steps, 10))
    # The split computed from the rates assumes the links keep those rates once the bytes move; they do not quite (the copy
    # engine and the zero-copy kernel share the helper's link, the box has one ceiling for all links): try the computed
    # share of forwarded rows and its neighbours in short loops (a small hill climb), keep the best.
    tried = {}
    if role:
        fwd0 = n - k
    def try_bias(bias):
        nonlocal k
        if role:
            k = max(0, min(n, (n - int(fwd0 * bias)) // 64 * 64))
        dt_b, _, _ = _timed_host_loop(step, 3, barrier, torch, dist)
        tried[bias] = dt_b / 3 * 1e3  # max over ranks: every rank takes the same decisions below

    try_bias(1.0)
    try_bias(1.2)
    if tried[1.2] < tried[1.0]:  # more forwarding helps: keep going while it does
        for bias in (1.4, 1.6, 1.8):
            try_bias(bias)
            if tried[bias] > min(v for b_, v in tried.items() if b_ != bias):
                break
    else:
        try_bias(0.8)
    best_bias = min(tried, key=tried.get)
    if role:
        k = max(0, min(n, (n - int(fwd0 * best_bias)) // 64 * 64))
    dt, _, by_rank = _timed_host_loop(step, steps, barrier, torch, dist)
    visible = int(np.concatenate([r.reshape(-1, 5)[:, 1] for r in res]).sum())
    # checks: same visible set as the unbalanced leg, no timeout, and the forwarded rows are the host rows (sampled)
    ok = visible == want_visible and int(status.download(np.uint32)[0]) == 0
    if role and role[1] == "owner" and k < n:
        got = priv.download(np.float32, offset=64 * (n - 1000), nbytes=64 * 1000).reshape(1000, 16)
        want = hs.model.host(np.float32).reshape(n, 12)[n - 1000:]
        ok = ok and np.array_equal(got.reshape(1000, 4, 4)[:, :, :3].reshape(1000, 12), want) and np.all(got[:, 3::4] == np.array([0, 0, 0, 1], np.float32))
    t_ok = torch.tensor([1 if ok else 0], device="cuda")
    dist.all_reduce(t_ok, op=dist.ReduceOp.MIN)
    roles = [None] * ws
    dist.all_gather_object(roles, {"rank": rank, "role": role[1] if role else "alone", "partner": role[0] if role else None, "own_rows": k})
    barrier()
    if hctx:
        hctx.host_release(host_ptr, everyone[role[0]]["bytes"])
        hctx.close()
    hs.close()
    return {"value": n * ws * steps / dt, "unit": "instances/s", "h2d_bytes_per_step": int(48 * n), "d2h_bytes_per_step": int(sum(r.nbytes for r in res)),
            "ms_per_step": dt / steps * 1e3, "ms_per_step_by_rank": by_rank, "steps": steps, "visible": visible, "verified": bool(int(t_ok.item())),
            "pairs": roles, "split_search_ms_per_step": tried, "upload": "copy engine for a rank's own rows, one zero-copy kernel (partner's PCIe link -> NVLink) for forwarded rows"}


def run_e2e_dirty(sc: DeviceScene, hs: HostScene, args, torch, barrier, dist, fractions=(0.10, 0.01)):
    """Per step, inside the timed region: flag a random `fraction` of the rows (the simulation's SetValueAt calls), EncodeSync (GPU pull),
    the hot-path step, read the indirect commands back."""
    from ravengine_b200.vram_vector import SYNC_PULL

    n = sc.n
    vec = hs.model
    lib, ctx = sc.ctx.lib, sc.ctx
    outs = [sc.world.material_outputs(m)[1] for m in sc.mats]
    lods = len(C2["lod_distances"])
    res = [np.zeros(lods * 5, np.uint32) for _ in outs]
    d2h = sum(r.nbytes for r in res)
    rng = np.random.default_rng(7)
    out = {}
    for frac in fractions:
        k = max(1, int(n * frac))
        sets = [np.sort(rng.choice(n, size=k, replace=False)) for _ in range(4)]

        def step(i):
            vec.tracking()[sets[i % 4]] = 1  # the simulation's SetValueAt calls of this frame
            vec.encode_sync(SYNC_PULL)
            hs.layers.encode_sync(SYNC_PULL)
            hs.flags.encode_sync(SYNC_PULL)
            hs.ids.encode_sync(SYNC_PULL)
            sc.step()
            for ind, r in zip(outs, res):
                rc = lib.rve_buffer_download(ctx.h, ind.handle, 0, r.ctypes.data, r.nbytes)  # synchronises the stream
                assert rc == 0

        steps = max(2, min(args.steps, 10))
        dt, ws, _ = _timed_host_loop(step, steps, barrier, torch, dist)
        out[f"{frac:g}"] = {"value": n * ws * steps / dt, "unit": "instances/s", "ms_per_step": dt / steps * 1e3,
                            "dirty_rows_per_step": k, "h2d_bytes_per_step": int(k * vec.elem_bytes + n),  # flagged rows + 1 B of flag per row
                            "sync": "RVS_SYNC_PULL", "d2h_bytes_per_step": int(d2h), "steps": steps}
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--instances", type=int, default=C2["instances"], help="instances per GPU")
    ap.add_argument("--cpu-sample", type=int, default=0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true", help="skip the host-buffer e2e legs (C5-sized shards: 9 GB of pinned host memory per rank)")
    ap.add_argument("--e2e-upload", default="copy", choices=["copy", "pull"],
                    help="100 %-dirty transform upload: copy engine into staging + expand kernels (RVS_SYNC_FULL) or one zero-copy kernel (RVS_SYNC_FULL_PULL)")
    ap.add_argument("--e2e-headline-only", action="store_true", help="skip the by_layout / by_dirty_fraction e2e variants")
    ap.add_argument("--placement", default="spread", choices=["spread", "packed"], help="rank -> GPU mapping when the box has more GPUs than ranks")
    ap.add_argument("--one-frame-in-flight", action="store_true",
                    help="N>1: one set of output buffers (r02c behaviour): every frame's encode waits for the previous frame's count exchange")
    ap.add_argument("--no-balance", action="store_true", help="N>1: never forward part of a slow rank's upload through a faster rank's link")
    ap.add_argument("--no-numa-bind", action="store_true", help="leave the process where the launcher put it (diagnostics)")
    ap.add_argument("--no-gather", action="store_true", help="N>1: skip the device-side id gather leg")
    ap.add_argument("--no-c5", action="store_true", help="skip the C5 sub-results (512 views / 1 B instances)")
    ap.add_argument("--no-stages", action="store_true", help="N=1: skip the per-stage block (C1, C2 shuffled, pyramid, C3, C4)")
    ap.add_argument("--c5-instances", type=int, default=1_000_000_000, help="total instances of the C5 instance-range scene")
    ap.add_argument("--c5-views", type=int, default=512)
    ap.add_argument("--force-exchange", action="store_true", help="N=1: run the count exchange machinery anyway (diagnostics)")
    ap.add_argument("--exchange", default="auto", choices=["auto", "p2p", "nccl"], help="N>1: per-slot count exchange")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world_size = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank)
        return
    run_ours(args, rank, world_size, local_rank)


if __name__ == "__main__":
    main()
