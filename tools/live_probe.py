#!/usr/bin/env python
"""Host-time split of a frame whose sparse-set-backed entity lists change (500 erases + 500 inserts per list): the set
operations, the batched EncodeSync of the four sets, the rvs_cull_static encode.  python tools/live_probe.py [sorted]"""
import sys, time
sys.path.insert(0, ".")
import numpy as np, torch
from ravengine_b200 import Context, SceneGenParams, scenes
from ravengine_b200.world import World, SORT_ENTITY_LISTS
from ravengine_b200.vram_vector import SparseSet, encode_sync_batch
st = torch.cuda.Stream(); ctx = Context(0, st.cuda_stream)
n = 10_000_000
dev = torch.device("cuda", 0)
with torch.cuda.stream(st):
    t_model = torch.empty((n, 16), dtype=torch.float32, device=dev); t_layers = torch.empty(n, dtype=torch.int32, device=dev)
    t_flags = torch.empty(n, dtype=torch.int16, device=dev); t_ids = torch.empty(n, dtype=torch.int32, device=dev)
model, layers, flags, ids = (ctx.wrap_tensor(t) for t in (t_model, t_layers, t_flags, t_ids))
ctx.gen_instances(SceneGenParams(20240, 500.0, 50.0, 0.5, 2.0, 0.05, 0.05, 0), 0, n, model, layers, flags, ids); ctx.synchronize()
cam = scenes.default_camera(z_near=0.1, z_far=1000.0)
w = World(ctx, SORT_ENTITY_LISTS if len(sys.argv) > 1 else 0); w.set_entity_buffers(model, layers, flags)
rng = np.random.default_rng(1); per = n // 4; sets = []; spare = []
for i in range(4):
    base = np.arange(i * per, (i + 1) * per, dtype=np.uint32); ss = SparseSet(ctx); live = base[: per - per // 50]; ss.emplace_many(live)
    sets.append(ss); spare.append(list(base[per - per // 50:]))
    mesh = w.mesh_collection(1.73, (0.0, 40.0, 120.0), [36, 24, 12], [0, 36, 60], [0, 24, 40]); w.add_command_set(w.material(), mesh, ss)
pyr = ctx.create_pyramid(2160); d = ctx.buffer_from_numpy(scenes.synth_depth(3840, 2160, cam)); ctx.build_pyramid(pyr, d, 3840, 2160)
for _ in range(2): w.cull_static(cam.view_proj_cm, cam.pos, pyr)
ctx.synchronize()
for f in range(5):
    t0 = time.perf_counter()
    for i, ss in enumerate(sets):
        members = ss.dense(); out = np.unique(members[rng.integers(0, len(members), size=500)]); ss.erase_many(out)
        take = [spare[i].pop() for _ in range(len(out))]; ss.emplace_many(np.asarray(take, dtype=np.uint32)); spare[i] = list(out) + spare[i]
    t1 = time.perf_counter()
    encode_sync_batch(sets)
    t2 = time.perf_counter()
    w.cull_static(cam.view_proj_cm, cam.pos, pyr)
    t3 = time.perf_counter()
    ctx.synchronize()
    print(f"frame {f}: set ops {1e3*(t1-t0):.2f} ms, batch sync {1e3*(t2-t1):.3f} ms, cull_static {1e3*(t3-t2):.3f} ms")
