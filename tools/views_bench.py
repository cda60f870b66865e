#!/usr/bin/env python
"""development aid / C5 'views' measurement: V views over N instances, one rve_cull_views launch (through the shim)
against V single-view culls. Prints one JSON line per configuration."""
import json
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import bench  # noqa: E402
import torch  # noqa: E402
from ravengine_b200 import scenes  # noqa: E402


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 10_000_000
    counts = [int(x) for x in sys.argv[2].split(",")] if len(sys.argv) > 2 else [4, 16, 32]
    mode = sys.argv[3] if len(sys.argv) > 3 else "inward"  # inward: cameras look at the scene centre (40 % visible);
    # grid: BASELINE C5's 32 x 16 camera grid, every camera looking down -Z like the C2 camera
    sc = bench.DeviceScene(n, 0, 0)
    ctx = sc.ctx
    sc.step()
    ctx.synchronize()
    grid = [(x, z) for z in np.linspace(-400, 400, 16) for x in np.linspace(-400, 400, 32)]  # the 512-view camera grid
    for V in counts:
        if mode == "grid":
            cams = [scenes.default_camera(z_near=0.1, z_far=1000.0, eye=(grid[i % 512][0], 5.0, grid[i % 512][1]),
                                          target=(grid[i % 512][0], 5.0, grid[i % 512][1] - 1.0)) for i in range(V)]
        else:
            cams = [scenes.default_camera(z_near=0.1, z_far=1000.0, eye=(grid[(i * 37) % 512][0], 5.0, grid[(i * 37) % 512][1]),
                                          target=(0.0, 0.0, 0.0) if i else (grid[0][0], 5.0, grid[0][1] - 1.0)) for i in range(V)]
        views_all = [(c.view_proj_cm, c.pos, 0xFFFFFFFF, False) for c in cams]
        chunks = [views_all[i:i + 128] for i in range(0, V, 128)]  # one rve_cull_views call takes up to 128 views
        pyrs = [sc.pyr] * 128  # the views share one pyramid here (per-view pyramids are supported; build cost is separate)
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
        reps = 3 if V <= 64 else 1
        for ch in chunks[:1]:
            sc.world.cull_static_views(ch, pyrs[:len(ch)])
        ev[0].record(sc.stream)
        vis = 0
        for _ in range(reps):
            for ch in chunks:
                sc.world.cull_static_views(ch, pyrs[:len(ch)])
        ev[1].record(sc.stream)
        ctx.synchronize()
        for ch in chunks:  # visible count (outside the timed region; the buffers are reused between chunks)
            sc.world.cull_static_views(ch, pyrs[:len(ch)])
            ctx.synchronize()
            vis += int(sum(sc.world.material_outputs(m)[1].download(np.uint32).reshape(-1, 5)[:, 1].sum() for m in sc.mats))
        for c in cams[:2]:
            sc.world.cull_static(c.view_proj_cm, c.pos, sc.pyr)
        ev[2].record(sc.stream)
        for _ in range(reps):
            for c in cams:
                sc.world.cull_static(c.view_proj_cm, c.pos, sc.pyr)
        ev[3].record(sc.stream)
        ctx.synchronize()
        batched, separate = ev[0].elapsed_time(ev[1]) / reps, ev[2].elapsed_time(ev[3]) / reps
        print(json.dumps({"stage": "cull_views", "instances": n, "views": V, "cameras": mode, "us_per_view_batched": batched / V * 1e3, "ms_batched": batched, "ms_separate_culls": separate,
                          "speedup": separate / batched, "g_instance_views_per_s": n * V / batched / 1e6,
                          "visible_instance_views": vis}), flush=True)


if __name__ == "__main__":
    main()
