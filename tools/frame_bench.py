#!/usr/bin/env python
"""Frame-level measurement: one camera + S shadow views over the C2 scene, the reference's pass
schedule run pass by pass through the shim (per camera: lit depth prepass cull + lit opaque cull over the same
material class, three light binnings; one cull per shadow view; src/RenderEngine_Draw.cpp:1769-1866, 2177-2209,
2567-2655) against rvs_frame_encode (one submission). Prints one JSON line per S.

  python tools/frame_bench.py [instances] [shadow view counts, e.g. 4,7,15] [lights]
"""
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
    counts = [int(x) for x in sys.argv[2].split(",")] if len(sys.argv) > 2 else [4, 7, 15]
    n_lights = int(sys.argv[3]) if len(sys.argv) > 3 else 1024
    sc = bench.DeviceScene(n, 0, 0)
    ctx, w, cam = sc.ctx, sc.world, sc.cam
    rng = np.random.default_rng(3)
    z = -rng.uniform(0.2, 200.0, size=n_lights)
    pos = np.stack([rng.uniform(-1, 1, size=n_lights) * (-z), rng.uniform(-1, 1, size=n_lights) * (-z) * 0.6, z], 1).astype(np.float32)
    inten = np.exp(rng.uniform(np.log(0.05), np.log(5.0), size=n_lights)).astype(np.float32)
    points = ctx.buffer_from_numpy(scenes.point_lights_bytes(pos, inten))
    grid, screen = (16, 9, 24), (3840, 2160)
    view = scenes.colmajor(np.eye(4))
    empty = ctx.create_pyramid(64)
    ring = [(np.cos(a) * 400.0, 150.0, np.sin(a) * 400.0) for a in np.linspace(0, 2 * np.pi, 16, endpoint=False)]
    for S in counts:
        shadow_cams = [scenes.default_camera(z_near=0.1, z_far=1000.0, eye=ring[i % 16], target=(0.0, 0.0, 0.0)) for i in range(S)]
        shadow_views = [(c.view_proj_cm, c.pos, 0xFFFFFFFF, True) for c in shadow_cams]
        cam_desc = dict(view_proj=cam.view_proj_cm, view=view, inv_proj=cam.inv_proj_cm, cam_pos=cam.pos, z_near=cam.z_near, z_far=cam.z_far,
                        pyramid=sc.pyr, depth=sc.depth, depth_wh=bench.C2["depth"], screen=screen)

        def reference_schedule():
            ctx.build_pyramid(sc.pyr, sc.depth, *bench.C2["depth"])
            for c in shadow_cams:  # shadow maps first (:1769-1866)
                w.cull_static(c.view_proj_cm, c.pos, empty, shadow=True)
            for _ in range(2):  # lit depth prepass + lit opaque: the same cull twice (:2177-2209)
                w.cull_static(cam.view_proj_cm, cam.pos, sc.pyr)
            for _ in range(3):  # renderFromPerspective<includeLighting>: prepass, opaque, transparent (:928-974)
                w.bin_lights(cam.inv_proj_cm, view, grid, screen, cam.z_near, cam.z_far, points, n_lights, None, 0)

        def one_submission():
            w.frame_encode([cam_desc], shadow_views, [empty] * S, points, n_lights, None, 0, grid)

        res = {}
        for name, fn in (("reference_schedule", reference_schedule), ("frame_encode", one_submission)):
            for _ in range(3):
                fn()
            ctx.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            l0 = ctx.launch_count
            reps = 5
            e0.record(sc.stream)
            for _ in range(reps):
                fn()
            e1.record(sc.stream)
            ctx.synchronize()
            res[name] = {"ms_per_frame": e0.elapsed_time(e1) / reps, "launches_per_frame": (ctx.launch_count - l0) / reps}
        print(json.dumps({"stage": "frame", "workload": f"C2 scene ({n} instances, 4 materials x 3 LODs), 1 camera + {S} shadow views, {n_lights} point lights",
                          "views": 1 + S, **res, "speedup": res["reference_schedule"]["ms_per_frame"] / res["frame_encode"]["ms_per_frame"]}), flush=True)


if __name__ == "__main__":
    main()
