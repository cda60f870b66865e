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
