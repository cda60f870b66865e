#!/usr/bin/env python
"""How much of "bit-exact vs the oracle" depends on the oracle's arithmetic contract? (parity item 1)

GLSL fixes neither contraction (a*b+c may or may not fuse), nor the evaluation order of dot(), nor how a division is
carried out, so the reference's own result differs from GPU to GPU. The oracle pins ONE legal evaluation (oracle.cpp
header) and the CUDA kernels reproduce exactly that one. This tool rebuilds the oracle under the OTHER legal evaluations
and counts, on the full C2 scene (10 M instances, 3840x2160 depth -> 1024^2 pyramid), how many per-instance decisions
(visible or not; LOD chosen) differ from the pinned evaluation:

  no_fma        every a*b+c rounded twice (no fused multiply-add anywhere)
  true_div      vector / scalar divides each component (the contract multiplies by the correctly rounded reciprocal)
  dot_reversed  dot() accumulates from the last component to the first
  all_three     the three together

CPU only (it runs the oracle; nothing here touches the product path):
  python tools/contract_sensitivity.py [--instances 10000000] > profiles/r02_contract_sensitivity.json
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import subprocess
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))

import bench  # noqa: E402
import oracle  # noqa: E402
from oracle import host_tables  # noqa: E402
from ravengine_b200 import scenes  # noqa: E402
from tests.helpers import make_ubo  # noqa: E402

VARIANTS = {"no_fma": ["-DORC_VARIANT_NO_FMA"], "true_div": ["-DORC_VARIANT_TRUE_DIV"], "dot_reversed": ["-DORC_VARIANT_DOT_REVERSED"],
            "all_three": ["-DORC_VARIANT_NO_FMA", "-DORC_VARIANT_TRUE_DIV", "-DORC_VARIANT_DOT_REVERSED"]}


def build_variant(name: str) -> C.CDLL:
    out = ROOT / "oracle" / "_variants"
    out.mkdir(exist_ok=True)
    so = out / f"liboracle_{name}.so"
    fma = ["-mfma"] if " fma " in Path("/proc/cpuinfo").read_text() else []
    cmd = ["g++", "-std=c++20", "-O3", "-fPIC", "-shared", "-ffp-contract=off", "-fno-fast-math", *fma, "-pthread", *VARIANTS[name], "-o", str(so),
           str(ROOT / "oracle" / "oracle.cpp")]
    subprocess.run(cmd, check=True)
    return C.CDLL(str(so))


def run(lib, ubo, t, model, layers, flags, depth, workers):
    """pyramid + cull through `lib` (a build of oracle.cpp); returns per-instance LOD (or -1 when culled)"""
    dim, levels = oracle.pyramid_dims(min(bench.C2["depth"]))
    pyr = np.zeros(int(oracle.lib().orc_pyramid_texel_count(C.c_uint32(dim))), np.float32)
    h, w = depth.shape
    lib.orc_pyramid_build_mt(C.c_void_p(depth.ctypes.data), C.c_uint32(w), C.c_uint32(h), C.c_void_p(pyr.ctypes.data), C.c_uint32(dim), C.c_uint32(levels),
                             C.c_int(workers))
    bl = [None if b is None else b.copy() for b in t.bindless]
    table = (C.c_void_p * len(bl))(*[(b.ctypes.data if b is not None else None) for b in bl])
    groups = (t.total_entities + 63) // 64
    lib.orc_cull(C.byref(ubo), C.c_void_p(t.cubos.ctypes.data), C.c_void_p(model.ctypes.data), C.c_void_p(layers.ctypes.data),
                 C.c_void_p(flags.ctypes.data), C.c_void_p(pyr.ctypes.data), C.c_uint32(dim), C.c_uint32(levels), table, C.c_uint32(groups),
                 C.c_int(workers))
    lod = np.full(model.shape[0], -1, np.int8)
    for cb in t.cubos:
        n = int(cb["numObjects"])
        if not n:
            continue
        ind, ids = bl[int(cb["indirectOutputBufferBindlessHandle"])], bl[int(cb["idOutputBufferBindlessHandle"])]
        for k in range(int(cb["numLODs"])):
            cnt = int(ind[int(cb["indirectBufferOffset"]) + k, 1])
            start = int(cb["cullingBufferOffset"]) + n * k
            lod[ids[start:start + cnt]] = k
    return lod, pyr


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--instances", type=int, default=bench.C2["instances"])
    args = ap.parse_args()
    n = args.instances
    workers = oracle.default_workers()
    cam, depth = bench.camera_and_depth()
    model = np.zeros((n, 16), np.float32)
    layers, flags, ids = np.zeros(n, np.uint32), np.zeros(n, np.uint16), np.zeros(n, np.uint32)
    oracle.gen_instances(bench.gen_params(bench.C2["seed"]), 0, n, model, layers, flags, ids)
    m = bench.C2["materials"]
    bounds = [n * i // m for i in range(m + 1)]
    mats = [scenes.MaterialDraws([scenes.DrawCommand(scenes.simple_mesh(1.7320508 if i % 2 == 0 else 1.74, bench.C2["lod_distances"]),
                                                     ids[bounds[i]:bounds[i + 1]])]) for i in range(m)]
    t = host_tables.build_static_tables(mats)
    ubo = make_ubo(cam.view_proj_cm, cam.pos, t.num_cubos)
    base, base_pyr = run(oracle.lib(), ubo, t, model, layers, flags, depth, workers)
    out = {"scene": f"C2: {n} instances, seed {bench.C2['seed']}, 3840x2160 depth -> 1024^2 pyramid", "visible_pinned": int((base >= 0).sum()),
           "variants": {}}
    # proof that the variant builds really evaluate differently: skinned positions of a small case, raw float differences
    from ravengine_b200 import SkinningUBO

    rng = np.random.default_rng(0)
    sk = scenes.skinning_inputs(rng, 1000, 2, 16)
    sk_ins = [sk[k] for k in ("pos", "nrm", "tan", "bit", "uv")]

    def skin(lib):
        outs = [np.zeros((2000, k), np.float32) for k in (3, 3, 3, 3, 2)]
        u = SkinningUBO(2, 1000, 16, 0, 0, 0)
        lib.orc_skin(C.byref(u), *[C.c_void_p(o.ctypes.data) for o in outs], *[C.c_void_p(i.ctypes.data) for i in sk_ins],
                     C.c_void_p(sk["pose"].ctypes.data), C.c_void_p(sk["weights"].ctypes.data), C.c_int(1))
        return outs[0]

    sk_base = skin(oracle.lib())
    for name in VARIANTS:
        vlib = build_variant(name)
        lod, pyr = run(vlib, ubo, t, model, layers, flags, depth, workers)
        vis_flip = int(((lod >= 0) != (base >= 0)).sum())
        lod_flip = int(((lod >= 0) & (base >= 0) & (lod != base)).sum())
        out["variants"][name] = {"visibility_flips": vis_flip, "lod_flips": lod_flip, "pyramid_texels_differing": int((pyr != base_pyr).sum()),
                                 "instances_affected": vis_flip + lod_flip, "fraction": (vis_flip + lod_flip) / n,
                                 "variant_is_live_skinned_floats_differing": int((skin(vlib) != sk_base).sum())}
    worst = max(v["instances_affected"] for v in out["variants"].values())
    out["reading"] = (f"at most {worst} of the {n} decisions of this scene depend on the contract (the worst variant); variant_is_live shows "
                      "that a variant build really evaluates differently (raw skinned floats; dot_reversed / true_div do not occur in the skinning sums)")
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
