#!/usr/bin/env python
"""Generates tests/golden/rvem.npz from oracle/_ref/libref_rvem.so - the reference mesh compiler's own writer
(SerializeMeshPart, tools/rvemc/rvemc.cpp:233-292, compiled unchanged by `make -C oracle refrvem` where /root/reference
exists; oracle/ref_rvem.cpp). Three assets: a static mesh, one with lightmap UVs, a skinned one. The .npz holds the
inputs and the exact bytes the reference wrote. Run from the repo root: `python tests/golden/make_golden_rvem.py`."""
from __future__ import annotations

import ctypes as C
import sys
import tempfile
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))
OUT = Path(__file__).resolve().parent
WEIGHT = np.dtype([("joint", "<u4"), ("influence", "<f4")])


def ref_lib() -> C.CDLL:
    p = ROOT / "oracle" / "_ref" / "libref_rvem.so"
    if not p.exists():
        raise FileNotFoundError(f"{p} missing: run `make -C oracle refrvem` in a container that has /root/reference")
    return C.CDLL(str(p))


def make_inputs(seed: int, nv: int, ni: int, lightmap: bool, skinned: bool) -> dict:
    rng = np.random.default_rng(seed)
    d = {k: rng.standard_normal((nv, 3)).astype(np.float32) for k in ("positions", "normals", "tangents", "bitangents")}
    d["uv0"] = rng.random((nv, 2), dtype=np.float32)
    d["indices"] = rng.integers(0, nv, ni).astype(np.uint32)
    d["lightmap"] = rng.random((nv, 2), dtype=np.float32) if lightmap else None
    if skinned:
        w = np.zeros((nv, 4), WEIGHT)
        w["joint"] = rng.integers(0, 64, (nv, 4))
        w["influence"] = rng.random((nv, 4), dtype=np.float32)
        d["weights"] = w
    else:
        d["weights"] = None
    return d


def run_ref(d: dict) -> bytes:
    VP = C.c_void_p
    p = lambda a: VP(a.ctypes.data) if a is not None else None  # noqa: E731
    with tempfile.TemporaryDirectory() as tmp:
        path = Path(tmp) / "sub" / "asset.rvem"  # the writer creates the parent directories itself
        rc = ref_lib().ref_rvem_write(str(path).encode(), C.c_uint32(len(d["positions"])), C.c_uint32(len(d["indices"])), p(d["positions"]),
                                      p(d["normals"]), p(d["tangents"]), p(d["bitangents"]), p(d["uv0"]), p(d["lightmap"]), p(d["indices"]),
                                      p(d["weights"].view(np.uint32)) if d["weights"] is not None else None, C.c_int(d["weights"] is not None))
        assert rc == 0
        return path.read_bytes()


CASES = {"static": (11, 37, 90, False, False), "lightmap": (12, 64, 201, True, False), "skinned": (13, 50, 150, False, True)}


def main():
    out = {}
    for name, (seed, nv, ni, lm, sk) in CASES.items():
        d = make_inputs(seed, nv, ni, lm, sk)
        out[f"{name}_bytes"] = np.frombuffer(run_ref(d), np.uint8)
        for k, v in d.items():
            if v is not None:
                out[f"{name}_{k}"] = v.view(np.uint32) if k == "weights" else v
    np.savez_compressed(OUT / "rvem.npz", **out)
    print({k: v.shape for k, v in out.items() if k.endswith("_bytes")})


if __name__ == "__main__":
    main()
