"""In-tree build of librve_b200.so (sm_100a only). Run: python -m ravengine_b200.build"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys
from pathlib import Path

HERE = Path(__file__).resolve().parent
CSRC = HERE / "csrc"
LIB = HERE / "librve_b200.so"
SOURCES = ["ctx.cu", "pyramid.cu", "cull.cu", "skinning.cu", "clusters.cu", "scenegen.cu", "frame.cu", "frame_cull.cu", "frame_skinned.cu", "sync.cu", "p2p.cu", "csm.cu", "mesh_arena.cu", "sort_ids.cu"]

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-lineinfo", "-O3", "-std=c++17",
    # parity with the oracle: no fused multiply-add, IEEE div/sqrt (nvcc defaults), no fast-math
    "-fmad=false", "-prec-div=true", "-prec-sqrt=true",
    "-Xcompiler", "-fPIC,-ffp-contract=off,-fvisibility=hidden,-O2",
    "-shared",
]


def nvcc_path() -> str:
    for cand in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", shutil.which("nvcc")):
        if cand and Path(cand).exists():
            return cand
    raise RuntimeError("nvcc not found; librve_b200.so cannot be built (there is no CPU fallback)")


def needs_build() -> bool:
    if not LIB.exists():
        return True
    t = LIB.stat().st_mtime
    deps = [CSRC / s for s in SOURCES] + [CSRC / "internal.h", CSRC / "frame_internal.h", HERE.parent / "include" / "rve_shim.h", HERE.parent / "include" / "rve_b200.h", Path(__file__)]
    return any(d.exists() and d.stat().st_mtime > t for d in deps)


def build_native(force: bool = False, verbose: bool = False, out: Path | None = None, defines: list[str] | None = None) -> Path:
    """`out` / `defines` build a tuning variant (e.g. -DRVE_CULL_STAGES=2) beside the product library."""
    if out is None and not force and not needs_build():
        return LIB
    out = out or LIB
    srcs = [str(CSRC / s) for s in SOURCES if (CSRC / s).exists()]
    cmd = [nvcc_path(), *NVCC_FLAGS, *(defines or []), *(["-Xptxas", "-v"] if verbose else []), "-o", str(out), *srcs]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        sys.stderr.write(r.stdout + r.stderr)
        raise RuntimeError("nvcc failed building librve_b200.so")
    if verbose:
        print(r.stdout + r.stderr)
    return out


if __name__ == "__main__":
    defs = [a for a in sys.argv[1:] if a.startswith("-D")]
    outs = [a for a in sys.argv[1:] if a.endswith(".so")]
    print(build_native(force=True, verbose="-v" in sys.argv, out=Path(outs[0]).resolve() if outs else None, defines=defs))
