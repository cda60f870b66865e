"""ctypes loader for the CPU oracle (TEST INFRASTRUCTURE ONLY — see oracle.h).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference leg may
import this package. PARITY UNPINNED: the reference has no tests/fixtures for this path.
"""
from __future__ import annotations

import ctypes as C
import subprocess
from pathlib import Path

import numpy as np

from ravengine_b200._abi import (CullingUBO, CullingUBOinstance, GridAssignUBO, GridBuildUBO, SceneGenParams,
                                 SkinningPrepareUBO, SkinningUBO)

HERE = Path(__file__).resolve().parent
LIB = HERE / "liboracle.so"
LIB_TF = HERE / "_ref" / "liboracle_tf.so"
VP, u32, u64, f32 = C.c_void_p, C.c_uint32, C.c_uint64, C.c_float


def _cpu_has_fma() -> bool:
    try:
        return " fma " in Path("/proc/cpuinfo").read_text()
    except OSError:
        return False


def build(force: bool = False) -> None:
    stale = not LIB.exists() or LIB.stat().st_mtime < max((HERE / f).stat().st_mtime for f in ("oracle.cpp", "oracle.h", "Makefile"))
    if force or stale:
        subprocess.run(["make", "-C", str(HERE)], check=True, capture_output=True)
    if not _cpu_has_fma():  # a prebuilt -mfma library would fault here: rebuild with libm's fmaf (same bits)
        marker = HERE / ".nofma_built"
        if not marker.exists():
            subprocess.run(["make", "-C", str(HERE), "-B", "FMA="], check=True, capture_output=True)
            marker.write_text("built without -mfma\n")


_libs: dict[bool, C.CDLL] = {}


def lib(taskflow: bool = False) -> C.CDLL:
    """taskflow=True loads the build parallelised with the reference's vendored taskflow (oracle/_ref)."""
    if taskflow and not LIB_TF.exists():
        taskflow = False
    if taskflow in _libs:
        return _libs[taskflow]
    build()
    l = C.CDLL(str(LIB_TF if taskflow else LIB))
    l.orc_pyramid_texel_count.restype = u64
    l.orc_pyramid_level_offset.restype = u64
    l.orc_sample_min.restype = f32
    l.orc_sample_min.argtypes = [VP, u32, u32, f32, f32]
    l.orc_ceil_log2.restype = f32
    l.orc_ceil_log2.argtypes = [f32]
    l.orc_gen_instances.argtypes = [C.POINTER(SceneGenParams), u64, u32, VP, VP, VP, VP]
    _libs[taskflow] = l
    return l


def _p(a: np.ndarray | None):
    return C.c_void_p(a.ctypes.data) if a is not None else None


def default_workers() -> int:
    return lib().orc_default_workers()


def pyramid_dims(width: int) -> tuple[int, int]:
    d, l = u32(), u32()
    lib().orc_pyramid_dims(u32(width), C.byref(d), C.byref(l))
    return d.value, l.value


def pyramid_level_offset(dim: int, level: int) -> int:
    return int(lib().orc_pyramid_level_offset(u32(dim), u32(level)))


def pyramid_build(depth: np.ndarray, width_for_dim: int | None = None, nthreads: int = 1, taskflow: bool = False) -> tuple[np.ndarray, int, int]:
    """depth: (h, w) float32. Returns (flat mip chain, dim, levels). nthreads > 1: rows dealt to worker threads (same values)."""
    h, w = depth.shape
    dim, levels = pyramid_dims(min(w, h) if width_for_dim is None else width_for_dim)
    out = np.zeros(int(lib().orc_pyramid_texel_count(u32(dim))), dtype=np.float32)
    depth = np.ascontiguousarray(depth, dtype=np.float32)
    if nthreads > 1:
        lib(taskflow).orc_pyramid_build_mt(_p(depth), u32(w), u32(h), _p(out), u32(dim), u32(levels), C.c_int(nthreads))
    else:
        lib().orc_pyramid_build(_p(depth), u32(w), u32(h), _p(out), u32(dim), u32(levels))
    return out, dim, levels


def cull(ubo: CullingUBO, cubos: np.ndarray, model: np.ndarray, layers: np.ndarray, flags: np.ndarray,
         pyramid: np.ndarray, dim: int, levels: int, bindless: list[np.ndarray | None], num_groups_x: int,
         nthreads: int = 1, taskflow: bool = False) -> None:
    """cubos: structured bytes (n*36). bindless[i] = host array for handle i (outputs are written in place)."""
    table = (VP * max(len(bindless), 1))(*[(b.ctypes.data if b is not None else None) for b in bindless])
    lib(taskflow).orc_cull(C.byref(ubo), _p(cubos), _p(model), _p(layers), _p(flags), _p(pyramid), u32(dim), u32(levels),
                           table, u32(num_groups_x), C.c_int(nthreads))


def skin(ubo: SkinningUBO, outs: list[np.ndarray], ins: list[np.ndarray], pose: np.ndarray, weights: np.ndarray,
         nthreads: int = 1) -> None:
    lib().orc_skin(C.byref(ubo), *[_p(o) for o in outs], *[_p(i) for i in ins], _p(pose), _p(weights), C.c_int(nthreads))


def skin_f64(ubo: SkinningUBO, outs: list[np.ndarray], ins: list[np.ndarray], pose: np.ndarray, weights: np.ndarray) -> None:
    lib().orc_skin_f64(C.byref(ubo), *[_p(o) for o in outs[:4]], *[_p(i) for i in ins[:4]], _p(pose), _p(weights))


def skinned_drawcall_prepare(ubo: SkinningPrepareUBO, commands: np.ndarray) -> None:
    lib().orc_skinned_drawcall_prepare(C.byref(ubo), _p(commands))


def cluster_build(ubo: GridBuildUBO, clusters: np.ndarray) -> None:
    lib().orc_cluster_build(C.byref(ubo), _p(clusters))


def cluster_assign(ubo: GridAssignUBO, clusters: np.ndarray, points: np.ndarray | None, spots: np.ndarray | None,
                   num_clusters: int, nthreads: int = 1) -> None:
    lib().orc_cluster_assign(C.byref(ubo), _p(clusters), _p(points), _p(spots), u32(num_clusters), C.c_int(nthreads))


def gen_instances(params: SceneGenParams, first_index: int, count: int, model: np.ndarray, layers: np.ndarray,
                  flags: np.ndarray, ids: np.ndarray) -> None:
    lib().orc_gen_instances(C.byref(params), u64(first_index), u32(count), _p(model), _p(layers), _p(flags), _p(ids))
