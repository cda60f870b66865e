"""Host-side mirror of the shared vertex arena and the `.rvem` container over the C-ABI
(include/rve_shim.h: rvs_arena_*, rvs_rvem_parse). Reference: src/RenderEngine_Allocate.cpp (AllocateMesh /
DeallocateMesh), include/RavEngine/Mesh.hpp:12-27 + src/MeshAsset.cpp:23-79 (reader), tools/rvemc/rvemc.cpp:246-287
(writer)."""
from __future__ import annotations

import ctypes as C

import numpy as np

from ._abi import load_library
from .api import Buffer, Context, _check

VERTEX_FREE, INDEX_FREE, VERTEX_ALLOCATED, INDEX_ALLOCATED = 0, 1, 2, 3
SKINNED, HAS_POSITIONS, HAS_NORMALS, HAS_TANGENTS, HAS_BITANGENTS, HAS_UV0, HAS_LIGHTMAP_UV = 1, 2, 4, 8, 16, 32, 64
WEIGHT = np.dtype([("joint", "<u4"), ("influence", "<f4")])  # VertexWeights::vweights::vw, four per vertex


class MeshView(C.Structure):  # rvs_mesh_view
    _fields_ = [(n, C.c_void_p) for n in ("positions", "normals", "tangents", "bitangents", "uv0", "lightmap_uvs", "indices")] + \
               [("num_vertices", C.c_uint32), ("num_indices", C.c_uint32)]


class RvemInfo(C.Structure):  # rvs_rvem_info
    _fields_ = [("num_vertices", C.c_uint32), ("num_indices", C.c_uint32), ("attributes", C.c_uint32), ("reserved", C.c_uint32)] + \
               [(n, C.c_uint64) for n in ("off_positions", "off_normals", "off_tangents", "off_bitangents", "off_uv0", "off_lightmap_uv",
                                          "off_indices", "mesh_bytes", "off_weights", "weights_bytes")]


def write_rvem(positions, normals, tangents, bitangents, uv0, indices, lightmap_uvs=None, weights=None) -> bytes:
    """The byte image tools/rvemc writes (rvemc.cpp:246-287): 16-byte header (the 13 meaningful bytes + padding),
    the five mandatory streams, the optional lightmap UVs, the indices, and for a skinned mesh VertexWeights[nVerts]."""
    nv = len(positions)
    attrs = HAS_POSITIONS | HAS_NORMALS | HAS_TANGENTS | HAS_BITANGENTS | HAS_UV0
    if lightmap_uvs is not None and len(lightmap_uvs):
        attrs |= HAS_LIGHTMAP_UV
    if weights is not None:
        attrs |= SKINNED
    head = b"rvem" + np.array([nv, len(indices)], "<u4").tobytes() + bytes([attrs, 0, 0, 0])
    parts = [head]
    for arr, w in ((positions, 3), (normals, 3), (tangents, 3), (bitangents, 3), (uv0, 2)):
        a = np.ascontiguousarray(arr, "<f4")
        assert a.shape == (nv, w)
        parts.append(a.tobytes())
    if attrs & HAS_LIGHTMAP_UV:
        parts.append(np.ascontiguousarray(lightmap_uvs, "<f4").tobytes())
    parts.append(np.ascontiguousarray(indices, "<u4").tobytes())
    if weights is not None:
        w = np.ascontiguousarray(weights, WEIGHT)
        assert w.shape == (nv, 4)
        parts.append(w.tobytes())
    return b"".join(parts)


def parse_rvem(blob: bytes) -> RvemInfo:
    """rvs_rvem_parse: validates the image and returns where its streams lie (host only, no device needed)."""
    lib = load_library()
    info = RvemInfo()
    buf = (C.c_char * len(blob)).from_buffer_copy(blob) if blob else (C.c_char * 1)()
    _check(lib, lib.rvs_rvem_parse(C.cast(buf, C.c_void_p), len(blob), C.byref(info)))
    return info


class MeshArena:
    """RenderEngine's shared mesh buffers + AllocateMesh / DeallocateMesh."""

    def __init__(self, ctx: Context, initial_verts: int = 0, initial_indices: int = 0):
        self.ctx, self.lib = ctx, ctx.lib
        h = C.c_void_p()
        _check(self.lib, self.lib.rvs_arena_create(ctx.h, initial_verts, initial_indices, C.byref(h)))
        self.h = h

    def destroy(self) -> None:
        if self.h:
            self.lib.rvs_arena_destroy(self.h)
            self.h = None

    def allocate(self, positions=None, normals=None, tangents=None, bitangents=None, uv0=None, lightmap_uvs=None, indices=()) -> int:
        keep = [None if a is None else np.ascontiguousarray(a, np.float32) for a in (positions, normals, tangents, bitangents, uv0, lightmap_uvs)]
        idx = np.ascontiguousarray(indices, np.uint32)
        nv = next((len(a) for a in keep if a is not None), 0)
        v = MeshView(*[a.ctypes.data if a is not None else None for a in keep], idx.ctypes.data if len(idx) else None, nv, len(idx))
        out = C.c_uint32()
        _check(self.lib, self.lib.rvs_arena_allocate_mesh(self.h, C.byref(v), C.byref(out)))
        return out.value

    def load_rvem(self, blob: bytes, skinned: bool = False):
        """-> (range, weights Buffer | None)"""
        buf = (C.c_char * len(blob)).from_buffer_copy(blob)
        rng, w = C.c_uint32(), C.c_uint32(0xFFFFFFFF)
        _check(self.lib, self.lib.rvs_arena_load_rvem(self.h, C.cast(buf, C.c_void_p), len(blob), 1 if skinned else 0, C.byref(rng), C.byref(w)))
        wb = None
        if skinned:
            wb = Buffer(self.ctx, w.value, int(self.lib.rve_buffer_size(self.ctx.h, w.value)))
        return rng.value, wb

    def deallocate(self, rng: int) -> None:
        _check(self.lib, self.lib.rvs_arena_deallocate_mesh(self.h, rng))

    def range(self, rng: int) -> tuple[int, int, int, int]:
        """(vertex start, vertex count, index start, index count) in elements"""
        v = [C.c_uint32() for _ in range(4)]
        _check(self.lib, self.lib.rvs_arena_range(self.h, rng, *[C.byref(x) for x in v]))
        return tuple(x.value for x in v)

    def buffers(self):
        """-> ([7 Buffers: pos, normal, tangent, bitangent, uv0, lightmapUV, index], vertex capacity, index capacity)"""
        hs = (C.c_uint32 * 7)()
        vc, ic = C.c_uint32(), C.c_uint32()
        _check(self.lib, self.lib.rvs_arena_buffers(self.h, hs, C.byref(vc), C.byref(ic)))
        return [Buffer(self.ctx, h, int(self.lib.rve_buffer_size(self.ctx.h, h))) for h in hs], vc.value, ic.value

    def list(self, which: int) -> list[tuple[int, int]]:
        n = self.lib.rvs_arena_list(self.h, which, None, 0)
        out = (C.c_uint32 * (2 * max(n, 1)))()
        self.lib.rvs_arena_list(self.h, which, out, n)
        return [(out[2 * i], out[2 * i + 1]) for i in range(n)]
