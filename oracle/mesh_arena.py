"""CPU oracle of the shared vertex arena and the `.rvem` container — TEST INFRASTRUCTURE ONLY.

Restates src/RenderEngine_Allocate.cpp:12-260 (AllocateMesh / DeallocateMesh / ReallocateGeneric) over plain Python
lists and numpy arrays, and src/MeshAsset.cpp:23-79 + src/MeshAssetSkinned.cpp:38-56 (the `.rvem` reader).
PARITY UNPINNED: the reference ships no `.rvem` file and no allocator test; its sources for this row need RGL and
the engine headers (unbuildable here). Pins we have: hand-worked allocation sequences (tests/test_mesh_arena_oracle.py).
One deliberate difference, shared with the product and stated in DESIGN.md: on a growth all six vertex streams are
re-packed from the true old starts (the reference re-packs streams 2..6 from already-updated starts).
"""
from __future__ import annotations

import ctypes
import ctypes.util

import numpy as np

_libm = ctypes.CDLL(ctypes.util.find_library("m") or "libm.so.6")
for _f in ("logf", "ceilf"):
    getattr(_libm, _f).restype = ctypes.c_float
    getattr(_libm, _f).argtypes = [ctypes.c_float]
_libm.powf.restype = ctypes.c_float
_libm.powf.argtypes = [ctypes.c_float, ctypes.c_float]

STRIDE_WORDS = (3, 3, 3, 3, 2, 2, 1)  # pos, normal, tangent, bitangent, uv0, lightmapUV (f32 words); index (u32)


def closest_power_of_1_5(v: int) -> int:
    """closest_power_of<float>(v, 1.5f) (include/RavEngine/Common3D.hpp:106-109) with libm's float functions."""
    f32 = np.float32
    q = f32(_libm.logf(float(f32(v)))) / f32(_libm.logf(1.5))
    return int(_libm.powf(1.5, _libm.ceilf(float(q))))


class Arena:
    def __init__(self, initial_verts: int = 1024, initial_indices: int = 1536):
        self.size = [initial_verts, initial_indices]                    # currentVertexSize, currentIndexSize
        self.free = [[[0, initial_verts]], [[0, initial_indices]]]      # [start, count] nodes, list order
        self.allocated = [[], []]
        self.streams = [np.zeros(initial_verts * w, np.uint32) for w in STRIDE_WORDS[:6]] + [np.zeros(initial_indices, np.uint32)]
        self.growths = 0

    # -- AllocateMesh helpers (src/RenderEngine_Allocate.cpp:19-80) --
    @staticmethod
    def _find_placement(requested, free):
        for node in free:
            if requested <= free[0][1]:  # sic: *bestRangeIndex is always begin()
                return node
        return None

    def _grow(self, fam: int, min_size: int):
        new_size = closest_power_of_1_5(min_size)
        old_size = self.size[fam]
        assert new_size > old_size
        ids = range(6) if fam == 0 else (6,)
        for s in ids:
            w = STRIDE_WORDS[s]
            fresh = np.zeros(new_size * w, np.uint32)
            off = 0
            for start, count in self.allocated[fam]:
                fresh[off * w:(off + count) * w] = self.streams[s][start * w:(start + count) * w]
                off += count
            self.streams[s] = fresh
        off = 0
        for node in self.allocated[fam]:
            node[0] = off
            off += node[1]
        free = self.free[fam]
        if not free:  # extendLastRange (:194-207)
            free.append([old_size, new_size - old_size])
        else:
            last = free[0]
            for node in free:
                if node[0] > last[0]:
                    last = node
            last[1] = (new_size - last[0]) & 0xFFFFFFFF
        self.size[fam] = new_size
        self.growths += 1

    def _location(self, fam: int, size: int):
        for _ in range(64):
            node = self._find_placement(size, self.free[fam])
            if node is not None:
                return node
            self._grow(fam, self.size[fam] + size)
        raise RuntimeError("placement loop does not terminate")

    def allocate(self, streams: list, indices):
        """streams: six arrays (or None) of n x 3 / n x 2 f32; indices: u32 array. Returns the [start, count] nodes."""
        n = next(len(s) for s in streams if s is not None) if any(s is not None for s in streams) else 0
        vnode = self._location(0, n)
        inode = self._location(1, len(indices))
        v = self._consume_identity(0, vnode, n)
        i = self._consume_identity(1, inode, len(indices))
        for s, arr in enumerate(streams):
            if arr is None or n == 0:
                continue
            w = STRIDE_WORDS[s]
            self.streams[s][v[0] * w:(v[0] + n) * w] = np.ascontiguousarray(arr, np.float32).reshape(-1).view(np.uint32)
        if len(indices):
            self.streams[6][i[0]:i[0] + len(indices)] = np.asarray(indices, np.uint32)
        return v, i

    def _consume_identity(self, fam, node, size):
        taken = [node[0], size]
        self.allocated[fam].append(taken)
        if node[1] == size:
            k = next(j for j, x in enumerate(self.free[fam]) if x is node)
            del self.free[fam][k]
        else:
            node[0] += size
            node[1] -= size
        return taken

    # -- DeallocateMesh (:101-145) --
    def _dealloc(self, fam: int, rng):
        start, count = rng
        found = [0, 0]
        for k, node in enumerate(self.allocated[fam]):
            if node[0] >= start and node[1] >= count:
                found = list(node)
                del self.allocated[fam][k]
                break
        overlap = False
        for f in self.free[fam]:
            if (f[0] + f[1]) & 0xFFFFFFFF == found[0]:
                f[1] = (f[1] + found[1]) & 0xFFFFFFFF
                overlap = True
            if (found[0] + found[1]) & 0xFFFFFFFF == f[0]:
                f[0] = (f[0] - found[1]) & 0xFFFFFFFF
                f[1] = (f[1] + found[1]) & 0xFFFFFFFF
                overlap = True
        if not overlap:
            self.free[fam].append([start, count])

    def deallocate(self, vnode, inode):
        self._dealloc(0, list(vnode))
        self._dealloc(1, list(inode))


# ---------------------------------------------------------------- .rvem ----------------------------------------------------------------
BITS = dict(skinned=1, positions=2, normals=4, tangents=8, bitangents=16, uv0=32, lightmap_uv=64)
ORDER = ("positions", "normals", "tangents", "bitangents", "uv0", "lightmap_uv")


def parse_rvem(blob: bytes) -> dict:
    """MeshAsset::DeserializeMeshFromMemory (src/MeshAsset.cpp:23-79) + the weight table of MeshAssetSkinned (:38-56)."""
    if len(blob) < 16 or blob[:4] != b"rvem":
        raise ValueError("Header does not match, data is not a mesh!")
    nv, ni = (int(x) for x in np.frombuffer(blob, "<u4", 2, 4))
    attrs = blob[12]
    at = 16
    out = dict(num_vertices=nv, num_indices=ni, attributes=attrs)
    for k, name in enumerate(ORDER):
        if attrs & (2 << k):
            w = STRIDE_WORDS[k]
            out[name] = np.frombuffer(blob, "<f4", nv * w, at).reshape(nv, w)
            at += nv * w * 4
        else:
            out[name] = None
    out["indices"] = np.frombuffer(blob, "<u4", ni, at)
    at += ni * 4
    out["mesh_bytes"] = at
    rest = len(blob) - at
    out["weights"] = None
    if rest >= 4:
        n = rest // 32
        out["weights"] = np.frombuffer(blob, np.dtype([("joint", "<u4"), ("influence", "<f4")]), n * 4, at).reshape(n, 4)
    return out
