"""Oracle restatement (numpy, integer/byte work) of the host table construction that feeds the
cull shader — src/RenderEngine_Draw.cpp:1223-1373 (static) and :1081-1187 (skinned).
TEST INFRASTRUCTURE ONLY.
"""
from __future__ import annotations

from dataclasses import dataclass

import numpy as np

CUBO_DTYPE = np.dtype([("indirectBufferOffset", "<u4"), ("numObjects", "<u4"), ("cullingBufferOffset", "<u4"),
                       ("radius", "<f4"), ("numLODs", "<u4"), ("idOutputBufferBindlessHandle", "<u4"),
                       ("entityIDInputBufferBindlessHandle", "<u4"), ("indirectOutputBufferBindlessHandle", "<u4"),
                       ("lodDistanceBufferBindlessHandle", "<u4")])
assert CUBO_DTYPE.itemsize == 36


@dataclass
class StaticTables:
    cubos: np.ndarray  # CUBO_DTYPE[numCubos] (entries past the written ones are zero)
    num_cubos: int  # CullingUBO.numCubos = sum of LOD counts (:1238), NOT the number of entries written
    total_entities: int
    indirect_init: list[np.ndarray]  # per material: (numLODs, 5) u32
    culling_slots: list[int]  # per material: numEntities * numLODs
    bindless: list  # handle -> host array


def build_static_tables(materials, quirk_slot: bool = False) -> StaticTables:
    """materials: list[scenes.MaterialDraws]. Allocates host arrays for every buffer and numbers the
    bindless handles in creation order (entity-id lists, lodDistances, then per material culling+indirect)."""
    bindless: list = []

    def handle(arr):
        bindless.append(arr)
        return len(bindless) - 1

    num_cubos = 0
    total = 0
    for m in materials:
        for c in m.commands:
            num_cubos += len(c.mesh.lods)
            total += len(c.entity_ids)
    cubos = np.zeros(num_cubos, dtype=CUBO_DTYPE)
    mesh_handles: dict[int, int] = {}
    cubo_idx = 0
    indirect_init, culling_slots = [], []
    for m in materials:
        num_lods = sum(len(c.mesh.lods) for c in m.commands)
        num_entities = sum(len(c.entity_ids) for c in m.commands)
        culling = np.full(max(num_entities * num_lods, 1), 0xFFFFFFFF, dtype=np.uint32)
        init = np.zeros((num_lods, 5), dtype=np.uint32)
        mesh_id = 0
        base_instance = 0
        lod_base = 0
        for c in m.commands:
            n = len(c.entity_ids)
            for lod_id, lm in enumerate(c.mesh.lods):
                slot = (mesh_id + lod_id) if quirk_slot else (lod_base + lod_id)
                init[slot] = (lm.index_count, 0, lm.index_start, lm.base_vertex, base_instance)
                base_instance += n
            lod_base += len(c.mesh.lods)
            mesh_id += 1
        indirect = init.copy()
        h_cull, h_ind = handle(culling), handle(indirect)
        ind_off = 0
        cull_off = 0
        for c in m.commands:
            lods = len(c.mesh.lods)
            key = id(c.mesh)
            if key not in mesh_handles:
                mesh_handles[key] = handle(np.asarray(c.mesh.lod_distances, dtype=np.float32))
            ids = np.ascontiguousarray(c.entity_ids, dtype=np.uint32)
            cubos[cubo_idx] = (ind_off, len(ids), cull_off, np.float32(c.mesh.radius), lods, h_cull, handle(ids), h_ind,
                               mesh_handles[key])
            cubo_idx += 1
            ind_off += lods
            cull_off += lods * len(ids)
        indirect_init.append(init)
        culling_slots.append(num_entities * num_lods)
    return StaticTables(cubos, num_cubos, total, indirect_init, culling_slots, bindless)


def sorted_slices(indirect: np.ndarray, culling: np.ndarray, cubos: np.ndarray, material_cubos: range) -> list[np.ndarray]:
    """Visible-id slices of one material, each sorted (append order is arrival order in the reference)."""
    out = []
    ind = indirect.reshape(-1, 5)
    for ci in material_cubos:
        cb = cubos[ci]
        for lod in range(int(cb["numLODs"])):
            cnt = int(ind[int(cb["indirectBufferOffset"]) + lod, 1])
            start = int(cb["numObjects"]) * lod + int(cb["cullingBufferOffset"])
            out.append(np.sort(culling[start:start + cnt]))
    return out
