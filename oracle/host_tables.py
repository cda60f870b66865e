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


# ---- skinned path (src/RenderEngine_Draw.cpp:546-574, 577-630, 1067-1218) --------------------------------------
@dataclass
class SkinnedMeshDesc:
    num_verts: int
    num_indices: int
    index_start: int
    vertex_start: int
    radius: float


@dataclass
class SkinnedCommandDesc:
    mesh: SkinnedMeshDesc
    num_bones: int
    entity_ids: np.ndarray


@dataclass
class SkinnedTables:
    cubos: list  # per material: CUBO_DTYPE[numCommands]
    dispatch_groups: list  # per material: ceil(LAST command's numObjects / 64)  (:1214)
    indirect_reset: list  # per material: (total_entities, 5) u32 staging table (:1081-1123)
    prepare_ubos: list  # per material: list of SkinningPrepareUBO field tuples, one per command (:552-570)
    skinning_ubos: list  # flat over all materials/commands: SkinningUBO field tuples (:597-625)
    bindless: list


def build_skinned_tables(materials) -> SkinnedTables:
    """materials: list[list[SkinnedCommandDesc]]. Handle numbering: per material culling, indirect; then per
    command its id list; one shared +inf lodDistance buffer per mesh object."""
    bindless: list = []

    def handle(arr):
        bindless.append(arr)
        return len(bindless) - 1

    cubos_all, groups, resets, prep_all, skin_all = [], [], [], [], []
    skeletal_vertex_offset = 0  # runs across materials (:1070)
    bone_read, vert_write = 0, 0  # SkinningUBO offsets run across all commands (:585, 623-624)
    lod_handles: dict[int, int] = {}
    for cmds in materials:
        total = sum(len(c.entity_ids) for c in cmds)
        reset = np.zeros((total, 5), np.uint32)
        for c in cmds:
            mesh_id = 0  # declared inside the per-command loop (:1088): every command starts at slot 0
            for i in range(len(c.entity_ids)):
                slot = mesh_id * 1 + 0 + i
                if slot < total:
                    reset[slot] = (c.mesh.num_indices, 0, c.mesh.index_start, skeletal_vertex_offset & 0xFFFFFFFF, i)
                skeletal_vertex_offset += c.mesh.num_verts
        culling = np.full(max(total, 1), 0xFFFFFFFF, np.uint32)
        h_cull, h_ind = handle(culling), handle(reset.copy())
        cubos = np.zeros(len(cmds), dtype=CUBO_DTYPE)
        ind_off = cull_off = 0
        last_n = 0
        prep = []
        vb_off = dc_off = bi_off = 0  # SkinningPrepareUBO is fresh per material (:551)
        for k, c in enumerate(cmds):
            n = len(c.entity_ids)
            key = id(c.mesh)
            if key not in lod_handles:
                lod_handles[key] = handle(np.array([np.inf], np.float32))
            ids = np.ascontiguousarray(c.entity_ids, dtype=np.uint32)
            cubos[k] = (ind_off, n, cull_off, np.float32(c.mesh.radius), 1, h_cull, handle(ids), h_ind, lod_handles[key])
            ind_off += 1
            cull_off += n
            last_n = n
            prep.append((c.mesh.index_start, vb_off, c.mesh.num_indices, c.mesh.num_verts, n, dc_off, bi_off))
            vb_off += c.mesh.num_verts
            dc_off += n
            bi_off += n
            skin_all.append((n, c.mesh.num_verts, c.num_bones, bone_read, vert_write, c.mesh.vertex_start))
            bone_read += c.num_bones * n
            vert_write += c.mesh.num_verts * n
        cubos_all.append(cubos)
        groups.append(int(np.ceil(np.float32(last_n) / np.float32(64.0))))
        resets.append(reset)
        prep_all.append(prep)
    return SkinnedTables(cubos_all, groups, resets, prep_all, skin_all, bindless)


def sync_ranges(tracking: np.ndarray, max_gap: int = 5) -> np.ndarray:
    """Copy ranges of BufferedVRAMStructureBase::EncodeSync (src/BufferedVRAMVector.cpp:106-139), restated loop for
    loop: dirty elements open / extend a range, `max_gap` consecutive clean ones close it at the index of the last
    of them. Returns (begin, end) pairs; clears the flags like the reference does."""
    out = []
    gap, begin = 0, 0
    n = len(tracking)
    for i in range(n):
        modified = bool(tracking[i])
        if modified and gap == 0:
            gap = max_gap
            begin = i
        elif modified:
            gap = max_gap
        if not modified:
            prev = gap
            gap = max(0, gap - 1)
            if gap == 0 and prev != 0:
                out.append((begin, i))
        if modified:
            tracking[i] = 0
    if gap != 0:
        out.append((begin, n))
    return np.array(out, np.uint32).reshape(-1, 2)


class SparseSetRef:
    """Restatement of UnorderedSparseSetGenericContainer + BufferedVRAMSparseSet's flag writes
    (include/RavEngine/SparseSet.hpp:22-47, include/RavEngine/BufferedVRAMVector.hpp:132-163) for index = value = u32:
    dense array with swap-remove erase, sparse -> dense map, reverse map, one modification flag per dense element."""

    INVALID = 0xFFFFFFFF

    def __init__(self):
        self.dense: list[int] = []
        self.sparse: list[int] = [self.INVALID]  # sparse_set{ default_index }
        self.reverse: list[int] = []
        self.flags: list[bool] = []  # syncTrackingBuffer

    def has(self, i: int) -> bool:
        return i < len(self.sparse) and self.sparse[i] != self.INVALID

    def emplace(self, i: int, value: int) -> None:
        if self.has(i):
            return
        self.dense.append(value)
        self.reverse.append(i)
        if i >= len(self.sparse):
            want = (i + 1 + 1) // 2 * 2  # closest_multiple_of<int>(i + 1, 2)
            self.sparse += [self.INVALID] * (want - len(self.sparse))
        self.sparse[i] = len(self.dense) - 1
        if len(self.dense) > len(self.flags):  # ResizeIfNeeded: new flags start true
            self.flags += [True] * (len(self.dense) - len(self.flags))
        self.flags[self.sparse[i]] = True

    def erase(self, i: int) -> None:
        assert self.has(i)
        d = self.sparse[i]
        self.flags[d] = True
        last = len(self.dense) - 1
        self.dense[d] = self.dense[last]  # unordered_vector::erase: swap with the last element, pop
        self.dense.pop()
        if d < len(self.dense):
            owner = self.reverse[-1]
            self.sparse[owner] = d
            self.reverse[d] = self.reverse[-1]
        self.reverse.pop()
        self.sparse[i] = self.INVALID
