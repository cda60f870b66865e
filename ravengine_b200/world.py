"""Python handle on the C++ shim (include/rve_shim.h): the World::RenderData stand-in plus the
encode blocks of RenderEngine::Draw for this path. All logic lives in the C++ shim; this only
marshals arguments."""
from __future__ import annotations

import ctypes as C

import numpy as np

from .api import Buffer, Context, Pyramid, _check

MAT_LIT_OPAQUE, MAT_UNLIT_OPAQUE, MAT_LIT_TRANSPARENT, MAT_UNLIT_TRANSPARENT, MAT_ALL = 1, 2, 4, 8, 0xF
QUIRK_INDIRECT_SLOT = 1
SORT_ENTITY_LISTS = 2


def _fp(a):
    a = np.ascontiguousarray(a, dtype=np.float32)
    return a, a.ctypes.data_as(C.POINTER(C.c_float))


def _up(a):
    a = np.ascontiguousarray(a, dtype=np.uint32)
    return a, a.ctypes.data_as(C.POINTER(C.c_uint32))


class World:
    def __init__(self, ctx: Context, flags: int = 0):
        self.ctx, self.lib = ctx, ctx.lib
        h = C.c_void_p()
        _check(self.lib, self.lib.rvs_world_create(ctx.h, flags, C.byref(h)))
        self.h = h
        self._keep = []  # host palettes the shim reads lazily

    def destroy(self):
        if self.h:
            self.lib.rvs_world_destroy(self.h)
            self.h = None

    def _buf(self, handle: int) -> Buffer | None:
        if handle == 0xFFFFFFFF:
            return None
        return Buffer(self.ctx, handle, self.lib.rve_buffer_size(self.ctx.h, handle))

    def set_entity_buffers(self, model: Buffer, layers: Buffer, flags: Buffer):
        _check(self.lib, self.lib.rvs_world_set_entity_buffers(self.h, model.handle, layers.handle, flags.handle))

    def mesh_collection(self, radius, lod_distances, index_counts, index_starts, base_vertices) -> int:
        out = C.c_uint32()
        k1, d = _fp(lod_distances)
        k2, a = _up(index_counts)
        k3, b = _up(index_starts)
        k4, c = _up(base_vertices)
        _check(self.lib, self.lib.rvs_mesh_collection_create(self.h, radius, len(k1), d, a, b, c, C.byref(out)))
        return out.value

    def material(self, kind: int = MAT_LIT_OPAQUE) -> int:
        out = C.c_uint32()
        _check(self.lib, self.lib.rvs_material_create(self.h, kind, C.byref(out)))
        return out.value

    def add_command(self, material: int, mesh: int, entity_ids: Buffer, count: int):
        _check(self.lib, self.lib.rvs_material_add_command(self.h, material, mesh, entity_ids.handle, count))

    def add_command_set(self, material: int, mesh: int, entities) -> None:
        """rvs_material_add_command_set: the command's entity list is a vram_vector.SparseSet, synced before every cull"""
        _check(self.lib, self.lib.rvs_material_add_command_set(self.h, material, mesh, entities.h))
        self._keep.append(entities)

    def set_command_entities(self, material: int, command: int, entity_ids: Buffer, count: int):
        _check(self.lib, self.lib.rvs_material_set_command_entities(self.h, material, command, entity_ids.handle, count))

    def cull_static(self, view_proj, cam_pos, pyramid: Pyramid, layers=0xFFFFFFFF, shadow=False, mask=MAT_ALL):
        k1, vp = _fp(view_proj)
        k2, cp = _fp(cam_pos)
        _check(self.lib, self.lib.rvs_cull_static(self.h, vp, cp, layers, 1 if shadow else 0, mask, pyramid.handle))

    def cull_static_with_pyramid(self, view_proj, cam_pos, pyramid: Pyramid, depth: Buffer, width: int, height: int, layers=0xFFFFFFFF,
                                 shadow=False, mask=MAT_ALL):
        """rvs_cull_static_with_pyramid: pyramid build + cull as one encode (tables first, cull overlapping the pyramid's tail)."""
        k1, vp = _fp(view_proj)
        k2, cp = _fp(cam_pos)
        _check(self.lib, self.lib.rvs_cull_static_with_pyramid(self.h, vp, cp, layers, 1 if shadow else 0, mask, pyramid.handle, depth.handle,
                                                               width, height))

    def cull_static_views(self, views, pyramids, mask=MAT_ALL):
        """views: list of (view_proj_cm[16], cam_pos[3], camera_layers, shadow) - all of them in ONE launch
        (rvs_cull_static_views); outputs hold len(views) consecutive copies of the single-view layout."""
        from ._abi import ShimView

        n = len(views)
        arr = (ShimView * n)()
        for a, (vp, cp, layers, shadow) in zip(arr, views):
            a.view_proj[:] = [float(x) for x in vp]
            a.cam_pos[:] = [float(x) for x in cp]
            a.camera_layers = layers
            a.shadow_pass = 1 if shadow else 0
        ph = (C.c_uint32 * n)(*[p.handle for p in pyramids])
        _check(self.lib, self.lib.rvs_cull_static_views(self.h, n, arr, ph, mask))

    def frame_encode(self, cameras, shadow_views=(), shadow_pyramids=(), points: Buffer | None = None, n_points: int = 0,
                     spots: Buffer | None = None, n_spots: int = 0, grid=(12, 12, 24), mask=MAT_ALL):
        """rvs_frame_encode. cameras: dicts(view_proj, view, inv_proj, cam_pos, z_near, z_far, layers, pyramid, depth (Buffer | None),
        depth_wh, screen); shadow_views: (view_proj_cm, cam_pos, layers, shadow) tuples as in cull_static_views."""
        from ._abi import FrameCamera, FrameDesc, ShimView

        inv = 0xFFFFFFFF
        cams = (FrameCamera * max(len(cameras), 1))()
        for c, d in zip(cams, cameras):
            c.view_proj[:] = [float(x) for x in d["view_proj"]]
            c.view[:] = [float(x) for x in d["view"]]
            c.inv_proj[:] = [float(x) for x in d["inv_proj"]]
            c.cam_pos[:] = [float(x) for x in d["cam_pos"]]
            c.z_near, c.z_far, c.layers = d["z_near"], d["z_far"], d.get("layers", 0xFFFFFFFF)
            c.pyramid = d["pyramid"].handle
            c.depth = d["depth"].handle if d.get("depth") is not None else inv
            c.depth_width, c.depth_height = d.get("depth_wh", (0, 0))
            c.screen_dim[:] = d["screen"]
        n = len(shadow_views)
        sv = (ShimView * max(n, 1))()
        for s, (vp, cp, layers, shadow) in zip(sv, shadow_views):
            s.view_proj[:] = [float(x) for x in vp]
            s.cam_pos[:] = [float(x) for x in cp]
            s.camera_layers, s.shadow_pass = layers, 1 if shadow else 0
        sp = (C.c_uint32 * max(n, 1))(*[p.handle for p in shadow_pyramids])
        f = FrameDesc(cams, len(cameras), sv, sp, n, points.handle if points else inv, n_points, spots.handle if spots else inv, n_spots,
                      (C.c_uint32 * 3)(*grid), mask)
        _check(self.lib, self.lib.rvs_frame_encode(self.h, C.byref(f)))

    def frame_cluster_buffer(self, camera: int) -> Buffer:
        out = C.c_uint32()
        _check(self.lib, self.lib.rvs_frame_cluster_buffer(self.h, camera, C.byref(out)))
        return self._buf(out.value)

    def material_outputs(self, material: int):
        c, i, l, e = C.c_uint32(), C.c_uint32(), C.c_uint32(), C.c_uint32()
        _check(self.lib, self.lib.rvs_material_outputs(self.h, material, C.byref(c), C.byref(i), C.byref(l), C.byref(e)))
        return self._buf(c.value), self._buf(i.value), l.value, e.value

    def last_cull_info(self):
        b, n, t = C.c_uint32(), C.c_uint32(), C.c_uint32()
        _check(self.lib, self.lib.rvs_last_cull_info(self.h, C.byref(b), C.byref(n), C.byref(t)))
        return self._buf(b.value), n.value, t.value

    def sorted_list_stats(self):
        """rvs_sorted_list_stats: (full rebuilds, delta updates) of the ascending list copies since the world was created"""
        f, d = C.c_uint32(), C.c_uint32()
        _check(self.lib, self.lib.rvs_sorted_list_stats(self.h, C.byref(f), C.byref(d)))
        return f.value, d.value

    # ---- skinned ----
    def skinned_mesh(self, num_verts, num_indices, index_start, vertex_start, radius, weights: Buffer) -> int:
        out = C.c_uint32()
        _check(self.lib, self.lib.rvs_skinned_mesh_create(self.h, num_verts, num_indices, index_start, vertex_start, radius,
                                                          weights.handle, C.byref(out)))
        return out.value

    def skinned_material(self, kind: int = MAT_LIT_OPAQUE) -> int:
        out = C.c_uint32()
        _check(self.lib, self.lib.rvs_skinned_material_create(self.h, kind, C.byref(out)))
        return out.value

    def add_skinned_command(self, material, mesh, num_bones, entity_ids: Buffer, count, palettes: np.ndarray):
        palettes = np.ascontiguousarray(palettes, dtype=np.float32)
        self._keep.append(palettes)
        _check(self.lib, self.lib.rvs_skinned_material_add_command(self.h, material, mesh, num_bones, entity_ids.handle, count,
                                                                   C.c_void_p(palettes.ctypes.data)))

    def set_mesh_arena(self, streams: list[Buffer]):
        arr = (C.c_uint32 * 5)(*[b.handle for b in streams])
        _check(self.lib, self.lib.rvs_world_set_mesh_arena(self.h, arr))

    def prepare_skeletal_mesh_buffers(self) -> bool:
        e = C.c_uint32()
        _check(self.lib, self.lib.rvs_prepare_skeletal_mesh_buffers(self.h, C.byref(e)))
        return bool(e.value)

    def pose_skeletal_meshes(self):
        _check(self.lib, self.lib.rvs_pose_skeletal_meshes(self.h))

    def prepare_skeletal_culling_buffer(self):
        _check(self.lib, self.lib.rvs_prepare_skeletal_culling_buffer(self.h))

    def cull_skeletal(self, view_proj, cam_pos, pyramid: Pyramid, layers=0xFFFFFFFF, shadow=False, mask=MAT_ALL):
        k1, vp = _fp(view_proj)
        k2, cp = _fp(cam_pos)
        _check(self.lib, self.lib.rvs_cull_skeletal(self.h, vp, cp, layers, 1 if shadow else 0, mask, pyramid.handle))

    def skinned_outputs(self):
        arr = (C.c_uint32 * 5)()
        p = C.c_uint32()
        _check(self.lib, self.lib.rvs_skinned_outputs(self.h, arr, C.byref(p)))
        return [self._buf(h) for h in arr], self._buf(p.value)

    def skinned_material_outputs(self, material: int):
        c, i, e = C.c_uint32(), C.c_uint32(), C.c_uint32()
        _check(self.lib, self.lib.rvs_skinned_material_outputs(self.h, material, C.byref(c), C.byref(i), C.byref(e)))
        return self._buf(c.value), self._buf(i.value), e.value

    def bin_lights(self, inv_proj, view_mat, grid, screen, z_near, z_far, points: Buffer | None, n_points: int,
                   spots: Buffer | None, n_spots: int) -> Buffer:
        k1, ip = _fp(inv_proj)
        k2, vm = _fp(view_mat)
        k3, g = _up(grid)
        k4, s = _up(screen)
        out = C.c_uint32()
        inv = 0xFFFFFFFF
        _check(self.lib, self.lib.rvs_bin_lights(self.h, ip, vm, g, s, z_near, z_far, points.handle if points else inv, n_points,
                                                 spots.handle if spots else inv, n_spots, C.byref(out)))
        return self._buf(out.value)
