"""Thin Python host layer over the C-ABI (ctypes). torch is used only for device memory,
streams and torch.distributed — every compute call goes to librve_b200.so.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass

import numpy as np

from . import _abi
from ._abi import (CullingUBO, CullingUBOinstance, GridAssignUBO, GridBuildUBO, SceneGenParams, SkinningPrepareUBO,
                   SkinningUBO)


class RveError(RuntimeError):
    pass


def _check(lib, rc: int) -> None:
    if rc != 0:
        raise RveError(f"rve error {rc}: {lib.rve_last_error_string().decode()}")


@dataclass
class Pyramid:
    handle: int
    dim: int
    levels: int
    texels: "Buffer"


class Buffer:
    """RGL::IBuffer stand-in: `handle` doubles as the bindless index a CUBO stores."""

    def __init__(self, ctx: "Context", handle: int, nbytes: int, keepalive=None):
        self.ctx, self.handle, self.nbytes, self._keepalive = ctx, handle, nbytes, keepalive

    @property
    def bindless(self) -> int:
        return self.handle

    @property
    def ptr(self) -> int:
        return self.ctx.lib.rve_buffer_device_ptr(self.ctx.h, self.handle)

    def upload(self, arr: np.ndarray, offset: int = 0) -> "Buffer":
        arr = np.ascontiguousarray(arr)
        _check(self.ctx.lib, self.ctx.lib.rve_buffer_upload(self.ctx.h, self.handle, offset, arr.ctypes.data, arr.nbytes))
        # pageable upload is staged before the call returns; pinned callers must keep `arr` alive
        return self

    def download(self, dtype=np.uint8, offset: int = 0, nbytes: int | None = None) -> np.ndarray:
        nbytes = self.nbytes - offset if nbytes is None else nbytes
        out = np.empty(nbytes // np.dtype(dtype).itemsize, dtype=dtype)
        _check(self.ctx.lib, self.ctx.lib.rve_buffer_download(self.ctx.h, self.handle, offset, out.ctypes.data, out.nbytes))
        return out

    def zero(self, offset: int = 0, nbytes: int | None = None) -> None:
        nbytes = self.nbytes - offset if nbytes is None else nbytes
        _check(self.ctx.lib, self.ctx.lib.rve_buffer_fill_zero(self.ctx.h, self.handle, offset, nbytes))

    def destroy(self) -> None:
        _check(self.ctx.lib, self.ctx.lib.rve_buffer_destroy(self.ctx.h, self.handle))
        self._keepalive = None


class Graph:
    """A captured frame (rve_graph): one launch replays every enqueue recorded by Context.capture."""

    def __init__(self, ctx: "Context", h):
        self.ctx, self.h = ctx, h

    def launch(self) -> None:
        _check(self.ctx.lib, self.ctx.lib.rve_graph_launch(self.ctx.h, self.h))

    def destroy(self) -> None:
        if self.h:
            self.ctx.lib.rve_graph_destroy(self.h)
            self.h = None


class Context:
    """One device + one in-order stream (the reference's device/queue pair)."""

    def __init__(self, device: int = 0, stream: int | None = None):
        self.lib = _abi.load_library()
        h = C.c_void_p()
        _check(self.lib, self.lib.rve_ctx_create(device, C.c_void_p(stream) if stream else None, C.byref(h)))
        self.h = h
        self.device = device

    def close(self) -> None:
        if self.h:
            self.lib.rve_ctx_destroy(self.h)
            self.h = None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    # ---- plumbing ----
    def synchronize(self) -> None:
        _check(self.lib, self.lib.rve_ctx_synchronize(self.h))

    @property
    def stream(self) -> int:
        return int(self.lib.rve_ctx_stream(self.h) or 0)

    @property
    def sm_count(self) -> int:
        return self.lib.rve_ctx_sm_count(self.h)

    @property
    def launch_count(self) -> int:
        return int(self.lib.rve_ctx_launch_count(self.h))

    def capture(self, record) -> "Graph":
        """Records the enqueues `record()` makes into a CUDA graph (rve_ctx_capture_begin / _end); replay with Graph.launch()."""
        _check(self.lib, self.lib.rve_ctx_capture_begin(self.h))
        try:
            record()
        finally:
            g = C.c_void_p()
            rc = self.lib.rve_ctx_capture_end(self.h, C.byref(g))
        _check(self.lib, rc)
        return Graph(self, g)

    def bind_host_to_device(self, cpus: bool = True, memory: bool = True) -> dict:
        """rve_ctx_bind_host_to_device: this thread onto the GPU's local CPUs, its next (pinned) allocations onto the GPU's NUMA node."""
        info = _abi.HostAffinity()
        _check(self.lib, self.lib.rve_ctx_bind_host_to_device(self.h, (1 if cpus else 0) | (2 if memory else 0), C.byref(info)))
        return {"numa_node": info.numa_node, "cpus_allowed": info.cpus_allowed, "cpus_local": info.cpus_local,
                "cpus_bound": bool(info.cpus_bound), "memory_bound": bool(info.memory_bound), "pci_bus_id": info.pci_bus_id.decode()}

    def set_profiling(self, on: bool) -> None:
        _check(self.lib, self.lib.rve_ctx_set_profiling(self.h, 1 if on else 0))

    def profile_read(self, kernel: int) -> tuple[float, int]:
        """(total ms, launches) of kernel 0 cull / 1 pyramid / 2 skinning / 3 cluster assign since the last read."""
        ms, n = C.c_double(), C.c_uint32()
        _check(self.lib, self.lib.rve_ctx_profile_read(self.h, kernel, C.byref(ms), C.byref(n)))
        return ms.value, n.value

    # ---- buffers ----
    def create_buffer(self, nbytes: int) -> Buffer:
        out = C.c_uint32()
        _check(self.lib, self.lib.rve_buffer_create(self.h, nbytes, C.byref(out)))
        return Buffer(self, out.value, nbytes)

    def buffer_from_numpy(self, arr: np.ndarray) -> Buffer:
        arr = np.ascontiguousarray(arr)
        return self.create_buffer(max(arr.nbytes, 16)).upload(arr)

    def wrap_tensor(self, t) -> Buffer:
        """Register a CUDA torch tensor's storage as a buffer (no copy)."""
        assert t.is_cuda and t.is_contiguous()
        nbytes = t.numel() * t.element_size()
        out = C.c_uint32()
        _check(self.lib, self.lib.rve_buffer_wrap(self.h, C.c_void_p(t.data_ptr()), nbytes, C.byref(out)))
        return Buffer(self, out.value, nbytes, keepalive=t)

    def copy(self, src: Buffer, dst: Buffer, nbytes: int, src_off: int = 0, dst_off: int = 0) -> None:
        _check(self.lib, self.lib.rve_buffer_copy(self.h, src.handle, src_off, dst.handle, dst_off, nbytes))

    # ---- stages ----
    def create_pyramid(self, width: int) -> Pyramid:
        out = C.c_uint32()
        _check(self.lib, self.lib.rve_depth_pyramid_create(self.h, width, C.byref(out)))
        dim, lv, tex = C.c_uint32(), C.c_uint32(), C.c_uint32()
        _check(self.lib, self.lib.rve_depth_pyramid_info(self.h, out.value, C.byref(dim), C.byref(lv), C.byref(tex)))
        nbytes = self.lib.rve_buffer_size(self.h, tex.value)
        return Pyramid(out.value, dim.value, lv.value, Buffer(self, tex.value, nbytes))

    def build_pyramid(self, pyr: Pyramid, depth: Buffer, width: int, height: int, deferred: bool = False) -> None:
        """rve_depth_pyramid_build; deferred=True: rve_depth_pyramid_build_deferred (the cull enqueued next finishes the upper levels)"""
        fn = self.lib.rve_depth_pyramid_build_deferred if deferred else self.lib.rve_depth_pyramid_build
        _check(self.lib, fn(self.h, pyr.handle, depth.handle, width, height))

    def pyramid_footmin(self, pyr: Pyramid) -> Buffer:
        """rve_depth_pyramid_footmin: the derived footprint-min chain the cull samples (inspection only)"""
        out = C.c_uint32()
        _check(self.lib, self.lib.rve_depth_pyramid_footmin(self.h, pyr.handle, C.byref(out)))
        return Buffer(self, out.value, self.lib.rve_buffer_size(self.h, out.value))

    def cull(self, ubo: CullingUBO, cubos: Buffer, model: Buffer, layers: Buffer, flags: Buffer, pyr: Pyramid,
             num_groups_x: int) -> None:
        _check(self.lib, self.lib.rve_cull(self.h, C.byref(ubo), cubos.handle, model.handle, layers.handle, flags.handle,
                                           pyr.handle, num_groups_x))

    def sort_entity_ids(self, ids: Buffer, count: int, id_limit: int, out: Buffer) -> None:
        """rve_sort_entity_ids: ascending copy of an entity-id list (the cull's result does not depend on list order)"""
        _check(self.lib, self.lib.rve_sort_entity_ids(self.h, ids.handle, count, id_limit, out.handle))

    def sort_entity_ids_batch(self, ids: list[Buffer], counts: list[int], id_limit: int, outs: list[Buffer]) -> None:
        """rve_sort_entity_ids_batch: ascending copies of several lists of one id space, two launches per 16 lists"""
        n = len(ids)
        a = (C.c_uint32 * n)(*[b.handle for b in ids])
        c = (C.c_uint32 * n)(*counts)
        o = (C.c_uint32 * n)(*[b.handle for b in outs])
        _check(self.lib, self.lib.rve_sort_entity_ids_batch(self.h, n, a, c, id_limit, o))

    def sort_entity_ids_static(self, ids: Buffer, count: int, id_limit: int, out: Buffer) -> None:
        """rve_sort_entity_ids_static: the same sort with its scratch in a fixed device array (no allocation; id_limit <=
        RVE_SORT_STATIC_MAX_IDS; calls that overlap on the GPU race)"""
        _check(self.lib, self.lib.rve_sort_entity_ids_static(self.h, ids.handle, count, id_limit, out.handle))

    def skin(self, ubo: SkinningUBO, outs: list[Buffer], ins: list[Buffer], pose: Buffer, weights: Buffer) -> None:
        o = (C.c_uint32 * 5)(*[b.handle for b in outs])
        i = (C.c_uint32 * 5)(*[b.handle for b in ins])
        _check(self.lib, self.lib.rve_skin(self.h, C.byref(ubo), o, i, pose.handle, weights.handle))

    def skinned_drawcall_prepare(self, ubo: SkinningPrepareUBO, indirect: Buffer) -> None:
        _check(self.lib, self.lib.rve_skinned_drawcall_prepare(self.h, C.byref(ubo), indirect.handle))

    def cluster_build(self, ubo: GridBuildUBO, clusters: Buffer) -> None:
        _check(self.lib, self.lib.rve_cluster_build(self.h, C.byref(ubo), clusters.handle))

    def cluster_assign(self, ubo: GridAssignUBO, clusters: Buffer, points: Buffer | None, spots: Buffer | None,
                       num_clusters: int) -> None:
        inv = 0xFFFFFFFF
        _check(self.lib, self.lib.rve_cluster_assign(self.h, C.byref(ubo), clusters.handle,
                                                     points.handle if points else inv, spots.handle if spots else inv,
                                                     num_clusters))

    def gen_instances(self, params: SceneGenParams, first_index: int, count: int, model: Buffer, layers: Buffer,
                      flags: Buffer, ids: Buffer) -> None:
        _check(self.lib, self.lib.rve_gen_instances(self.h, C.byref(params), first_index, count, model.handle,
                                                    layers.handle, flags.handle, ids.handle))

    def shard_offsets(self, counts_all: Buffer, world: int, slots: int, offsets: Buffer, totals: Buffer) -> None:
        _check(self.lib, self.lib.rve_shard_offsets(self.h, counts_all.handle, world, slots, offsets.handle, totals.handle))

    # ---- peer-memory count exchange (csrc/p2p.cu) ----
    def ipc_export(self, buf: Buffer) -> bytes:
        h = (C.c_uint8 * 64)()
        _check(self.lib, self.lib.rve_ipc_export(self.h, buf.handle, h))
        return bytes(h)

    def ipc_import(self, handle: bytes, nbytes: int) -> Buffer:
        raw = (C.c_uint8 * 64).from_buffer_copy(handle)
        out = C.c_uint32()
        _check(self.lib, self.lib.rve_ipc_import(self.h, raw, nbytes, C.byref(out)))
        return Buffer(self, out.value, nbytes)

    def mailbox_bytes(self, world: int, slots: int) -> int:
        return int(self.lib.rve_count_mailbox_bytes(world, slots))

    def exchange_counts_p2p(self, dense_counts: Buffer, slots: int, mailboxes: list[Buffer], rank: int, world: int, epoch: int,
                            offsets: Buffer, totals: Buffer, status: Buffer) -> None:
        m = (C.c_uint32 * world)(*[b.handle for b in mailboxes])
        _check(self.lib, self.lib.rve_exchange_counts_p2p(self.h, None, None, None, 0, dense_counts.handle, slots, m, rank, world, epoch,
                                                          offsets.handle, totals.handle, status.handle))

    def exchange_counts_p2p_indirect(self, indirect: list[Buffer], first_slots: list[int], slots: list[int], mailboxes: list[Buffer],
                                     rank: int, world: int, epoch: int, offsets: Buffer, totals: Buffer, status: Buffer) -> None:
        """Same exchange, the kernel reading the instanceCount fields straight out of the indirect buffers."""
        n = len(indirect)
        h = (C.c_uint32 * n)(*[b.handle for b in indirect])
        f = (C.c_uint32 * n)(*first_slots)
        c = (C.c_uint32 * n)(*slots)
        m = (C.c_uint32 * world)(*[b.handle for b in mailboxes])
        _check(self.lib, self.lib.rve_exchange_counts_p2p(self.h, h, f, c, n, 0xFFFFFFFF, 0, m, rank, world, epoch, offsets.handle,
                                                          totals.handle, status.handle))

    def gather_ids_p2p(self, src: list[Buffer], plan, dst: list[list], consumers: int, id_base: int, offsets: Buffer, totals: Buffer,
                       mailboxes: list[Buffer], rank: int, world: int, epoch: int, status: Buffer) -> None:
        """rve_gather_ids_p2p: `plan` is a sharding.GatherPlan, dst[q][b] rank q's global culling buffer b (None for non-consumers)."""
        nb, ns = len(src), len(plan.slot_buffer)
        sh = (C.c_uint32 * nb)(*[x.handle for x in src])
        inv = 0xFFFFFFFF
        dh = (C.c_uint32 * (world * nb))(*[(dst[q][b].handle if dst[q] and dst[q][b] is not None else inv) for q in range(world) for b in range(nb)])
        sb = (C.c_uint32 * ns)(*plan.slot_buffer)
        ss = (C.c_uint32 * ns)(*plan.src_start)
        ds = (C.c_uint32 * ns)(*plan.dst_start)
        m = (C.c_uint32 * world)(*[x.handle for x in mailboxes])
        _check(self.lib, self.lib.rve_gather_ids_p2p(self.h, sh, nb, sb, ss, ds, ns, dh, consumers, id_base, offsets.handle, totals.handle, m,
                                                     rank, world, epoch, status.handle))

    # ---- upload forwarding between ranks (csrc/sync.cu) ----
    def host_import(self, pid: int, fd: int, nbytes: int) -> int:
        """rve_host_import: maps and pins another rank's shared host block; returns the host pointer"""
        p = C.c_void_p()
        _check(self.lib, self.lib.rve_host_import(self.h, pid, fd, nbytes, C.byref(p)))
        return p.value

    def host_release(self, ptr: int, nbytes: int) -> None:
        _check(self.lib, self.lib.rve_host_release(self.h, C.c_void_p(ptr), nbytes))

    def flag_signal(self, flags: Buffer, word: int, value: int) -> None:
        _check(self.lib, self.lib.rve_flag_signal(self.h, flags.handle, word, value & 0xFFFFFFFF))

    def flag_wait(self, flags: Buffer, word: int, value: int, status: Buffer) -> None:
        _check(self.lib, self.lib.rve_flag_wait(self.h, flags.handle, word, value & 0xFFFFFFFF, status.handle))

    def forward_rows(self, host_ptr: int, row_bytes: int, affine: bool, first_row: int, rows: int, dst: Buffer) -> None:
        """rvs_forward_rows: rows of an imported host block -> the same rows of `dst` (a peer's private buffer)"""
        _check(self.lib, self.lib.rvs_forward_rows(self.h, C.c_void_p(host_ptr), row_bytes, 1 if affine else 0, first_row, rows, dst.handle))

    def wrap_pointer(self, ptr: int, nbytes: int, keepalive=None) -> Buffer:
        """Register device memory owned elsewhere (another context of the same device) as a buffer of this context."""
        out = C.c_uint32()
        _check(self.lib, self.lib.rve_buffer_wrap(self.h, C.c_void_p(ptr), nbytes, C.byref(out)))
        return Buffer(self, out.value, nbytes, keepalive=keepalive)

    def gather_instance_counts(self, indirect: Buffer, first_slot: int, slots: int, out: Buffer, out_off: int = 0) -> None:
        _check(self.lib, self.lib.rve_gather_instance_counts(self.h, indirect.handle, first_slot, slots, out.handle, out_off))

    def gather_instance_counts_multi(self, indirect: list[Buffer], first_slots: list[int], slots: list[int], out: Buffer, out_off: int = 0) -> None:
        n = len(indirect)
        h = (C.c_uint32 * n)(*[b.handle for b in indirect])
        f = (C.c_uint32 * n)(*first_slots)
        s = (C.c_uint32 * n)(*slots)
        _check(self.lib, self.lib.rve_gather_instance_counts_multi(self.h, h, f, s, n, out.handle, out_off))


def cull_views(ctx: "Context", ubos: list, pyramids: list, strides, cubos: Buffer, model: Buffer, layers: Buffer, flags: Buffer,
               num_groups_x: int) -> None:
    """rve_cull_views: one launch, every instance read once and tested against all views. `strides` = (numCubos, 2)
    uint32 array of (indirect, culling) strides between consecutive views."""
    n = len(ubos)
    arr = (CullingUBO * n)(*ubos)
    ph = (C.c_uint32 * n)(*[p.handle for p in pyramids])
    st = np.ascontiguousarray(strides, dtype=np.uint32)
    _check(ctx.lib, ctx.lib.rve_cull_views(ctx.h, arr, n, ph, st.ctypes.data, cubos.handle, model.handle, layers.handle, flags.handle,
                                           num_groups_x))


def cull_views_matrices(ctx: "Context", ubos: list, matrices: Buffer, matrix_index, pyramids: list, strides, cubos: Buffer, model: Buffer,
                        layers: Buffer, flags: Buffer, num_groups_x: int) -> None:
    """rve_cull_views_matrices: as cull_views, the view-projection of view v read from `matrices` (device) at mat4 index
    matrix_index[v] - e.g. the lightViewProj output of csm_cascades."""
    n = len(ubos)
    arr = (CullingUBO * n)(*ubos)
    ph = (C.c_uint32 * n)(*[p.handle for p in pyramids])
    ix = (C.c_uint32 * n)(*[int(i) for i in matrix_index])
    st = np.ascontiguousarray(strides, dtype=np.uint32)
    _check(ctx.lib, ctx.lib.rve_cull_views_matrices(ctx.h, arr, n, matrices.handle, ix, ph, st.ctypes.data, cubos.handle, model.handle,
                                                    layers.handle, flags.handle, num_groups_x))


# rve_b200.h RveCsmCamera / RveCsmLight as numpy record types (host staging of rve_csm_cascades)
CSM_CAMERA = np.dtype([("viewOnly", "<f4", 16), ("zNear", "<f4"), ("zFar", "<f4"), ("tanHalfFovy", "<f4"), ("aspect", "<f4")])
CSM_LIGHT = np.dtype([("direction", "<f4", 3), ("castsShadows", "<u4"), ("numCascades", "<u4"), ("shadowCascades", "<f4", 4),
                      ("maxShadowDistanceFactor", "<f4")])


def csm_camera_params(fov_degrees: float, width: int, height: int) -> tuple[float, float]:
    """rve_csm_camera_params: (tanHalfFovy, aspect) as src/RenderEngine_Draw.cpp:220 derives them (host, libm tanf)."""
    from ._abi import load_library
    t, a = C.c_float(), C.c_float()
    load_library().rve_csm_camera_params(C.c_float(fov_degrees), width, height, C.byref(t), C.byref(a))
    return t.value, a.value


def csm_cascades(ctx: "Context", cameras: Buffer, n_cameras: int, lights: Buffer, n_lights: int, view_proj: Buffer,
                 view: Buffer | None = None, proj: Buffer | None = None, distances: Buffer | None = None) -> None:
    """rve_csm_cascades: the cascade fit of src/RenderEngine_Draw.cpp:149-293 on the device."""
    inv = 0xFFFFFFFF
    _check(ctx.lib, ctx.lib.rve_csm_cascades(ctx.h, cameras.handle, n_cameras, lights.handle, n_lights, view_proj.handle,
                                             view.handle if view is not None else inv, proj.handle if proj is not None else inv,
                                             distances.handle if distances is not None else inv))
