"""ctypes mirror of include/rve_b200.h (struct layouts + function prototypes).

The library is loaded from the in-tree build (ravengine_b200/librve_b200.so). There is no
CPU fallback: if the library is missing or no CUDA device exists, calls fail loudly.
"""
from __future__ import annotations

import ctypes as C
from pathlib import Path

import os

# RVE_B200_LIB selects a tuning variant built with `python -m ravengine_b200.build -D... path.so` (development only)
LIB_PATH = Path(os.environ.get("RVE_B200_LIB") or Path(__file__).resolve().parent / "librve_b200.so")

u32, u64, f32, i32 = C.c_uint32, C.c_uint64, C.c_float, C.c_int32


class IndirectIndexedCommand(C.Structure):
    _pack_ = 1
    _fields_ = [("indexCount", u32), ("instanceCount", u32), ("indexStart", u32), ("baseVertex", u32), ("baseInstance", u32)]


class CullingUBO(C.Structure):
    _pack_ = 1
    _fields_ = [("viewProj", f32 * 16), ("camPos", f32 * 3), ("numCubos", u32), ("cameraRenderLayers", u32),
                ("singleInstanceModeAndShadowMode", u32)]


class CullingUBOinstance(C.Structure):
    _pack_ = 1
    _fields_ = [("indirectBufferOffset", u32), ("numObjects", u32), ("cullingBufferOffset", u32), ("radius", f32),
                ("numLODs", u32), ("idOutputBufferBindlessHandle", u32), ("entityIDInputBufferBindlessHandle", u32),
                ("indirectOutputBufferBindlessHandle", u32), ("lodDistanceBufferBindlessHandle", u32)]


class SkinningUBO(C.Structure):
    _pack_ = 1
    _fields_ = [("numObjects", u32), ("numVertices", u32), ("numBones", u32), ("boneReadOffset", u32),
                ("vertexWriteOffset", u32), ("vertexReadOffset", u32)]


class SkinningPrepareUBO(C.Structure):
    _pack_ = 1
    _fields_ = [("indexBufferOffset", u32), ("vertexBufferOffset", u32), ("nIndicesInThisMesh", u32),
                ("nVerticesInThisMesh", u32), ("nTotalObjects", u32), ("drawCallBufferOffset", u32),
                ("baseInstanceOffset", u32)]


class GridBuildUBO(C.Structure):
    _pack_ = 1
    _fields_ = [("invProj", f32 * 16), ("gridSize", u32 * 3), ("zNear", f32), ("screenDim", u32 * 2), ("zFar", f32)]


class GridAssignUBO(C.Structure):
    _pack_ = 1
    _fields_ = [("viewMat", f32 * 16), ("pointLightCount", u32), ("spotLightCount", u32)]


class SceneGenParams(C.Structure):
    _fields_ = [("seed", u64), ("extent_xy", f32), ("extent_z", f32), ("scale_min", f32), ("scale_max", f32),
                ("frac_layer_zero", f32), ("frac_no_cull_bits", f32), ("id_base", u32)]


class HostAffinity(C.Structure):  # rve_host_affinity
    _fields_ = [("numa_node", i32), ("cpus_allowed", u32), ("cpus_local", u32), ("cpus_bound", u32), ("memory_bound", u32),
                ("pci_bus_id", C.c_char * 32)]


assert C.sizeof(IndirectIndexedCommand) == 20
assert C.sizeof(CullingUBO) == 88
assert C.sizeof(CullingUBOinstance) == 36
assert C.sizeof(SkinningUBO) == 24
assert C.sizeof(SkinningPrepareUBO) == 28
assert C.sizeof(GridBuildUBO) == 92
assert C.sizeof(GridAssignUBO) == 72

CLUSTER_BYTES = 160
POINT_LIGHT_BYTES = 512
SPOT_LIGHT_BYTES = 168

# every symbol include/rve_b200.h declares: name -> (restype, argtypes)
VP = C.c_void_p
PROTOTYPES = {
    "rve_ctx_create": (C.c_int, [C.c_int, VP, C.POINTER(VP)]),
    "rve_ctx_destroy": (C.c_int, [VP]),
    "rve_ctx_synchronize": (C.c_int, [VP]),
    "rve_ctx_stream": (VP, [VP]),
    "rve_last_error_string": (C.c_char_p, []),
    "rve_version": (C.c_char_p, []),
    "rve_ctx_sm_count": (C.c_int, [VP]),
    "rve_ctx_launch_count": (u64, [VP]),
    "rve_ctx_bad_cubo_handles": (u32, [VP]),
    "rve_ctx_capture_begin": (C.c_int, [VP]),
    "rve_ctx_capture_end": (C.c_int, [VP, C.POINTER(VP)]),
    "rve_graph_launch": (C.c_int, [VP, VP]),
    "rve_graph_destroy": (C.c_int, [VP]),
    "rve_ctx_bind_host_to_device": (C.c_int, [VP, u32, VP]),
    "rve_ctx_set_profiling": (C.c_int, [VP, C.c_int]),
    "rve_ctx_profile_read": (C.c_int, [VP, C.c_int, C.POINTER(C.c_double), C.POINTER(u32)]),
    "rve_buffer_create": (C.c_int, [VP, C.c_size_t, C.POINTER(u32)]),
    "rve_buffer_wrap": (C.c_int, [VP, VP, C.c_size_t, C.POINTER(u32)]),
    "rve_buffer_destroy": (C.c_int, [VP, u32]),
    "rve_buffer_upload": (C.c_int, [VP, u32, C.c_size_t, VP, C.c_size_t]),
    "rve_buffer_upload_batch": (C.c_int, [VP, VP, u32]),
    "rve_buffer_download": (C.c_int, [VP, u32, C.c_size_t, VP, C.c_size_t]),
    "rve_buffer_copy": (C.c_int, [VP, u32, C.c_size_t, u32, C.c_size_t, C.c_size_t]),
    "rve_buffer_fill_zero": (C.c_int, [VP, u32, C.c_size_t, C.c_size_t]),
    "rve_buffer_device_ptr": (VP, [VP, u32]),
    "rve_buffer_size": (C.c_size_t, [VP, u32]),
    "rve_buffer_bindless_index": (u32, [VP, u32]),
    "rve_depth_pyramid_create": (C.c_int, [VP, u32, C.POINTER(u32)]),
    "rve_depth_pyramid_destroy": (C.c_int, [VP, u32]),
    "rve_depth_pyramid_info": (C.c_int, [VP, u32, C.POINTER(u32), C.POINTER(u32), C.POINTER(u32)]),
    "rve_depth_pyramid_footmin": (C.c_int, [VP, u32, C.POINTER(u32)]),
    "rve_depth_pyramid_build": (C.c_int, [VP, u32, u32, u32, u32]),
    "rve_depth_pyramid_build_deferred": (C.c_int, [VP, u32, u32, u32, u32]),
    "rve_depth_pyramid_texels_changed": (C.c_int, [VP, u32]),
    "rve_cull": (C.c_int, [VP, C.POINTER(CullingUBO), u32, u32, u32, u32, u32, u32]),
    "rve_sort_entity_ids": (C.c_int, [VP, u32, u32, u32, u32]),
    "rve_sort_entity_ids_batch": (C.c_int, [VP, u32, C.POINTER(u32), C.POINTER(u32), u32, C.POINTER(u32)]),
    "rve_sort_entity_ids_static": (C.c_int, [VP, u32, u32, u32, u32]),
    "rve_cull_views": (C.c_int, [VP, C.POINTER(CullingUBO), u32, C.POINTER(u32), VP, u32, u32, u32, u32, u32]),
    "rve_cull_views_matrices": (C.c_int, [VP, C.POINTER(CullingUBO), u32, u32, C.POINTER(u32), C.POINTER(u32), VP, u32, u32, u32, u32, u32]),
    "rve_skin": (C.c_int, [VP, C.POINTER(SkinningUBO), C.POINTER(u32), C.POINTER(u32), u32, u32]),
    "rve_skinned_drawcall_prepare": (C.c_int, [VP, C.POINTER(SkinningPrepareUBO), u32]),
    "rve_cluster_build": (C.c_int, [VP, C.POINTER(GridBuildUBO), u32]),
    "rve_cluster_assign": (C.c_int, [VP, C.POINTER(GridAssignUBO), u32, u32, u32, u32]),
    "rve_shard_offsets": (C.c_int, [VP, u32, u32, u32, u32, u32]),
    "rve_gather_instance_counts": (C.c_int, [VP, u32, u32, u32, u32, C.c_size_t]),
    "rve_gather_instance_counts_multi": (C.c_int, [VP, C.POINTER(u32), C.POINTER(u32), C.POINTER(u32), u32, u32, C.c_size_t]),
    "rve_count_mailbox_bytes": (C.c_size_t, [u32, u32]),
    "rve_ipc_export": (C.c_int, [VP, u32, VP]),
    "rve_ipc_import": (C.c_int, [VP, VP, C.c_size_t, C.POINTER(u32)]),
    "rve_exchange_counts_p2p": (C.c_int, [VP, C.POINTER(u32), C.POINTER(u32), C.POINTER(u32), u32, u32, u32, C.POINTER(u32), u32, u32, u32,
                                          u32, u32, u32]),
    "rve_gather_ids_p2p": (C.c_int, [VP, C.POINTER(u32), u32, C.POINTER(u32), C.POINTER(u32), C.POINTER(u32), u32, C.POINTER(u32), u32, u32,
                                     u32, u32, C.POINTER(u32), u32, u32, u32, u32]),
    "rve_scatter_rows": (C.c_int, [VP, u32, u32, u32, u64, u64, u32]),
    "rve_csm_camera_params": (None, [f32, u32, u32, C.POINTER(f32), C.POINTER(f32)]),
    "rve_csm_cascades": (C.c_int, [VP, u32, u32, u32, u32, u32, u32, u32, u32]),
    "rve_gen_instances": (C.c_int, [VP, C.POINTER(SceneGenParams), u64, u32, u32, u32, u32, u32]),
}

FP = C.POINTER(f32)
UP = C.POINTER(u32)
# include/rve_shim.h
PROTOTYPES.update({
    "rvs_world_create": (C.c_int, [VP, u32, C.POINTER(VP)]),
    "rvs_world_destroy": (C.c_int, [VP]),
    "rvs_world_set_entity_buffers": (C.c_int, [VP, u32, u32, u32]),
    "rvs_mesh_collection_create": (C.c_int, [VP, f32, u32, FP, UP, UP, UP, UP]),
    "rvs_material_create": (C.c_int, [VP, u32, UP]),
    "rvs_material_add_command": (C.c_int, [VP, u32, u32, u32, u32]),
    "rvs_material_set_command_entities": (C.c_int, [VP, u32, u32, u32, u32]),
    "rvs_cull_static": (C.c_int, [VP, FP, FP, u32, u32, u32, u32]),
    "rvs_cull_static_with_pyramid": (C.c_int, [VP, FP, FP, u32, u32, u32, u32, u32, u32, u32]),
    "rvs_cull_static_views": (C.c_int, [VP, u32, VP, UP, u32]),
    "rvs_material_outputs": (C.c_int, [VP, u32, UP, UP, UP, UP]),
    "rvs_last_cull_info": (C.c_int, [VP, UP, UP, UP]),
    "rvs_sorted_list_stats": (C.c_int, [VP, UP, UP]),
    "rvs_skinned_mesh_create": (C.c_int, [VP, u32, u32, u32, u32, f32, u32, UP]),
    "rvs_skinned_material_create": (C.c_int, [VP, u32, UP]),
    "rvs_skinned_material_add_command": (C.c_int, [VP, u32, u32, u32, u32, u32, VP]),
    "rvs_world_set_mesh_arena": (C.c_int, [VP, UP]),
    "rvs_prepare_skeletal_mesh_buffers": (C.c_int, [VP, UP]),
    "rvs_pose_skeletal_meshes": (C.c_int, [VP]),
    "rvs_prepare_skeletal_culling_buffer": (C.c_int, [VP]),
    "rvs_cull_skeletal": (C.c_int, [VP, FP, FP, u32, u32, u32, u32]),
    "rvs_skinned_outputs": (C.c_int, [VP, UP, UP]),
    "rvs_skinned_material_outputs": (C.c_int, [VP, u32, UP, UP, UP]),
    "rvs_frame_encode": (C.c_int, [VP, VP]),
    "rvs_frame_cluster_buffer": (C.c_int, [VP, u32, UP]),
    "rvs_arena_create": (C.c_int, [VP, u32, u32, C.POINTER(VP)]),
    "rvs_arena_destroy": (C.c_int, [VP]),
    "rvs_arena_allocate_mesh": (C.c_int, [VP, VP, UP]),
    "rvs_arena_deallocate_mesh": (C.c_int, [VP, u32]),
    "rvs_arena_range": (C.c_int, [VP, u32, UP, UP, UP, UP]),
    "rvs_arena_buffers": (C.c_int, [VP, UP, UP, UP]),
    "rvs_arena_list": (u32, [VP, u32, UP, u32]),
    "rvs_rvem_parse": (C.c_int, [VP, u64, VP]),
    "rvs_arena_load_rvem": (C.c_int, [VP, VP, u64, u32, UP, UP]),
    "rvs_bin_lights": (C.c_int, [VP, FP, FP, UP, UP, f32, f32, u32, u32, u32, u32, UP]),
})



class ShimView(C.Structure):
    _fields_ = [("view_proj", f32 * 16), ("cam_pos", f32 * 3), ("camera_layers", u32), ("shadow_pass", u32)]


class FrameCamera(C.Structure):  # rvs_frame_camera
    _fields_ = [("view_proj", f32 * 16), ("view", f32 * 16), ("inv_proj", f32 * 16), ("cam_pos", f32 * 3), ("z_near", f32), ("z_far", f32),
                ("layers", u32), ("pyramid", u32), ("depth", u32), ("depth_width", u32), ("depth_height", u32), ("screen_dim", u32 * 2)]


class FrameDesc(C.Structure):  # rvs_frame_desc
    _fields_ = [("cameras", C.POINTER(FrameCamera)), ("num_cameras", u32), ("shadow_views", C.POINTER(ShimView)),
                ("shadow_pyramids", C.POINTER(u32)), ("num_shadow_views", u32), ("point_lights", u32), ("point_light_count", u32),
                ("spot_lights", u32), ("spot_light_count", u32), ("grid_size", u32 * 3), ("pass_mask", u32)]


class SyncStats(C.Structure):
    _fields_ = [("ranges", u32), ("rows_in_ranges", u32), ("direct_copies", u32), ("packed_rows", u32),
                ("direct_bytes", u64), ("packed_bytes", u64)]


SYNC_RANGES, SYNC_PACKED, SYNC_PULL, SYNC_FULL, SYNC_FULL_PULL = 0, 1, 2, 3, 4
VECTOR_AFFINE_ROWS = 1
VECTOR_SHARED_HOST = 2
PROTOTYPES.update({
    "rvs_vector_create": (C.c_int, [VP, u32, u32, C.POINTER(VP)]),
    "rvs_vector_create_ex": (C.c_int, [VP, u32, u32, u32, C.POINTER(VP)]),
    "rvs_vector_destroy": (C.c_int, [VP]),
    "rvs_vector_resize": (C.c_int, [VP, u32]),
    "rvs_vector_size": (u32, [VP]),
    "rvs_vector_host": (VP, [VP]),
    "rvs_vector_tracking": (VP, [VP]),
    "rvs_vector_set": (C.c_int, [VP, u32, VP]),
    "rvs_vector_mark": (C.c_int, [VP, u32, u32]),
    "rvs_vector_encode_sync": (C.c_int, [VP, u32, u32, C.POINTER(SyncStats)]),
    "rvs_vector_private_buffer": (u32, [VP]),
    "rvs_vector_encode_sync_rows": (C.c_int, [VP, u32, u32, u32, C.POINTER(SyncStats)]),
    "rvs_vectors_encode_sync_batch": (C.c_int, [C.POINTER(VP), u32, u32, C.POINTER(SyncStats)]),
    "rvs_vector_share": (C.c_int, [VP, C.POINTER(i32), C.POINTER(i32), C.POINTER(u64)]),
    "rve_host_import": (C.c_int, [VP, i32, i32, u64, C.POINTER(VP)]),
    "rve_host_release": (C.c_int, [VP, VP, u64]),
    "rve_flag_signal": (C.c_int, [VP, u32, u32, u32]),
    "rve_flag_wait": (C.c_int, [VP, u32, u32, u32, u32]),
    "rvs_forward_rows": (C.c_int, [VP, VP, u32, u32, u64, u64, u32]),
    "rvs_vector_pulled_rows": (u64, [VP]),
    "rvs_vector_last_ranges": (u32, [VP, UP, u32]),
    "rvs_sync_ranges_from_flags": (u32, [C.POINTER(C.c_uint8), u32, UP, u32]),
    "rvs_sync_ranges_from_indices": (u32, [UP, u32, u32, UP, u32]),
})

PROTOTYPES.update({
    "rvs_sparse_set_create": (C.c_int, [VP, C.POINTER(VP)]),
    "rvs_sparse_set_destroy": (C.c_int, [VP]),
    "rvs_sparse_set_emplace": (C.c_int, [VP, u32, u32]),
    "rvs_sparse_set_erase": (C.c_int, [VP, u32]),
    "rvs_sparse_set_emplace_many": (C.c_int, [VP, VP, VP, u32]),
    "rvs_sparse_set_erase_many": (C.c_int, [VP, VP, u32]),
    "rvs_sparse_set_has": (C.c_int, [VP, u32]),
    "rvs_sparse_set_dense_size": (u32, [VP]),
    "rvs_sparse_set_dense": (VP, [VP]),
    "rvs_sparse_set_reverse_map": (VP, [VP]),
    "rvs_sparse_set_vector": (VP, [VP]),
    "rvs_material_add_command_set": (C.c_int, [VP, u32, u32, VP]),
})

_lib = None


def load_library() -> C.CDLL:
    """Load librve_b200.so (building it in-tree first if nvcc is available and it is stale)."""
    global _lib
    if _lib is not None:
        return _lib
    if not LIB_PATH.exists():
        from .build import build_native

        build_native()
    if not LIB_PATH.exists():
        raise RuntimeError(f"{LIB_PATH} is missing: run `python -m ravengine_b200.build` (no CPU fallback exists)")
    lib = C.CDLL(str(LIB_PATH))
    for name, (res, args) in PROTOTYPES.items():
        fn = getattr(lib, name)  # AttributeError = ABI mismatch, fail loudly
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib
