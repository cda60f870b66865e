/*
 * rve_shim.h — host-side mirror of the reference's per-frame encode blocks for the render-prep
 * path, as a C API over rve_b200.h. Each function reproduces one lambda of RenderEngine::Draw
 * (src/RenderEngine_Draw.cpp): the same table construction (CUBO table, indirect-init table,
 * offsets, dispatch sizes), the same buffer (re)allocation rules, then the rve_* call that
 * replaces "bind + SetComputeBytes + DispatchCompute". The data it walks is a stand-in for
 * World::RenderData (include/RavEngine/World.hpp:329-449).
 */
#ifndef RVE_SHIM_H
#define RVE_SHIM_H

#include "rve_b200.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct rvs_world rvs_world;

/* material kinds, used like filterRenderData(lightingFilter, materialInstance)
 * (src/RenderEngine_Draw.cpp:1232, 1080): a pass keeps a material iff (kind & pass_mask) != 0 */
#define RVS_MAT_LIT_OPAQUE 0x1u
#define RVS_MAT_UNLIT_OPAQUE 0x2u
#define RVS_MAT_LIT_TRANSPARENT 0x4u
#define RVS_MAT_UNLIT_TRANSPARENT 0x8u
#define RVS_MAT_ALL 0xFu

/* flags for rvs_world_create */
#define RVS_QUIRK_INDIRECT_SLOT 0x1u /* write the indirect-init entry at (meshID + lodID) exactly like
                                        src/RenderEngine_Draw.cpp:1314 (collides when a material holds several
                                        multi-LOD collections); default: running LOD offset = what the shader reads */

#define RVS_SORT_ENTITY_LISTS 0x2u   /* static culls read an ascending device copy of every command's entity list
                                        (rve_sort_entity_ids), rebuilt when the list changed: a list bound with
                                        rvs_material_add_command_set after a sync that uploaded something, a caller-owned
                                        buffer after rvs_material_set_command_entities. Results are identical (compared as
                                        sets per slice); skinned (single-instance) culls never use it */
RVE_API int rvs_world_create(rve_ctx* ctx, uint32_t flags, rvs_world** out);
RVE_API int rvs_world_destroy(rvs_world* w);

/* World::RenderData::worldTransforms / renderLayers / perObjectAttributes private buffers
 * (include/RavEngine/World.hpp:440-446), indexed by entity id. */
RVE_API int rvs_world_set_entity_buffers(rvs_world* w, rve_buffer model_matrices, rve_buffer render_layers,
                                         rve_buffer per_object_flags);

/* MeshCollectionStatic (include/RavEngine/MeshCollection.hpp:24-59): bounding radius, per-LOD
 * distance thresholds (uploaded to a device buffer like lodDistances.GetPrivateBuffer()) and the
 * per-LOD mesh ranges in the shared arena (totalIndices, getIndexRangeStart, getVertexRangeStart). */
RVE_API int rvs_mesh_collection_create(rvs_world* w, float radius, uint32_t num_lods, const float* lod_distances,
                                       const uint32_t* index_counts, const uint32_t* index_starts,
                                       const uint32_t* base_vertices, uint32_t* out_mesh);

/* staticMeshRenderData: ordered map material -> MDIICommand{commands[]} (World.hpp:339-349).
 * Materials are visited in creation order. */
RVE_API int rvs_material_create(rvs_world* w, uint32_t kind, uint32_t* out_material);
/* one command = (mesh collection, dense entity-id list); ids live in a device buffer
 * (command.entities.GetPrivateBuffer()). */
RVE_API int rvs_material_add_command(rvs_world* w, uint32_t material, uint32_t mesh, rve_buffer entity_ids,
                                     uint32_t entity_count);
/* update DenseSize()/buffer of an existing command */
RVE_API int rvs_material_set_command_entities(rvs_world* w, uint32_t material, uint32_t command_index,
                                              rve_buffer entity_ids, uint32_t entity_count);

/* cullTheRenderData(staticMeshRenderData) — src/RenderEngine_Draw.cpp:1221-1407. */
RVE_API int rvs_cull_static(rvs_world* w, const float view_proj[16], const float cam_pos[3], uint32_t camera_layers,
                            uint32_t shadow_pass, uint32_t pass_mask, rve_pyramid pyramid);
/* generatePyramid for the camera's target (src/RenderEngine_Draw.cpp:2475-2514) and cullTheRenderData for that camera
 * (:1221-1407) as ONE encode of two launches: a persistent pyramid kernel that builds the tile levels (mips 0..5) and
 * writes the pass tables it carries as a parameter block, then the cull, whose first CTAs derive the pyramid's upper
 * levels while the first frustum phases run (every warp checks that they are done before its first Hi-Z batch).
 * Same results as rve_depth_pyramid_build followed by rvs_cull_static - including the complete pyramid, which other
 * consumers may read once this call's work has run. RVE_PYRAMID_DEFER=0 keeps upload, pyramid (two launches) and cull apart. */
RVE_API int rvs_cull_static_with_pyramid(rvs_world* w, const float view_proj[16], const float cam_pos[3], uint32_t camera_layers,
                                         uint32_t shadow_pass, uint32_t pass_mask, rve_pyramid pyramid, rve_buffer depth,
                                         uint32_t depth_width, uint32_t depth_height);
/* All views of a frame in one pass: the reference calls cullTheRenderData once per view - main camera,
 * shadow cascades, spot / point shadow faces (src/RenderEngine_Draw.cpp:1769-1866, 1875-1987) - re-walking the scene
 * each time and reusing the material buffers between culls. Here the tables are built once and rve_cull_views tests
 * every instance against all views in one launch; each material's cullingBuffer / indirectBuffer then hold num_views
 * consecutive copies of the single-view layout (view v at v * numEntities*numLODs ids / v * numLODs commands).
 * Batches of fewer than 8 views run pass by pass (one rve_cull per view, CUBO offsets displaced per view) into the same
 * layout: below that the batched kernel's fixed cost loses to separate launches (profiles/r01r_views.jsonl). */
typedef struct rvs_view {
    float view_proj[16];
    float cam_pos[3];
    uint32_t camera_layers;
    uint32_t shadow_pass;
} rvs_view;
RVE_API int rvs_cull_static_views(rvs_world* w, uint32_t num_views, const rvs_view* views, const rve_pyramid* pyramids,
                                  uint32_t pass_mask);
/* drawcommand.cullingBuffer / indirectBuffer of a material after rvs_cull_static */
RVE_API int rvs_material_outputs(rvs_world* w, uint32_t material, rve_buffer* culling_buffer, rve_buffer* indirect_buffer,
                                 uint32_t* num_lods_total, uint32_t* num_entities_total);
/* the world's CUBO buffer and the numCubos / totalEntities of the last static cull (for inspection) */
RVE_API int rvs_last_cull_info(rvs_world* w, rve_buffer* cubo_buffer, uint32_t* num_cubos, uint32_t* total_entities);
/* RVS_SORT_ENTITY_LISTS bookkeeping since the world was created (for inspection): lists whose ascending copy was rebuilt from
 * the whole list, and lists whose copy followed a small change of their sparse set through the command's own bitmap (only the
 * ids that entered / left the set applied, copy re-expanded) */
RVE_API int rvs_sorted_list_stats(rvs_world* w, uint32_t* full_rebuilds, uint32_t* delta_updates);

/* ---- skinned path ---- */
/* MeshAssetSkinned: vertex/index ranges in the shared arena, radius, weights buffer */
RVE_API int rvs_skinned_mesh_create(rvs_world* w, uint32_t num_verts, uint32_t num_indices, uint32_t index_start,
                                    uint32_t vertex_start, float radius, rve_buffer weights, uint32_t* out_mesh);
RVE_API int rvs_skinned_material_create(rvs_world* w, uint32_t kind, uint32_t* out_material);
/* MDIICommandSkinned::command (World.hpp:351-362). `palettes` = entity_count * num_bones mat4 in
 * entities.GetReverseMap() order; caller-owned host memory read at rvs_pose_skeletal_meshes time
 * (AnimatorComponent::GetSkinningMats, src/RenderEngine_Draw.cpp:611-619). */
RVE_API int rvs_skinned_material_add_command(rvs_world* w, uint32_t material, uint32_t mesh, uint32_t num_bones,
                                             rve_buffer entity_ids, uint32_t entity_count, const float* palettes);
/* sharedPosition/Normal/Tangent/Bitangent/UV0 buffers of the mesh arena (src/RenderEngine_Allocate.cpp:168-177) */
RVE_API int rvs_world_set_mesh_arena(rvs_world* w, const rve_buffer in_streams[5]);
/* prepareSkeletalMeshBuffers — :479-544 (pow2 growth of shared skinned streams, palette buffer,
 * per-material indirect + culling buffers). Returns skeletalMeshesExist in *exist. */
RVE_API int rvs_prepare_skeletal_mesh_buffers(rvs_world* w, uint32_t* exist);
/* poseSkeletalMeshes — :577-630 (palette upload + one rve_skin per command) */
RVE_API int rvs_pose_skeletal_meshes(rvs_world* w);
/* prepareSkeletalCullingBuffer — :546-574 (one rve_skinned_drawcall_prepare per command) */
RVE_API int rvs_prepare_skeletal_culling_buffer(rvs_world* w);
/* cullSkeletalMeshes — :1067-1218 (single-instance mode; dispatch sized by the LAST command, :1214) */
RVE_API int rvs_cull_skeletal(rvs_world* w, const float view_proj[16], const float cam_pos[3], uint32_t camera_layers,
                              uint32_t shadow_pass, uint32_t pass_mask, rve_pyramid pyramid);
RVE_API int rvs_skinned_outputs(rvs_world* w, rve_buffer out_streams[5], rve_buffer* pose_buffer);
RVE_API int rvs_skinned_material_outputs(rvs_world* w, uint32_t material, rve_buffer* culling_buffer,
                                         rve_buffer* indirect_buffer, uint32_t* num_entities_total);

/* ---- clustered lights: renderFromPerspective<includeLighting> prologue, :928-974 ---- */
RVE_API int rvs_bin_lights(rvs_world* w, const float inv_proj[16], const float view_mat[16], const uint32_t grid_size[3],
                           const uint32_t screen_dim[2], float z_near, float z_far, rve_buffer point_lights,
                           uint32_t point_light_count, rve_buffer spot_lights, uint32_t spot_light_count,
                           rve_buffer* out_cluster_buffer);

/* ---- one frame in one submission: the pass schedule of RenderEngine::Draw without its redundancy ---- */
/* One camera of the frame (RenderTargetCollection.hpp:45-54 camData + its target's depth pyramid) */
typedef struct rvs_frame_camera {
    float view_proj[16], view[16], inv_proj[16];
    float cam_pos[3];
    float z_near, z_far;
    uint32_t layers;
    rve_pyramid pyramid;  /* target.depthPyramid */
    rve_buffer depth;     /* last frame's depth, rebuilt into the pyramid first; RVE_BUFFER_INVALID keeps the pyramid */
    uint32_t depth_width, depth_height;
    uint32_t screen_dim[2];
} rvs_frame_camera;
typedef struct rvs_frame_desc {
    const rvs_frame_camera* cameras;
    uint32_t num_cameras;
    const rvs_view* shadow_views;       /* cascades (rve_csm_cascades), spot lights, point-light faces */
    const rve_pyramid* shadow_pyramids; /* one per shadow view */
    uint32_t num_shadow_views;
    rve_buffer point_lights;
    uint32_t point_light_count;
    rve_buffer spot_lights;
    uint32_t spot_light_count;
    uint32_t grid_size[3];              /* cluster grid */
    uint32_t pass_mask;                 /* RVS_MAT_* classes to cull; 0 = all */
} rvs_frame_desc;
/* Replaces, for one frame, the per-pass culls and light binning of src/RenderEngine_Draw.cpp:1769-1866 (shadow views),
 * :2177-2209 and :2567-2655 (six renderFromPerspective passes per camera, each with a cullTheRenderData over its
 * material class; three of them re-bin the lights): pyramids built once, view-independent skinned work once, ALL
 * views culled in one rve_cull_views launch (view v of every material's buffers: cameras first, then the shadow
 * views, laid out as rvs_cull_static_views describes), lights binned once per camera. */
RVE_API int rvs_frame_encode(rvs_world* w, const rvs_frame_desc* frame);
/* the cluster buffer rvs_frame_encode filled for camera `camera` */
RVE_API int rvs_frame_cluster_buffer(rvs_world* w, uint32_t camera, rve_buffer* out);

/* ---- dirty-range upload: BufferedVRAMVector / EncodeSync (include/RavEngine/BufferedVRAMVector.hpp:33-209,
 * src/BufferedVRAMVector.cpp:27-144) - the step before the render-prep path ---- */
typedef struct rvs_vector rvs_vector;
typedef struct rvs_sync_stats {
    uint32_t ranges;         /* ranges the reference's scan produces (maxGapSize 5) */
    uint32_t rows_in_ranges; /* elements they cover */
    uint32_t direct_copies;  /* ranges sent as one host->device copy each */
    uint32_t packed_rows;    /* modified rows of short ranges, sent through the packed staging block + scatter kernel */
    uint64_t direct_bytes;
    uint64_t packed_bytes;   /* staging block size (rows + u32 indices) */
} rvs_sync_stats;
#define RVS_SYNC_RANGES 0u /* the reference's behaviour: one copy per range */
#define RVS_SYNC_PACKED 1u /* ranges shorter than direct_min_rows: ONE staging upload + ONE scatter kernel */
#define RVS_SYNC_PULL 2u   /* no host scan at all: one kernel reads the flags and the flagged rows from the pinned host
                              arrays over PCIe, writes the rows to the private buffer and clears the flags. The flag
                              block is double-buffered: rvs_vector_tracking() changes at every such sync. Like the
                              reference's shared-buffer copies, rows must not be rewritten before the stream got there. */

#define RVS_SYNC_FULL 3u   /* every row is uploaded, flags are not consulted (and end up clear): the first sync after a
                              Resize (src/BufferedVRAMVector.cpp:31-81 copies the whole buffer) and frames in which
                              every entity moved. One copy-engine copy; affine vectors: chunked copies into device
                              staging, each chunk expanded on a second stream while the next crosses PCIe */
#define RVS_SYNC_FULL_PULL 4u /* the same by ONE kernel that reads the pinned host rows over PCIe (no staging at all) */

RVE_API int rvs_vector_create(rve_ctx* ctx, uint32_t elem_bytes, uint32_t count, rvs_vector** out);
/* World::RenderData::worldTransforms is a BufferedVRAMVector<matrix4> (include/RavEngine/World.hpp:440) written once per
 * moved entity by SetValueAt(id, GetWorldMatrix()) (src/World.cpp:346,385). A world matrix is affine: its fourth row is
 * (0,0,0,1). With RVS_VECTOR_AFFINE_ROWS the HOST rows are glm::mat4x3 (elem_bytes must be 48: columns c0..c3, xyz each)
 * while the private buffer keeps the reference's 64-B mat4 rows (w = 0,0,0,1 filled in on the device), so every consumer
 * of the private buffer is unchanged and 25 % fewer bytes cross PCIe. Every sync mode works; byte-identical to a 64-B
 * vector holding the same matrices. */
#define RVS_VECTOR_AFFINE_ROWS 0x1u
/* The host block is a memfd mapping (pinned with cudaHostRegister) instead of a cudaHostAlloc block, so that ANOTHER rank
 * of the same box can map and pin it too (rvs_vector_share / rve_host_import) and carry part of this rank's upload over
 * its own PCIe link (rvs_forward_rows). The block is allocated once: such a vector cannot grow past its first size. */
#define RVS_VECTOR_SHARED_HOST 0x2u
RVE_API int rvs_vector_create_ex(rve_ctx* ctx, uint32_t elem_bytes, uint32_t count, uint32_t flags, rvs_vector** out);
RVE_API int rvs_vector_destroy(rvs_vector* v);
/* Resize: new elements are zero and start out modified (BufferedVRAMVector.hpp:64-69) */
RVE_API int rvs_vector_resize(rvs_vector* v, uint32_t count);
RVE_API uint32_t rvs_vector_size(rvs_vector* v);
/* GetHostBuffer(): pinned host memory, count * elem_bytes; valid until the next growth */
RVE_API void* rvs_vector_host(rvs_vector* v);
/* syncTrackingBuffer: one byte per element, non-zero = modified since the last sync. A vector nobody wrote to (no
 * rvs_vector_set / _mark / _resize and no call of this accessor since the last sync) syncs without touching its flags. */
RVE_API uint8_t* rvs_vector_tracking(rvs_vector* v);
/* SetValueAt / GetValueAtForWriting (hpp:113-121) */
RVE_API int rvs_vector_set(rvs_vector* v, uint32_t index, const void* value);
RVE_API int rvs_vector_mark(rvs_vector* v, uint32_t begin, uint32_t end);
/* EncodeSync: enqueues the copies on the context's stream; afterwards the private buffer equals the host buffer */
RVE_API int rvs_vector_encode_sync(rvs_vector* v, uint32_t mode, uint32_t direct_min_rows, rvs_sync_stats* stats);
/* RVS_SYNC_FULL / RVS_SYNC_FULL_PULL restricted to rows [row_begin, row_end): the owner's share of a forwarded upload */
/* EncodeSync of n vectors in one go (mode RVS_SYNC_PACKED): vectors with small changes share ONE staging copy and ONE scatter
 * launch per 16 vectors - a frame in which many BufferedVRAM structures change a little (entity lists, layers, attributes; the
 * reference syncs each one by itself, src/RenderEngine_Draw.cpp:112-146) pays one copy and one launch instead of one pair per
 * structure; the others take their own rvs_vector_encode_sync. stats: n entries or null. */
RVE_API int rvs_vectors_encode_sync_batch(rvs_vector* const* vectors, uint32_t n, uint32_t direct_min_rows, rvs_sync_stats* stats);
RVE_API int rvs_vector_encode_sync_rows(rvs_vector* v, uint32_t mode, uint32_t row_begin, uint32_t row_end, rvs_sync_stats* stats);
/* (pid, fd, mapped bytes) of a RVS_VECTOR_SHARED_HOST vector's host block: what a peer passes to rve_host_import */
RVE_API int rvs_vector_share(rvs_vector* v, int32_t* pid, int32_t* fd, uint64_t* bytes);

/* ---- upload forwarding between ranks of one box. The reference has one GPU: its EncodeSync
 * copies (src/BufferedVRAMVector.cpp:87-104) all cross that GPU's PCIe link. With
 * one process per GPU the links are not equal (on the pool's 8-GPU boxes four GPUs share one path to host memory), and a
 * rank on a fast link can carry the tail of a slow rank's rows: it maps the owner's host block, reads the rows over its
 * own link with one kernel and writes them - expanded to 64-B rows when `affine` - straight into the owner's private
 * buffer over NVLink peer memory. ---- */
/* maps and pins the host block (pid, fd) of another process (or this one); *out_ptr is usable by kernels of ctx's device */
RVE_API int rve_host_import(rve_ctx* ctx, int32_t pid, int32_t fd, uint64_t bytes, void** out_ptr);
RVE_API int rve_host_release(rve_ctx* ctx, void* ptr, uint64_t bytes);
/* flag words (u32) in a buffer both ranks have mapped (rve_ipc_export / rve_ipc_import): signal = release-store of `value`
 * behind everything enqueued so far; wait = spin until the word reaches `value` (wrap-safe), ~2 s timeout -> *status |= 1 */
RVE_API int rve_flag_signal(rve_ctx* ctx, rve_buffer flags, uint32_t word, uint32_t value);
RVE_API int rve_flag_wait(rve_ctx* ctx, rve_buffer flags, uint32_t word, uint32_t value, rve_buffer status);
/* rows [first_row, first_row + rows) of the imported host block (row_bytes each; affine != 0: 48-B rows expanded to 64-B
 * rows) -> the same rows of `dst` (the owner's private buffer, imported), on ctx's stream */
RVE_API int rvs_forward_rows(rve_ctx* ctx, const void* host_rows, uint32_t row_bytes, uint32_t affine, uint64_t first_row, uint64_t rows,
                             rve_buffer dst);

/* GetPrivateBuffer(): may change at every rvs_vector_encode_sync after a growth, like the reference's */
RVE_API rve_buffer rvs_vector_private_buffer(rvs_vector* v);
/* rows copied by RVS_SYNC_PULL syncs since creation (reads a device counter: synchronises the stream) */
RVE_API uint64_t rvs_vector_pulled_rows(rvs_vector* v);
/* (begin, end) pairs of the last sync, for inspection; returns their number */
RVE_API uint32_t rvs_vector_last_ranges(rvs_vector* v, uint32_t* out_pairs, uint32_t max_pairs);
/* The range derivation by itself, host only (no device needed): from a flag array by the reference's scan
 * (src/BufferedVRAMVector.cpp:106-139, maxGapSize 5), and from the ascending list of flagged indices - the shortcut a sync
 * takes when every flag of the frame went through SetValueAt / mark / Resize. Both return the number of (begin, end) pairs. */
RVE_API uint32_t rvs_sync_ranges_from_flags(const uint8_t* flags, uint32_t count, uint32_t* out_pairs, uint32_t max_pairs);
RVE_API uint32_t rvs_sync_ranges_from_indices(const uint32_t* indices, uint32_t n, uint32_t count, uint32_t* out_pairs, uint32_t max_pairs);

/* ---- entity lists: BufferedVRAMSparseSet<entity_id_t, entity_id_t> (include/RavEngine/World.hpp:342,355: MDIICommand::command::
 * entities; include/RavEngine/SparseSet.hpp:6-104; include/RavEngine/BufferedVRAMVector.hpp:129-170). The dense array is what the
 * cull reads (entityIDInputBufferBindlessHandle); erase is a swap-remove, so a live engine's lists are dense runs with
 * out-of-order entries wherever something was removed. ---- */
typedef struct rvs_sparse_set rvs_sparse_set;
RVE_API int rvs_sparse_set_create(rve_ctx* ctx, rvs_sparse_set** out);
RVE_API int rvs_sparse_set_destroy(rvs_sparse_set* s);
/* Emplace(sparse_index, value): no-op when the index has a record; the new dense element is flagged modified */
RVE_API int rvs_sparse_set_emplace(rvs_sparse_set* s, uint32_t sparse_index, uint32_t value);
/* EraseAtSparseIndex: the last dense element moves into the hole (flagged modified there); an unknown index is an error
 * (the reference asserts) */
RVE_API int rvs_sparse_set_erase(rvs_sparse_set* s, uint32_t sparse_index);
RVE_API int rvs_sparse_set_emplace_many(rvs_sparse_set* s, const uint32_t* sparse_indices, const uint32_t* values, uint32_t n);
RVE_API int rvs_sparse_set_erase_many(rvs_sparse_set* s, const uint32_t* sparse_indices, uint32_t n);
RVE_API int rvs_sparse_set_has(rvs_sparse_set* s, uint32_t sparse_index);
RVE_API uint32_t rvs_sparse_set_dense_size(rvs_sparse_set* s);
/* GetDenseData() / GetReverseMap(): host views, valid until the next mutation */
RVE_API const uint32_t* rvs_sparse_set_dense(rvs_sparse_set* s);
RVE_API const uint32_t* rvs_sparse_set_reverse_map(rvs_sparse_set* s);
/* the dense array as a BufferedVRAMVector stand-in (rvs_vector_encode_sync / rvs_vector_private_buffer) */
RVE_API rvs_vector* rvs_sparse_set_vector(rvs_sparse_set* s);
/* MDIICommand::command with its entities held in a sparse set: every cull over the world first syncs the set (EncodeSync,
 * packed mode) and reads its private buffer and DenseSize() - rvs_material_add_command with a live list */
RVE_API int rvs_material_add_command_set(rvs_world* w, uint32_t material, uint32_t mesh, rvs_sparse_set* entities);

/* ---- shared vertex arena + .rvem container: the data formats on the input side of skinning ---- */
/* RenderEngine's shared mesh buffers (src/RenderEngine_Allocate.cpp:168-177): positions, normals, tangents, bitangents
 * (12 B), uv0, lightmapUV (8 B), u32 indices; offsets and sizes are ELEMENTS (include/RavEngine/MeshAllocation.hpp:27-52). */
typedef struct rvs_arena rvs_arena;
/* MeshPartView (include/RavEngine/Mesh.hpp:35-60): host pointers, any vertex stream may be NULL (absent attribute) */
typedef struct rvs_mesh_view {
    const float *positions, *normals, *tangents, *bitangents, *uv0, *lightmap_uvs;
    const uint32_t* indices;
    uint32_t num_vertices, num_indices;
} rvs_mesh_view;
/* initial sizes 0 = the reference's (1024 vertices, 1536 indices; include/RavEngine/RenderEngine.hpp:87) */
RVE_API int rvs_arena_create(rve_ctx* ctx, uint32_t initial_verts, uint32_t initial_indices, rvs_arena** out);
RVE_API int rvs_arena_destroy(rvs_arena* a);
/* RenderEngine::AllocateMesh (src/RenderEngine_Allocate.cpp:12-100): placement, growth and upload; out_range is the
 * MeshRange (it follows re-packs; query it with rvs_arena_range after any later allocation) */
RVE_API int rvs_arena_allocate_mesh(rvs_arena* a, const rvs_mesh_view* mesh, uint32_t* out_range);
/* RenderEngine::DeallocateMesh (src/RenderEngine_Allocate.cpp:101-145) */
RVE_API int rvs_arena_deallocate_mesh(rvs_arena* a, uint32_t range);
/* MeshRange::getVertexRangeStart / getIndexRangeStart (+ counts) */
RVE_API int rvs_arena_range(rvs_arena* a, uint32_t range, uint32_t* vertex_start, uint32_t* vertex_count, uint32_t* index_start,
                            uint32_t* index_count);
/* sharedPositionBuffer .. sharedLightmapUVBuffer, sharedIndexBuffer and currentVertexSize / currentIndexSize; the
 * handles change when the arena grows (the reference re-creates its buffers too) */
RVE_API int rvs_arena_buffers(rvs_arena* a, rve_buffer out_streams[7], uint32_t* vertex_capacity, uint32_t* index_capacity);
#define RVS_ARENA_VERTEX_FREE 0u
#define RVS_ARENA_INDEX_FREE 1u
#define RVS_ARENA_VERTEX_ALLOCATED 2u
#define RVS_ARENA_INDEX_ALLOCATED 3u
/* (start, count) pairs of vertexFreeList / indexFreeList / vertexAllocatedList / indexAllocatedList in list order, for
 * inspection; returns the list length */
RVE_API uint32_t rvs_arena_list(rvs_arena* a, uint32_t which, uint32_t* out_pairs, uint32_t max_pairs);

/* SerializedMeshDataHeader attribute bits (include/RavEngine/Mesh.hpp:18-26) */
#define RVS_RVEM_SKINNED 0x01u
#define RVS_RVEM_HAS_POSITIONS 0x02u
#define RVS_RVEM_HAS_NORMALS 0x04u
#define RVS_RVEM_HAS_TANGENTS 0x08u
#define RVS_RVEM_HAS_BITANGENTS 0x10u
#define RVS_RVEM_HAS_UV0 0x20u
#define RVS_RVEM_HAS_LIGHTMAP_UV 0x40u
/* Where each stream of a .rvem image lies (byte offsets into the image; ~0 = absent) */
typedef struct rvs_rvem_info {
    uint32_t num_vertices, num_indices, attributes, reserved;
    uint64_t off_positions, off_normals, off_tangents, off_bitangents, off_uv0, off_lightmap_uv, off_indices;
    uint64_t mesh_bytes;    /* end of the index stream = MeshAsset::DeserializeMeshFromMemory's second result */
    uint64_t off_weights;   /* VertexWeights table of a skinned mesh (32 B per vertex), ~0 when fewer than 4 bytes follow */
    uint64_t weights_bytes; /* whole entries only */
} rvs_rvem_info;
/* MeshAsset::DeserializeMeshFromMemory (src/MeshAsset.cpp:23-79), host only, zero-copy: validates the header and sizes */
RVE_API int rvs_rvem_parse(const void* data, uint64_t size, rvs_rvem_info* out);
/* MeshAsset(name) / MeshAssetSkinned(path) (src/MeshAsset.cpp:81-87, src/MeshAssetSkinned.cpp:25-70): parse, AllocateMesh,
 * and for skinned != 0 the weights buffer (VertexWeights[num_vertices], binding 21 of skinning_cs.csh) */
RVE_API int rvs_arena_load_rvem(rvs_arena* a, const void* data, uint64_t size, uint32_t skinned, uint32_t* out_range,
                                rve_buffer* out_weights);

#ifdef __cplusplus
}
#endif
#endif
