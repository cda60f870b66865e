/*
 * rve_b200.h — C-ABI of the B200-native render-prep path (drop-in boundary).
 *
 * One header, plain C: POD layouts that are byte-identical to the reference's
 * push-constant / storage-buffer structs, and the entry points a reference
 * maintainer would bind instead of "bind + SetComputeBytes + DispatchCompute"
 * on the six compute pipelines of RenderEngine::Draw.
 *
 * Every declaration cites the reference interface it replaces
 * (paths relative to the reference checkout).
 *
 * Conventions
 *   - all functions return 0 on success, a negative rve_status on failure,
 *     never throw across the boundary; rve_last_error_string() describes the
 *     last failure of the calling thread (reference: Debug::Fatal aborts,
 *     include/RavEngine/Debug.hpp — we return codes instead).
 *   - every rve_* compute call is an asynchronous enqueue on the context's
 *     stream (the reference's in-order queue, src/RenderEngine.cpp:304-307),
 *     exactly like ICommandBuffer::DispatchCompute
 *     (deps/RGL/include/RGL/CommandBuffer.hpp:99-184). Push-constant structs
 *     are copied at enqueue time.
 *   - mat4 is column-major 16 x f32 (glm / GLSL): m[col*4 + row].
 *   - buffer sizes/offsets are 64-bit (the reference uses u32,
 *     deps/RGL/include/RGL/Buffer.hpp:22,75); instance ids stay u32.
 */
#ifndef RVE_B200_H
#define RVE_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define RVE_API __attribute__((visibility("default")))
#else
#define RVE_API
#endif

/* ------------------------------------------------------------------ */
/* status codes                                                        */
/* ------------------------------------------------------------------ */
typedef enum rve_status {
    RVE_OK = 0,
    RVE_ERR_INVALID_ARGUMENT = -1,
    RVE_ERR_OUT_OF_MEMORY = -2,
    RVE_ERR_CUDA = -3,
    RVE_ERR_BAD_HANDLE = -4,
    RVE_ERR_NO_DEVICE = -5,
    RVE_ERR_UNSUPPORTED = -6
} rve_status;

/* ------------------------------------------------------------------ */
/* POD layouts (bit-identical to the reference)                        */
/* ------------------------------------------------------------------ */
#pragma pack(push, 1)

/* RGL::IndirectIndexedCommand — deps/RGL/include/RGL/CommandBuffer.hpp:55-61
 * (consumed by vkCmdDrawIndexedIndirect with stride 20,
 *  deps/RGL/src/VkCommandBuffer.cpp:790-799). */
typedef struct RveIndirectIndexedCommand {
    uint32_t indexCount;
    uint32_t instanceCount;
    uint32_t indexStart;
    uint32_t baseVertex;
    uint32_t baseInstance;
} RveIndirectIndexedCommand;

/* CullingUBO — include/RavEngine/RenderEngine.hpp:224-230,
 * shaders/defaultcull.csh:19-25 (push constants, scalar layout, 88 B). */
typedef struct RveCullingUBO {
    float viewProj[16];
    float camPos[3];
    uint32_t numCubos;
    uint32_t cameraRenderLayers;
    uint32_t singleInstanceModeAndShadowMode; /* bit0 single-instance (skinned), bit1 shadow pass */
} RveCullingUBO;

/* CullingUBOinstance ("CUBO") — include/RavEngine/RenderEngine.hpp:212-222,
 * shaders/defaultcull.csh:7-17 (binding 0, scalar array, 36 B each). */
typedef struct RveCullingUBOinstance {
    uint32_t indirectBufferOffset;
    uint32_t numObjects;
    uint32_t cullingBufferOffset;
    float radius;
    uint32_t numLODs;
    uint32_t idOutputBufferBindlessHandle;
    uint32_t entityIDInputBufferBindlessHandle;
    uint32_t indirectOutputBufferBindlessHandle;
    uint32_t lodDistanceBufferBindlessHandle;
} RveCullingUBOinstance;

/* SkinningUBO — include/RavEngine/RenderEngine.hpp:178-185,
 * shaders/skinning_cs.csh:83-90 (24 B). */
typedef struct RveSkinningUBO {
    uint32_t numObjects;
    uint32_t numVertices;
    uint32_t numBones;
    uint32_t boneReadOffset;
    uint32_t vertexWriteOffset;
    uint32_t vertexReadOffset;
} RveSkinningUBO;

/* SkinningPrepareUBO — include/RavEngine/RenderEngine.hpp:232-240,
 * shaders/skinned_mesh_drawcall.csh:1-9 (28 B). */
typedef struct RveSkinningPrepareUBO {
    uint32_t indexBufferOffset;
    uint32_t vertexBufferOffset;
    uint32_t nIndicesInThisMesh;
    uint32_t nVerticesInThisMesh;
    uint32_t nTotalObjects;
    uint32_t drawCallBufferOffset;
    uint32_t baseInstanceOffset;
} RveSkinningPrepareUBO;

/* GridBuildUBO — include/RavEngine/RenderEngine.hpp:103-112 (pack(1), 92 B),
 * shaders/cluster_build_grid.csh:8-14. */
typedef struct RveGridBuildUBO {
    float invProj[16];
    uint32_t gridSize[3];
    float zNear;
    uint32_t screenDim[2];
    float zFar;
} RveGridBuildUBO;

/* GridAssignUBO — include/RavEngine/RenderEngine.hpp:114-117 (72 B),
 * shaders/cluster_assign_lights.csh:22-26. */
typedef struct RveGridAssignUBO {
    float viewMat[16];
    uint32_t pointLightCount;
    uint32_t spotLightCount;
} RveGridAssignUBO;

/* PyramidCopyUBO — include/RavEngine/RenderEngine.hpp:89-91,
 * shaders/depthpyramidcopy.fsh:1-3. */
typedef struct RvePyramidCopyUBO {
    uint32_t size;
} RvePyramidCopyUBO;

#define RVE_CLUSTER_MAX_POINTS 16 /* shaders/cluster_defs.h:1 */
#define RVE_CLUSTER_MAX_SPOTS 16  /* shaders/cluster_defs.h:2 */
#define RVE_LIGHT_MIN_INFLUENCE 0.01f /* shaders/cluster_defs.h:3 */

/* Cluster — shaders/cluster_shared.glsl:3-11 (scalar layout, 160 B). */
typedef struct RveCluster {
    float minPoint[3];
    float maxPoint[3];
    uint32_t pointLightCount;
    uint32_t spotLightCount;
    uint32_t pointLightIndices[RVE_CLUSTER_MAX_POINTS];
    uint32_t spotLightIndices[RVE_CLUSTER_MAX_SPOTS];
} RveCluster;

/* PointLight — shaders/cluster_shared.glsl:14-25, upload struct
 * include/RavEngine/World.hpp:393-405 (scalar layout, 512 B). */
typedef struct RvePointLight {
    float viewMats[6][16];
    float projMat[16];
    float position[3];
    float color[3];
    float intensity;
    int32_t castsShadows;
    uint32_t shadowRenderLayers;
    uint32_t illuminationLayers;
    uint32_t shadowmapBindlessIndices[6];
} RvePointLight;

/* SpotLight — shaders/cluster_shared.glsl:27-38, upload struct
 * include/RavEngine/World.hpp:406-416 (scalar layout, 168 B). */
typedef struct RveSpotLight {
    float lightViewProj[16];
    float worldTransform[16];
    float color[3];
    float intensity;
    float coneAngle;
    float penumbraAngle;
    int32_t castsShadows;
    uint32_t shadowmapBindlessIndex;
    uint32_t shadowRenderLayers;
    uint32_t illuminationLayers;
} RveSpotLight;

/* VertexJointBinding — shaders/skinning_cs.csh:66-75,
 * include/RavEngine/Skeleton.hpp:50-60 (std430, 32 B). */
typedef struct RveJointInfluence {
    uint32_t joint_index;
    float influence;
} RveJointInfluence;
typedef struct RveVertexJointBinding {
    RveJointInfluence weights[4];
} RveVertexJointBinding;

#pragma pack(pop)

/* per-object attribute bits — include/RavEngine/Layer.hpp:8-13 */
#define RVE_ATTR_FRUSTUM_CULL 0x1u
#define RVE_ATTR_OCCLUSION_CULL 0x2u
#define RVE_ATTR_CASTS_SHADOWS 0x4u
/* CullingUBO.singleInstanceModeAndShadowMode bits — defaultcull.csh:186-187 */
#define RVE_CULL_SINGLE_INSTANCE 0x1u
#define RVE_CULL_SHADOW_MODE 0x2u

#ifdef __cplusplus
static_assert(sizeof(RveIndirectIndexedCommand) == 20, "layout");
static_assert(sizeof(RveCullingUBO) == 88, "layout");
static_assert(sizeof(RveCullingUBOinstance) == 36, "layout");
static_assert(sizeof(RveSkinningUBO) == 24, "layout");
static_assert(sizeof(RveSkinningPrepareUBO) == 28, "layout");
static_assert(sizeof(RveGridBuildUBO) == 92, "layout");
static_assert(sizeof(RveGridAssignUBO) == 72, "layout");
static_assert(sizeof(RveCluster) == 160, "layout");
static_assert(sizeof(RvePointLight) == 512, "layout");
static_assert(sizeof(RveSpotLight) == 168, "layout");
static_assert(sizeof(RveVertexJointBinding) == 32, "layout");
#endif

/* ------------------------------------------------------------------ */
/* context                                                             */
/* ------------------------------------------------------------------ */
typedef struct rve_ctx rve_ctx;

/* One context per (thread, device) = the reference's single device + single
 * in-order queue (src/RenderEngine.cpp:304-307). `cuda_stream` may be NULL
 * (the context creates its own non-blocking stream) or an existing
 * cudaStream_t to enqueue on. */
RVE_API int rve_ctx_create(int device_ordinal, void* cuda_stream, rve_ctx** out_ctx);
RVE_API int rve_ctx_destroy(rve_ctx* ctx);
/* ICommandBuffer::BlockUntilCompleted — src/RenderEngine_Draw.cpp:471,2680 */
RVE_API int rve_ctx_synchronize(rve_ctx* ctx);
RVE_API void* rve_ctx_stream(rve_ctx* ctx);
RVE_API const char* rve_last_error_string(void);
RVE_API const char* rve_version(void);
/* number of SMs of the context's device (grid sizing information) */
RVE_API int rve_ctx_sm_count(rve_ctx* ctx);
/* CUBO entries dropped by rve_cull since the context was created because a bindless handle did not name a live buffer
 * (the reference would fault in defaultcull.csh:56-59); synchronises the stream */
RVE_API uint32_t rve_ctx_bad_cubo_handles(rve_ctx* ctx);
/* how many kernels this context has launched so far (bench.py "gpu_launches") */
RVE_API uint64_t rve_ctx_launch_count(rve_ctx* ctx);

/* One frame = one submission. The reference records its whole frame into one command buffer and commits it once
 * (src/App_priv.hpp:475-483); every rve_* / rvs_* call here is an enqueue on the context's stream, so a frame's calls can be
 * recorded into a CUDA graph and replayed with a single launch - what a small scene needs, whose step is bound by launch
 * latency rather than bandwidth (C1: 100 k instances). Between _begin and _end nothing executes; host tables passed to the
 * calls are copied into pinned blocks owned by the graph. Capture a frame after it has run once normally (buffers and
 * scratch sized, bindless table uploaded); replaying is valid while the scene's buffers, counts and views are the ones
 * recorded - per-entity data (transforms, layers, attributes, depth) may change freely, it is read at replay time. Calls that
 * synchronise (rve_buffer_download, rve_ctx_synchronize, profile reads) and the profiling events are not capturable. */
typedef struct rve_graph rve_graph;
RVE_API int rve_ctx_capture_begin(rve_ctx* ctx);
RVE_API int rve_ctx_capture_end(rve_ctx* ctx, rve_graph** out);
RVE_API int rve_graph_launch(rve_ctx* ctx, rve_graph* graph);
RVE_API int rve_graph_destroy(rve_graph* graph);

/* Host placement for the upload path that feeds this one (the reference's per-frame EncodeSync of the host-shared
 * worldTransforms / renderLayers / perObjectAttributes arrays, src/BufferedVRAMVector.cpp:27-144, from Shared-access
 * RGL buffers, deps/RGL/include/RGL/Buffer.hpp:25-35): binds the CALLING thread (threads it spawns inherit) to the CPUs
 * local to the context's GPU and makes the GPU's NUMA node the preferred source of the pages it allocates next - i.e.
 * of the pinned host arrays. One process per GPU on a multi-socket box otherwise pins wherever the scheduler put it and
 * every frame's upload crosses the socket interconnect. Linux sysfs only (no libnuma / NVML); a box without NUMA
 * information is left as it is (numa_node -1, nothing bound). Call it before creating rvs_vector objects. */
typedef struct rve_host_affinity {
    int32_t numa_node;      /* node of the GPU's PCIe root, -1 when the kernel does not say */
    uint32_t cpus_allowed;  /* CPUs the thread could use before the call */
    uint32_t cpus_local;    /* ... of which local to the GPU */
    uint32_t cpus_bound;    /* 1 when the thread's affinity was narrowed to the local CPUs */
    uint32_t memory_bound;  /* 1 when the preferred-node memory policy was set */
    char pci_bus_id[32];
} rve_host_affinity;
#define RVE_BIND_CPUS 0x1u
#define RVE_BIND_MEMORY 0x2u
RVE_API int rve_ctx_bind_host_to_device(rve_ctx* ctx, uint32_t flags, rve_host_affinity* out);

/* Per-kernel device timing (bench.py roofline): while enabled, the dominant kernel of each
 * stage is bracketed by CUDA events on the context's stream. kernel: 0 = cull, 1 = depth pyramid,
 * 2 = skinning, 3 = cluster assign. rve_ctx_profile_read synchronises, returns the summed
 * milliseconds and launch count since the last read, and clears them. */
#define RVE_PROF_CULL 0
#define RVE_PROF_PYRAMID 1
#define RVE_PROF_SKIN 2
#define RVE_PROF_CLUSTER_ASSIGN 3
RVE_API int rve_ctx_set_profiling(rve_ctx* ctx, int enabled);
RVE_API int rve_ctx_profile_read(rve_ctx* ctx, int kernel, double* total_ms, uint32_t* launches);

/* ------------------------------------------------------------------ */
/* buffers — RGL::IBuffer (deps/RGL/include/RGL/Buffer.hpp:50-81)      */
/* ------------------------------------------------------------------ */
/* A buffer handle IS its bindless index: the value a CUBO carries in its
 * *BindlessHandle fields (IBuffer::GetReadonlyBindlessGPUHandle /
 * GetReadwriteBindlessGPUHandle, deps/RGL/src/VkBuffer.cpp:67-88,150-157).
 * The library keeps a device-side pointer table so the cull kernel resolves
 * cubo.*BindlessHandle like the shader does (defaultcull.csh:56-59). */
typedef uint32_t rve_buffer;
#define RVE_BUFFER_INVALID 0xFFFFFFFFu

/* IDevice::CreateBuffer (Private access). Zero-initialised. */
RVE_API int rve_buffer_create(rve_ctx* ctx, size_t bytes, rve_buffer* out);
/* Register device memory the caller owns (e.g. a torch tensor) as a buffer. */
RVE_API int rve_buffer_wrap(rve_ctx* ctx, void* device_ptr, size_t bytes, rve_buffer* out);
/* Deferred to stream order (reference: gcBuffers queue, src/RenderEngine.cpp:2428-2442). */
RVE_API int rve_buffer_destroy(rve_ctx* ctx, rve_buffer buf);
/* IBuffer::SetBufferData / UpdateBufferData + staging CopyBufferToBuffer. */
RVE_API int rve_buffer_upload(rve_ctx* ctx, rve_buffer buf, size_t dst_offset, const void* host_src, size_t bytes);
/* Several small uploads in one stream operation (the reference records one CopyBufferToBuffer per table,
 * src/RenderEngine_Draw.cpp:1321-1329, 1367-1375). Data is consumed before the call returns. */
typedef struct rve_upload_segment {
    rve_buffer dst;
    size_t dst_offset;
    const void* src;
    size_t bytes;
} rve_upload_segment;
RVE_API int rve_buffer_upload_batch(rve_ctx* ctx, const rve_upload_segment* segs, uint32_t n);
RVE_API int rve_buffer_download(rve_ctx* ctx, rve_buffer buf, size_t src_offset, void* host_dst, size_t bytes);
/* ICommandBuffer::CopyBufferToBuffer — deps/RGL/include/RGL/CommandBuffer.hpp */
RVE_API int rve_buffer_copy(rve_ctx* ctx, rve_buffer src, size_t src_offset, rve_buffer dst, size_t dst_offset, size_t bytes);
RVE_API int rve_buffer_fill_zero(rve_ctx* ctx, rve_buffer buf, size_t offset, size_t bytes);
RVE_API void* rve_buffer_device_ptr(rve_ctx* ctx, rve_buffer buf);
RVE_API size_t rve_buffer_size(rve_ctx* ctx, rve_buffer buf);
RVE_API uint32_t rve_buffer_bindless_index(rve_ctx* ctx, rve_buffer buf);

/* ------------------------------------------------------------------ */
/* a-1 depth pyramid                                                   */
/* ------------------------------------------------------------------ */
/* DepthPyramid::DepthPyramid(width) — src/DepthPyramid.cpp:8-28:
 * dim = 2^floor(log2(min(width,1024))), numLevels = 1+log2(dim), R32F mip
 * chain. Ours: one linear f32 buffer, level L starts at texel offset
 * 4*(dim^2 - (dim>>L)^2)/3, row-major (dim>>L)^2 texels. */
typedef uint32_t rve_pyramid;
RVE_API int rve_depth_pyramid_create(rve_ctx* ctx, uint32_t width, rve_pyramid* out);
RVE_API int rve_depth_pyramid_destroy(rve_ctx* ctx, rve_pyramid pyr);
RVE_API int rve_depth_pyramid_info(rve_ctx* ctx, rve_pyramid pyr, uint32_t* dim, uint32_t* num_levels, rve_buffer* texels);
/* generatePyramid — src/RenderEngine_Draw.cpp:2475-2514: depthpyramidcopy.fsh
 * (mip 0, min-reduction sample of the w x h f32 depth buffer) followed by
 * depthpyramid.csh for mips 1..N-1. Whole chain in one call. */
RVE_API int rve_depth_pyramid_build(rve_ctx* ctx, rve_pyramid pyr, rve_buffer depth, uint32_t width, uint32_t height);
/* The same build for a caller that enqueues the camera's cull next (generatePyramid is followed by cullTheRenderData for
 * that camera, src/RenderEngine_Draw.cpp:2475-2514 then :1221-1407): the launch may stop after the tile levels (mips 0..5)
 * and leave the upper levels to the first CTAs of that cull, off the critical path. The library completes the
 * pyramid whichever of ITS calls touches it next - rve_cull / rvs_cull_static* inside the cull kernel, rve_cull_views and
 * rve_depth_pyramid_footmin, rve_buffer_download and rve_buffer_copy (of the texel buffer) with a launch of their own, the
 * next build by rewriting everything; only code that reads mips >= 6 of the texel buffer through its raw device pointer
 * BEFORE any of those must use rve_depth_pyramid_build instead. */
RVE_API int rve_depth_pyramid_build_deferred(rve_ctx* ctx, rve_pyramid pyr, rve_buffer depth, uint32_t width, uint32_t height);
/* The library keeps a derived copy of the chain for rve_cull (the min over each 2x2 sampler footprint, one load per
 * sample). It follows rve_depth_pyramid_build and rve_buffer_upload/copy/fill_zero on the texel buffer automatically;
 * call this after writing the texels any other way (own kernel through rve_buffer_device_ptr). */
RVE_API int rve_depth_pyramid_texels_changed(rve_ctx* ctx, rve_pyramid pyr);
/* The derived copy itself, for inspection (tests, debugging): per level of n x n texels an (n+1)^2 image
 * F[j][i] = min over texels {i-1,i} x {j-1,j} clamped to the level, levels back to back (level L starts at
 * 4*(dim^2 - n^2)/3 + 4*(dim - n) + L floats). No reference counterpart: the shader samples the mip chain with a
 * min-reduction sampler (src/RenderEngine.cpp:323-331), this buffer holds what each such sample returns. */
RVE_API int rve_depth_pyramid_footmin(rve_ctx* ctx, rve_pyramid pyr, rve_buffer* footmin);

/* ------------------------------------------------------------------ */
/* a-2 cull + LOD + indirect compaction — shaders/defaultcull.csh      */
/* ------------------------------------------------------------------ */
/* Replaces the encode at src/RenderEngine_Draw.cpp:1380-1405 (static) and
 * :1186-1215 (skinned): bindings Cubo=0, modelMatrix=1, renderLayer=2,
 * perObject=7, depthPyramid=8 (+sampler 9), bindless sets 3-6
 * (include/RavEngine/RenderEngine.hpp:95-102), push constants = CullingUBO,
 * DispatchCompute(num_groups_x,1,1, 64,1,1): num_groups_x*64 invocations, those
 * past sum(cubo.numObjects) exit (defaultcull.csh:150-164).
 * Indirect buffers must have been initialised by the caller (instanceCount=0),
 * as the reference does with CopyBufferToBuffer from its staging table.
 * The CUBO array may be passed as the reference builds it: `numCubos` counts LODs while one entry per command is written
 * (src/RenderEngine_Draw.cpp:1238), so its tail is stale staging memory. Entries that cannot be selected (numObjects 0, or
 * starting past the dispatch) are never resolved; a selectable entry whose handles do not name live buffers is dropped
 * (treated as numObjects 0) and counted - rve_ctx_bad_cubo_handles() - instead of being followed. */
RVE_API int rve_cull(rve_ctx* ctx, const RveCullingUBO* ubo, rve_buffer cubos,
                     rve_buffer model_matrices, rve_buffer render_layers,
                     rve_buffer per_object_flags, rve_pyramid pyramid,
                     uint32_t num_groups_x);

/* List maintenance for rve_cull: writes the ids of `entity_ids[0..count)` to `sorted_out` in ascending order. The cull maps
 * invocation -> list position -> entity id (shaders/defaultcull.csh:150-170) and appends in arrival order, so - outside
 * single-instance mode, where the position is the output slot - its result does not depend on the order of a command's
 * list; the engine's lists are swap-remove dense arrays (include/RavEngine/SparseSet.hpp:34-47) and lose their runs as
 * entities come and go. A CUBO may name the sorted copy as its entityIDInputBufferBindlessHandle; rebuild it when the list
 * changed. Ids must be unique and < id_limit (the entity id space, e.g. the element count of worldTransforms); ids past
 * the limit are dropped. Bitmap pass (set bits; block counts + scan + expand in one look-back pass): two launches, no
 * comparison sort. */
RVE_API int rve_sort_entity_ids(rve_ctx* ctx, rve_buffer entity_ids, uint32_t count, uint32_t id_limit, rve_buffer sorted_out);
/* The same for n_lists lists of one id space (the commands of a world whose lists changed this frame) with TWO launches
 * per 16 lists instead of two per list: one sets the bits of every list, one counts, scans (decoupled look-back) and expands. */
RVE_API int rve_sort_entity_ids_batch(rve_ctx* ctx, uint32_t n_lists, const rve_buffer* entity_ids, const uint32_t* counts, uint32_t id_limit,
                                      const rve_buffer* sorted_out);

/* rve_sort_entity_ids without any allocation. This entry point has no counterpart in the reference; it was added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving
 * loops, harnesses that forbid it). Same arguments, same result as rve_sort_entity_ids, but the scratch (bitmap, block
 * sums, out-of-range counter) is a fixed file-scope device array in csrc/sort_ids.cu sized for id spaces of up to
 * RVE_SORT_STATIC_MAX_IDS ids (1 MiB + 1 KiB, held by every GPU that loads the library until the process exits). The call
 * makes no cudaMalloc / cudaFree / allocating library call: it enqueues one memset and four kernels on the context's
 * stream (the first call on a device also resolves the array's device address with cudaGetSymbolAddress - make one call
 * before starting a capture). id_limit > RVE_SORT_STATIC_MAX_IDS returns RVE_ERR_UNSUPPORTED with nothing enqueued (use
 * rve_sort_entity_ids). All calls on a device share the one array, from every context and stream: it is cleared at the start
 * of each call, and calls that overlap on the GPU (different streams, or graphs replayed concurrently) race - serialise
 * them. */
#define RVE_SORT_STATIC_MAX_IDS 8388608u
RVE_API int rve_sort_entity_ids_static(rve_ctx* ctx, rve_buffer entity_ids, uint32_t count, uint32_t id_limit, rve_buffer sorted_out);

/* View-batched cull (SURVEY 8e partition B, 8f-2): the reference runs cullTheRenderData once per view of a frame
 * (main camera, every shadow cascade / spot / point face: src/RenderEngine_Draw.cpp:1769-1866, 1875-1987), each
 * pass re-reading every transform. Here one launch reads each instance ONCE and tests it against all `num_views`
 * views (<= 128): ubos[v] / pyramids[v] are what the v-th rve_cull call would take (same numCubos everywhere, no
 * single-instance mode). View v appends into the same buffers the CUBOs name, displaced by v * strides[c]:
 * indirect slot + v * indirect_stride commands, culling-buffer index + v * culling_stride slots, per CUBO c (host
 * array of numCubos entries; normally the material's numLODs and numEntities * numLODs). The caller sizes and
 * initialises the buffers for num_views views. Results are bit-identical to num_views separate rve_cull calls. */
typedef struct RveViewStride {
    uint32_t indirect_stride; /* IndirectIndexedCommands between consecutive views */
    uint32_t culling_stride;  /* u32 id slots between consecutive views */
} RveViewStride;
RVE_API int rve_cull_views(rve_ctx* ctx, const RveCullingUBO* ubos, uint32_t num_views, const rve_pyramid* pyramids,
                           const RveViewStride* strides, rve_buffer cubos, rve_buffer model_matrices,
                           rve_buffer render_layers, rve_buffer per_object_flags, uint32_t num_groups_x);
/* rve_cull_views with the view-projection matrices taken from DEVICE memory: view v uses the 16 floats (column-major)
 * at view_proj_matrices[matrix_index[v] * 16]; ubos[v].viewProj is ignored, every other field of ubos[v] applies.
 * Lets the shadow views of a frame consume rve_csm_cascades' output (the lightViewProj the reference uploads in
 * src/RenderEngine_Draw.cpp:282 and reads back per shadow pass at :1769-1866) without a host round trip; the frustum
 * planes are derived on the device with the same arithmetic as on the host (bit-identical results). */
RVE_API int rve_cull_views_matrices(rve_ctx* ctx, const RveCullingUBO* ubos, uint32_t num_views, rve_buffer view_proj_matrices,
                                    const uint32_t* matrix_index, const rve_pyramid* pyramids, const RveViewStride* strides,
                                    rve_buffer cubos, rve_buffer model_matrices, rve_buffer render_layers,
                                    rve_buffer per_object_flags, uint32_t num_groups_x);

/* ------------------------------------------------------------------ */
/* a-3 skinning — shaders/skinning_cs.csh                              */
/* ------------------------------------------------------------------ */

/* Replaces src/RenderEngine_Draw.cpp:577-630: out bindings 0-4
 * (pos,normal,tangent,bitangent vec3; uv vec2), in bindings 10-14, pose 20,
 * weights 21; DispatchCompute(ceil(nObj/8), ceil(nVerts/32),1, 8,32,1). */
RVE_API int rve_skin(rve_ctx* ctx, const RveSkinningUBO* ubo,
                     const rve_buffer out_streams[5], const rve_buffer in_streams[5],
                     rve_buffer pose, rve_buffer weights);

/* ------------------------------------------------------------------ */
/* a-4 skinned draw-call prep — shaders/skinned_mesh_drawcall.csh      */
/* ------------------------------------------------------------------ */
/* Replaces src/RenderEngine_Draw.cpp:546-574 (binding 0 = indirect buffer). */
RVE_API int rve_skinned_drawcall_prepare(rve_ctx* ctx, const RveSkinningPrepareUBO* ubo, rve_buffer indirect);

/* ------------------------------------------------------------------ */
/* a-5 / a-6 clustered lights                                          */
/* ------------------------------------------------------------------ */
/* cluster_build_grid.csh, encode src/RenderEngine_Draw.cpp:935-950:
 * DispatchCompute(gridSize.x, gridSize.y, gridSize.z, 1,1,1). */
RVE_API int rve_cluster_build(rve_ctx* ctx, const RveGridBuildUBO* ubo, rve_buffer clusters);
/* cluster_assign_lights.csh, encode src/RenderEngine_Draw.cpp:953-970:
 * DispatchCompute(numClusters/128,1,1, 128,1,1); bindings 0 clusters,
 * 1 point lights, 2 spot lights. num_clusters = invocations (groups*128). */
RVE_API int rve_cluster_assign(rve_ctx* ctx, const RveGridAssignUBO* ubo, rve_buffer clusters,
                               rve_buffer point_lights, rve_buffer spot_lights, uint32_t num_clusters);

/* ------------------------------------------------------------------ */
/* f-3 cascaded-shadow-map matrices                                    */
/* ------------------------------------------------------------------ */
/* One camera of the frame as the cascade fit sees it (src/RenderEngine_Draw.cpp:161-176): viewOnly is the
 * column-major view matrix, tanHalfFovy = tan(deg_to_rad(FOV * 1.5) / 2) computed by the caller (:219 widens the
 * fov by 1.5), aspect = float(targetWidth / targetHeight), the reference's INTEGER ratio (:219). */
typedef struct RveCsmCamera {
    float viewOnly[16];
    float zNear, zFar;
    float tanHalfFovy, aspect;
} RveCsmCamera;
/* One directional light (World::DirLightUploadData + its shadow settings, src/RenderEngine_Draw.cpp:178-180,
 * 205-216): shadowCascades are the split fractions in [0,1], numCascades <= 4. */
typedef struct RveCsmLight {
    float direction[3];
    uint32_t castsShadows;
    uint32_t numCascades;
    float shadowCascades[4];
    float maxShadowDistanceFactor;
} RveCsmLight;
/* Host helper (no device work): tanHalfFovy and aspect exactly as src/RenderEngine_Draw.cpp:220 and glm's
 * perspectiveRH_ZO derive them (libm tanf). A zero target height yields aspect 0 instead of the reference's
 * division fault. */
RVE_API void rve_csm_camera_params(float fov_degrees, uint32_t target_width, uint32_t target_height, float* tan_half_fovy,
                                   float* aspect);
/* Replaces the CPU loop of src/RenderEngine_Draw.cpp:149-293: for every camera c, shadow-casting light l and
 * cascade k < numCascades writes the light view-projection (and optionally the view and the projection) as 16
 * column-major floats at matrix index (c + l) * 4 + k — the reference's passIndex + i slot, collisions included —
 * and the cascade far distances at out_distances[l * 4 + k] (last camera wins, as light.cascadeDistances does).
 * out_view / out_proj / out_distances may be RVE_BUFFER_INVALID. f32, glm's operation order, no contraction. */
RVE_API int rve_csm_cascades(rve_ctx* ctx, rve_buffer cameras, uint32_t num_cameras, rve_buffer lights, uint32_t num_lights,
                             rve_buffer out_view_proj, rve_buffer out_view, rve_buffer out_proj, rve_buffer out_distances);

/* ------------------------------------------------------------------ */
/* 8(e) multi-GPU helpers (instance-range sharding)                    */
/* ------------------------------------------------------------------ */
/* Scatter of a packed dirty list (the device half of the dirty-range upload that replaces the reference's
 * per-range CopyBufferToBuffer, src/BufferedVRAMVector.cpp:87-104): staging holds `count` rows of elem_bytes at
 * rows_offset and their u32 destination row indices at indices_offset; dst[indices[i]] = rows[i]. */
RVE_API int rve_scatter_rows(rve_ctx* ctx, rve_buffer dst, uint32_t elem_bytes, rve_buffer staging, uint64_t indices_offset,
                             uint64_t rows_offset, uint32_t count);

/* After every rank culled its shard into rank-local slices and the per-slot
 * counts were all-gathered (NCCL, done by the host plumbing), compute on the
 * device: offsets[r][s] = sum_{q<r} counts[q][s] and total[s] = sum_q counts[q][s].
 * counts_all: u32[world][num_slots]; offsets_out: u32[world][num_slots];
 * totals_out: u32[num_slots]. */
RVE_API int rve_shard_offsets(rve_ctx* ctx, rve_buffer counts_all, uint32_t world, uint32_t num_slots,
                              rve_buffer offsets_out, rve_buffer totals_out);
/* Gather this rank's instanceCount fields (stride 20 B) of `num_slots`
 * commands into a dense u32 array (the all-gather payload). */
RVE_API int rve_gather_instance_counts(rve_ctx* ctx, rve_buffer indirect, uint32_t first_slot, uint32_t num_slots,
                                       rve_buffer counts_out, size_t counts_offset_bytes);

/* Same for several indirect buffers in ONE launch (one per material): slots of buffer i land after those of
 * buffers < i. n_buffers <= 64 per call. */
RVE_API int rve_gather_instance_counts_multi(rve_ctx* ctx, const rve_buffer* indirect, const uint32_t* first_slot,
                                             const uint32_t* num_slots, uint32_t n_buffers, rve_buffer counts_out,
                                             size_t counts_offset_bytes);

/* The exchange step as ONE kernel over NVLink peer memory (no NCCL call): every rank owns a mailbox buffer of
 * rve_count_mailbox_bytes(world, slots) bytes, zero-filled once, created with rve_buffer_create and made visible
 * to its peers with rve_ipc_export / rve_ipc_import (CUDA IPC; one process per GPU). The kernel takes this rank's
 * counts (from `n_buffers` indirect buffers like rve_gather_instance_counts_multi, or - n_buffers == 0 - from the
 * dense u32[dense_slots] array `dense_counts`), stores them into every rank's mailbox, signals, waits for every
 * peer's signal of the same `epoch` (1, 2, 3, ... - one per exchange, the same on all ranks) and writes
 * offsets_out[world][slots] (exclusive scan over ranks) and totals_out[slots]. mailboxes[q] is rank q's mailbox
 * (mailboxes[rank] = the local one). status (u32): bit 0 is set if a peer did not arrive within ~2 s, bit 1 if a peer
 * announced a different slot count; offsets / totals then hold 0xFFFFFFFF instead of values derived from stale counts.
 * The mailbox geometry (64 header words, then counts) does not depend on the slot count of a call. world <= 8. */
RVE_API size_t rve_count_mailbox_bytes(uint32_t world, uint32_t total_slots);
RVE_API int rve_ipc_export(rve_ctx* ctx, rve_buffer buf, unsigned char out_handle[64]);
RVE_API int rve_ipc_import(rve_ctx* ctx, const unsigned char handle[64], size_t bytes, rve_buffer* out);
RVE_API int rve_exchange_counts_p2p(rve_ctx* ctx, const rve_buffer* indirect, const uint32_t* first_slot,
                                    const uint32_t* num_slots, uint32_t n_buffers, rve_buffer dense_counts,
                                    uint32_t dense_slots, const rve_buffer* mailboxes, uint32_t rank, uint32_t world,
                                    uint32_t epoch, rve_buffer offsets_out, rve_buffer totals_out, rve_buffer status);

/* Device-side gather of the compacted ids for a single consumer (north_star: "all-gathered over NVLink only when a
 * single consumer needs them"): after rve_exchange_counts_p2p of the same `epoch` produced offsets / totals, every rank
 * writes its visible ids straight into the consumers' GLOBAL culling buffers over the mapped peer memory, at the slice
 * the single-GPU cull would have used - cubo.cullingBufferOffset + numObjects * lod of the GLOBAL command
 * (shaders/defaultcull.csh:305-312; consumed by the draw at src/RenderEngine_Draw.cpp:1537-1545) - displaced by
 * offsets[rank][slot]; order inside a slice is rank order, then this rank's arrival order. Enqueue it behind the
 * exchange on the same stream. Per slot s (<= 64): slot_buffer[s] indexes src_culling (<= 16 local culling buffers),
 * src_start[s] / dst_start[s] are the first id slots of the local / global slice. dst_culling[q * n_buffers + b] is
 * rank q's global buffer b, mapped with rve_ipc_import (own rank: the local handle); entries of ranks outside the
 * `consumers` bit mask are ignored. id_base is added to every id (local -> global entity ids). A consumer's kernel
 * returns when every rank's ids have landed; status bit 0 = timeout. NVLink traffic: 4 B x visible per remote consumer. */
RVE_API int rve_gather_ids_p2p(rve_ctx* ctx, const rve_buffer* src_culling, uint32_t n_buffers, const uint32_t* slot_buffer,
                               const uint32_t* src_start, const uint32_t* dst_start, uint32_t slots, const rve_buffer* dst_culling,
                               uint32_t consumers, uint32_t id_base, rve_buffer offsets, rve_buffer totals, const rve_buffer* mailboxes,
                               uint32_t rank, uint32_t world, uint32_t epoch, rve_buffer status);

/* ------------------------------------------------------------------ */
/* synthetic scene generation on the device (bench / C5)  */
/* ------------------------------------------------------------------ */
/* Fills transforms (T*R*S), layers, flags and identity ids for `count`
 * instances whose global indices start at `first_index`, procedurally from
 * (seed, global index) so a 1 B-instance scene needs no host->device copy.
 * The oracle has the same generator (oracle: orc_gen_instances). */
typedef struct RveSceneGenParams {
    uint64_t seed;
    float extent_xy;     /* positions uniform in [-extent_xy, extent_xy]^2 */
    float extent_z;      /* ... x [-extent_z, extent_z] (engine Y up: y uses extent_z) */
    float scale_min, scale_max;
    float frac_layer_zero;   /* fraction of instances with layer mask 0 */
    float frac_no_cull_bits; /* fraction with frustum+occlusion bits cleared */
    uint32_t id_base;        /* entity id written = id_base + local index */
} RveSceneGenParams;
RVE_API int rve_gen_instances(rve_ctx* ctx, const RveSceneGenParams* p, uint64_t first_index, uint32_t count,
                              rve_buffer model_matrices, rve_buffer render_layers, rve_buffer per_object_flags,
                              rve_buffer entity_ids);

#ifdef __cplusplus
}
#endif
#endif /* RVE_B200_H */
