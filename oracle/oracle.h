/*
 * oracle.h — C interface of the CPU oracle.
 *
 * TEST INFRASTRUCTURE ONLY. This is a CPU restatement of the reference's GLSL
 * compute shaders for the render-prep path; it exists to check the CUDA
 * kernels and to be timed as the CPU baseline. Only tests/, the smoke test
 * and bench.py's cpu_baseline / --impl reference leg may load it. The product
 * (ravengine_b200/) never links or calls it.
 *
 * PARITY UNPINNED: the reference ships no test, golden vector or fixture for
 * this path and its own implementation only runs through a
 * graphics API that does not exist on these machines, so the oracle is pinned
 * by hand-computed known-answer cases (tests/test_oracle_known_answers.py) and
 * committed golden fixtures generated from it (tests/golden/), not by
 * reference outputs.
 */
#ifndef RVE_ORACLE_H
#define RVE_ORACLE_H

#include "../include/rve_b200.h"

#ifdef __cplusplus
extern "C" {
#endif

/* DepthPyramid::DepthPyramid — src/DepthPyramid.cpp:10-17 */
void orc_pyramid_dims(uint32_t width, uint32_t* dim, uint32_t* num_levels);
uint64_t orc_pyramid_texel_count(uint32_t dim);
uint64_t orc_pyramid_level_offset(uint32_t dim, uint32_t level);
/* depthpyramidcopy.fsh:10-20 + depthpyramid.csh:9-28 */
void orc_pyramid_build(const float* depth, uint32_t width, uint32_t height, float* pyramid, uint32_t dim,
                       uint32_t num_levels);
void orc_pyramid_build_mt(const float* depth, uint32_t width, uint32_t height, float* pyramid, uint32_t dim, uint32_t num_levels,
                          int nthreads);
/* the min-reduction sampler model (src/RenderEngine.cpp:323-331), exposed for unit tests */
float orc_sample_min(const float* image, uint32_t w, uint32_t h, float u, float v);
float orc_ceil_log2(float v);

/* defaultcull.csh:148-313. `bindless` maps handle -> host pointer. nthreads<=1:
 * serial, invocation order (deterministic append). nthreads>1: 64-invocation
 * chunks on worker threads with an atomic append, like the shader. */
void orc_cull(const RveCullingUBO* ubo, const RveCullingUBOinstance* cubos, const float* model_matrices,
              const uint32_t* render_layers, const uint16_t* per_object_flags, const float* pyramid,
              uint32_t pyr_dim, uint32_t pyr_levels, void* const* bindless, uint32_t num_groups_x, int nthreads);

/* skinning_cs.csh:95-138 */
void orc_skin(const RveSkinningUBO* ubo, float* out_pos, float* out_nrm, float* out_tan, float* out_bitan,
              float* out_uv, const float* in_pos, const float* in_nrm, const float* in_tan, const float* in_bitan,
              const float* in_uv, const float* pose, const RveVertexJointBinding* weights, int nthreads);
/* same arithmetic evaluated in double (tolerance reference for the 1e-5 bar) */
void orc_skin_f64(const RveSkinningUBO* ubo, double* out_pos, double* out_nrm, double* out_tan, double* out_bitan,
                  const float* in_pos, const float* in_nrm, const float* in_tan, const float* in_bitan,
                  const float* pose, const RveVertexJointBinding* weights);

/* skinned_mesh_drawcall.csh:25-39 */
void orc_skinned_drawcall_prepare(const RveSkinningPrepareUBO* ubo, RveIndirectIndexedCommand* commands);

/* cluster_build_grid.csh:16-79 */
void orc_cluster_build(const RveGridBuildUBO* ubo, RveCluster* clusters);
/* cluster_assign_lights.csh:28-88 (+ cluster_shared.glsl:41-44) */
void orc_cluster_assign(const RveGridAssignUBO* ubo, RveCluster* clusters, const RvePointLight* points,
                        const RveSpotLight* spots, uint32_t num_clusters, int nthreads);

/* synthetic instance generator, bit-identical to the device one (rve_gen_instances) */
void orc_gen_instances(const RveSceneGenParams* p, uint64_t first_index, uint32_t count, float* model_matrices,
                       uint32_t* render_layers, uint16_t* per_object_flags, uint32_t* entity_ids);

/* worker count the engine would use: max(hardware_concurrency-2, 2),
 * include/RavEngine/App.hpp:112-118 */
int orc_default_workers(void);
/* 1 when built against the reference's vendored taskflow (oracle/_ref build) */
int orc_uses_taskflow(void);

#ifdef __cplusplus
}
#endif
#endif
