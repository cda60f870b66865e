// frame_skinned.cu — the skinned-mesh encode blocks of RenderEngine::Draw restated over the C-ABI:
// prepareSkeletalMeshBuffers (:479-544), poseSkeletalMeshes (:577-630),
// prepareSkeletalCullingBuffer (:546-574), cullSkeletalMeshes (:1067-1218). Host C++ only.
#include <cmath>
#include <cstring>
#include <limits>

#include "frame_internal.h"
#include "internal.h"

using namespace rve;
using namespace rvs;

namespace {

// closest_power_of<uint32_t>(n, 2): next power of two >= n (0 stays 0)
uint64_t closest_pow2(uint64_t n) {
    if (n == 0) return 0;
    uint64_t p = 1;
    while (p < n) p <<= 1;
    return p;
}

// resizeSkeletonBuffer (:486-508): grow-only, to the next power of two elements
int resize_skeleton_buffer(rvs_world* w, rve_buffer* buf, size_t* count, size_t stride, uint64_t needed) {
    if (*buf != RVE_BUFFER_INVALID && *count >= needed) return 0;
    const uint64_t newCount = closest_pow2(needed);
    if (newCount == 0) return 0;
    if (*buf != RVE_BUFFER_INVALID) {
        int rc = rve_buffer_destroy(w->ctx, *buf);
        if (rc) return rc;
        *buf = RVE_BUFFER_INVALID;
    }
    *count = newCount;
    return rve_buffer_create(w->ctx, newCount * stride, buf);
}

}  // namespace

extern "C" {

int rvs_skinned_mesh_create(rvs_world* w, uint32_t num_verts, uint32_t num_indices, uint32_t index_start,
                            uint32_t vertex_start, float radius, rve_buffer weights, uint32_t* out_mesh) {
    if (!w || !out_mesh) return set_error(RVE_ERR_INVALID_ARGUMENT, "null argument");
    SkinnedMesh m{num_verts, num_indices, index_start, vertex_start, radius, weights, RVE_BUFFER_INVALID};
    const float inf = std::numeric_limits<float>::infinity();  // single LOD (src/MeshCollection.cpp:24)
    int rc = rve_buffer_create(w->ctx, sizeof(float), &m.lodDistances);
    if (rc) return rc;
    rc = rve_buffer_upload(w->ctx, m.lodDistances, 0, &inf, sizeof(float));
    if (rc) return rc;
    w->skinnedMeshes.push_back(m);
    *out_mesh = uint32_t(w->skinnedMeshes.size() - 1);
    return 0;
}

int rvs_skinned_material_create(rvs_world* w, uint32_t kind, uint32_t* out_material) {
    if (!w || !out_material) return set_error(RVE_ERR_INVALID_ARGUMENT, "null argument");
    SkinnedMaterial m;
    m.kind = kind;
    w->skinnedMaterials.push_back(m);
    *out_material = uint32_t(w->skinnedMaterials.size() - 1);
    return 0;
}

int rvs_skinned_material_add_command(rvs_world* w, uint32_t material, uint32_t mesh, uint32_t num_bones,
                                     rve_buffer entity_ids, uint32_t entity_count, const float* palettes) {
    if (!w) return set_error(RVE_ERR_INVALID_ARGUMENT, "world is null");
    RVE_REQUIRE(material < w->skinnedMaterials.size() && mesh < w->skinnedMeshes.size(), "bad material / mesh index");
    w->skinnedMaterials[material].commands.push_back({mesh, num_bones, entity_ids, entity_count, palettes});
    return 0;
}

int rvs_world_set_mesh_arena(rvs_world* w, const rve_buffer in_streams[5]) {
    if (!w || !in_streams) return set_error(RVE_ERR_INVALID_ARGUMENT, "null argument");
    for (int i = 0; i < 5; ++i) w->arena[i] = in_streams[i];
    return 0;
}

int rvs_prepare_skeletal_mesh_buffers(rvs_world* w, uint32_t* exist) {
    if (!w) return set_error(RVE_ERR_INVALID_ARGUMENT, "world is null");
    uint64_t totalVertsToSkin = 0, totalJointsToSkin = 0, totalObjectsToSkin = 0;
    for (auto& m : w->skinnedMaterials) {
        uint64_t totalEntitiesForThisCommand = 0;
        for (auto& c : m.commands) {
            totalObjectsToSkin += c.count;
            totalEntitiesForThisCommand += c.count;
            totalVertsToSkin += uint64_t(w->skinnedMeshes[c.mesh].numVerts) * c.count;
            totalJointsToSkin += uint64_t(c.numBones) * c.count;
        }
        m.numEntities = uint32_t(totalEntitiesForThisCommand);
        int rc = resize_skeleton_buffer(w, &m.indirectBuffer, &m.indirectCount, sizeof(RveIndirectIndexedCommand), totalEntitiesForThisCommand);
        if (rc) return rc;
        rc = resize_skeleton_buffer(w, &m.cullingBuffer, &m.cullingCount, sizeof(uint32_t), totalEntitiesForThisCommand);
        if (rc) return rc;
    }
    RVE_REQUIRE(totalVertsToSkin <= 0xFFFFFFFFull, "more than 2^32 skinned vertices (u32 offsets in SkinningUBO)");
    int rc = resize_skeleton_buffer(w, &w->poseBuffer, &w->poseCount, 64, totalJointsToSkin);
    if (rc) return rc;
    // the five shared skinned streams grow together (:536-540)
    if (totalVertsToSkin > w->skinnedOutCount || w->skinnedOut[0] == RVE_BUFFER_INVALID) {
        size_t cnt = 0;
        for (int i = 0; i < 5; ++i) {
            cnt = w->skinnedOutCount;
            rc = resize_skeleton_buffer(w, &w->skinnedOut[i], &cnt, i == 4 ? 8 : 12, totalVertsToSkin);
            if (rc) return rc;
        }
        w->skinnedOutCount = cnt;
    }
    if (exist) *exist = (totalObjectsToSkin > 0 && totalVertsToSkin > 0) ? 1u : 0u;
    return 0;
}

int rvs_pose_skeletal_meshes(rvs_world* w) {
    if (!w) return set_error(RVE_ERR_INVALID_ARGUMENT, "world is null");
    RveSkinningUBO subo{};
    for (auto& m : w->skinnedMaterials) {
        for (auto& c : m.commands) {
            const SkinnedMesh& mesh = w->skinnedMeshes[c.mesh];
            subo.numObjects = c.count;
            subo.numVertices = mesh.numVerts;
            subo.numBones = c.numBones;
            subo.vertexReadOffset = mesh.vertexStart;
            // :608-619 write joint matrices into the (mapped, in the reference) palette buffer
            const size_t bytes = size_t(c.count) * c.numBones * 64;
            if (bytes) {
                RVE_REQUIRE(c.palettes != nullptr, "skinned command has no palette pointer");
                int rc = rve_buffer_upload(w->ctx, w->poseBuffer, size_t(subo.boneReadOffset) * 64, c.palettes, bytes);
                if (rc) return rc;
            }
            // :621-622 DispatchCompute(ceil(nObj/8), ceil(nVerts/32), 1, 8, 32, 1)
            int rc = rve_skin(w->ctx, &subo, w->skinnedOut, w->arena, w->poseBuffer, mesh.weights);
            if (rc) return rc;
            subo.boneReadOffset += subo.numBones * subo.numObjects;
            subo.vertexWriteOffset += subo.numVertices * subo.numObjects;
        }
    }
    return 0;
}

int rvs_prepare_skeletal_culling_buffer(rvs_world* w) {
    if (!w) return set_error(RVE_ERR_INVALID_ARGUMENT, "world is null");
    for (auto& m : w->skinnedMaterials) {
        RveSkinningPrepareUBO ubo{};  // fresh per material (:551)
        for (auto& c : m.commands) {
            const SkinnedMesh& mesh = w->skinnedMeshes[c.mesh];
            ubo.nVerticesInThisMesh = mesh.numVerts;
            ubo.nTotalObjects = c.count;
            ubo.indexBufferOffset = mesh.indexStart;
            ubo.nIndicesInThisMesh = mesh.numIndices;
            int rc = rve_skinned_drawcall_prepare(w->ctx, &ubo, m.indirectBuffer);
            if (rc) return rc;
            ubo.vertexBufferOffset += mesh.numVerts;
            ubo.drawCallBufferOffset += c.count;
            ubo.baseInstanceOffset += c.count;
        }
    }
    return 0;
}

int rvs_cull_skeletal(rvs_world* w, const float view_proj[16], const float cam_pos[3], uint32_t camera_layers,
                      uint32_t shadow_pass, uint32_t pass_mask, rve_pyramid pyramid) {
    RVE_PROFILE_FN_N("Cull Skeletal Meshes");
    if (!w || !view_proj || !cam_pos) return set_error(RVE_ERR_INVALID_ARGUMENT, "null argument");
    rve_ctx* ctx = w->ctx;
    // :1069-1123 reset the indirect buffers from a host staging table
    uint32_t skeletalVertexOffset = 0;
    std::vector<RveIndirectIndexedCommand> staging;
    for (auto& m : w->skinnedMaterials) {
        if (!(m.kind & pass_mask)) continue;
        uint32_t total_entities = 0;
        for (auto& c : m.commands) total_entities += c.count;
        staging.assign(total_entities, RveIndirectIndexedCommand{0, 0, 0, 0, 0});
        for (auto& c : m.commands) {
            const uint32_t meshID = 0;  // declared inside the per-command loop in the reference (:1088)
            const SkinnedMesh& mesh = w->skinnedMeshes[c.mesh];
            for (uint32_t i = 0; i < c.count; i++) {
                const RveIndirectIndexedCommand initData{mesh.numIndices, 0, mesh.indexStart, skeletalVertexOffset, i};
                const uint32_t slot = meshID * 1 + 0 + i;  // (:1103) every command starts again at slot 0
                if (slot < total_entities) staging[slot] = initData;
                skeletalVertexOffset += mesh.numVerts;
            }
        }
        if (total_entities) {
            int rc = rve_buffer_upload(ctx, m.indirectBuffer, 0, staging.data(), total_entities * sizeof(RveIndirectIndexedCommand));
            if (rc) return rc;
        }
    }
    // :1127-1216 one dispatch per material
    for (auto& m : w->skinnedMaterials) {
        if (!(m.kind & pass_mask)) continue;
        RveCullingUBO globalCubo{};
        std::memcpy(globalCubo.viewProj, view_proj, 64);
        std::memcpy(globalCubo.camPos, cam_pos, 12);
        globalCubo.cameraRenderLayers = camera_layers;
        globalCubo.singleInstanceModeAndShadowMode = RVE_CULL_SINGLE_INSTANCE | (shadow_pass ? RVE_CULL_SHADOW_MODE : 0u);
        globalCubo.numCubos = uint32_t(m.commands.size());  // += GetNumLods() == 1 per command
        if (globalCubo.numCubos == 0) continue;
        int rc = realloc_buffer(w, &m.cuboBuffer, &m.cuboBytes, globalCubo.numCubos, sizeof(RveCullingUBOinstance));
        if (rc) return rc;
        std::vector<RveCullingUBOinstance> cubos(globalCubo.numCubos);
        RveCullingUBOinstance cubo{};
        cubo.numLODs = 1;
        uint32_t i = 0;
        for (auto& c : m.commands) {
            const SkinnedMesh& mesh = w->skinnedMeshes[c.mesh];
            const uint32_t lodsForThisMesh = 1;
            cubo.numObjects = c.count;
            cubo.radius = mesh.radius;
            cubo.idOutputBufferBindlessHandle = rve_buffer_bindless_index(ctx, m.cullingBuffer);
            cubo.indirectOutputBufferBindlessHandle = rve_buffer_bindless_index(ctx, m.indirectBuffer);
            cubo.lodDistanceBufferBindlessHandle = rve_buffer_bindless_index(ctx, mesh.lodDistances);
            cubo.entityIDInputBufferBindlessHandle = rve_buffer_bindless_index(ctx, c.entityIds);
            cubos[i++] = cubo;
            cubo.indirectBufferOffset += lodsForThisMesh;
            cubo.cullingBufferOffset += lodsForThisMesh * c.count;
        }
        rc = rve_buffer_upload(ctx, m.cuboBuffer, 0, cubos.data(), cubos.size() * sizeof(RveCullingUBOinstance));
        if (rc) return rc;
        // :1214 the dispatch is sized by the LAST command's numObjects only
        const uint32_t groups = uint32_t(std::ceil(cubo.numObjects / 64.f));
        rc = rve_cull(ctx, &globalCubo, m.cuboBuffer, w->model, w->layers, w->flagsBuf, pyramid, groups);
        if (rc) return rc;
    }
    return 0;
}

int rvs_skinned_outputs(rvs_world* w, rve_buffer out_streams[5], rve_buffer* pose_buffer) {
    if (!w) return set_error(RVE_ERR_INVALID_ARGUMENT, "world is null");
    if (out_streams)
        for (int i = 0; i < 5; ++i) out_streams[i] = w->skinnedOut[i];
    if (pose_buffer) *pose_buffer = w->poseBuffer;
    return 0;
}

int rvs_skinned_material_outputs(rvs_world* w, uint32_t material, rve_buffer* culling_buffer, rve_buffer* indirect_buffer,
                                 uint32_t* num_entities_total) {
    if (!w) return set_error(RVE_ERR_INVALID_ARGUMENT, "world is null");
    RVE_REQUIRE(material < w->skinnedMaterials.size(), "bad material index");
    const SkinnedMaterial& m = w->skinnedMaterials[material];
    if (culling_buffer) *culling_buffer = m.cullingBuffer;
    if (indirect_buffer) *indirect_buffer = m.indirectBuffer;
    if (num_entities_total) *num_entities_total = m.numEntities;
    return 0;
}

}  // extern "C"
