// frame.cu — host-side mirror of RenderEngine::Draw's encode blocks for static meshes and lights
// (include/rve_shim.h). Pure host C++; every device action goes through the rve_* C-ABI.
#include <cmath>
#include <cstring>

#include "../../include/rve_shim.h"
#include "frame_internal.h"
#include "internal.h"

using namespace rve;

namespace rvs {

// reallocBuffer (src/RenderEngine_Draw.cpp, used at :1250,1289-1291): recreate when the element
// count changes; old buffer goes to the deferred-release queue (here: stream-ordered destroy)
int realloc_buffer(rvs_world* w, rve_buffer* buf, size_t* cur_bytes, size_t count, size_t stride) {
    const size_t bytes = count * stride;
    if (*buf != RVE_BUFFER_INVALID && *cur_bytes == bytes) return 0;
    if (*buf != RVE_BUFFER_INVALID) {
        int rc = rve_buffer_destroy(w->ctx, *buf);
        if (rc) return rc;
        *buf = RVE_BUFFER_INVALID;
    }
    *cur_bytes = bytes;
    if (bytes == 0) return 0;
    return rve_buffer_create(w->ctx, bytes, buf);
}

}  // namespace rvs

using namespace rvs;

extern "C" {

int rvs_world_create(rve_ctx* ctx, uint32_t flags, rvs_world** out) {
    if (!ctx || !out) return set_error(RVE_ERR_INVALID_ARGUMENT, "null argument");
    auto* w = new rvs_world();
    w->ctx = ctx;
    w->flags = flags;
    *out = w;
    return 0;
}

int rvs_world_destroy(rvs_world* w) {
    if (!w) return 0;
    auto kill = [&](rve_buffer b) {
        if (b != RVE_BUFFER_INVALID) rve_buffer_destroy(w->ctx, b);
    };
    for (auto& m : w->meshes) kill(m.lodDistances);
    for (auto& m : w->materials) {
        kill(m.cullingBuffer);
        kill(m.indirectBuffer);
        for (auto& c : m.commands) {
            kill(c.sortedIds);
            kill(c.sortBitmap);
        }
    }
    for (auto& m : w->skinnedMaterials) {
        kill(m.cullingBuffer);
        kill(m.indirectBuffer);
        kill(m.cuboBuffer);
    }
    kill(w->cuboBuffer);
    kill(w->clusterBuffer);
    for (auto& c : w->frameClusters) kill(c.first);
    kill(w->poseBuffer);
    for (auto b : w->skinnedOut) kill(b);
    delete w;
    return 0;
}

int rvs_world_set_entity_buffers(rvs_world* w, rve_buffer model, rve_buffer layers, rve_buffer flags) {
    if (!w) return set_error(RVE_ERR_INVALID_ARGUMENT, "world is null");
    w->model = model;
    w->layers = layers;
    w->flagsBuf = flags;
    return 0;
}

int rvs_mesh_collection_create(rvs_world* w, float radius, uint32_t num_lods, const float* lod_distances,
                               const uint32_t* index_counts, const uint32_t* index_starts, const uint32_t* base_vertices,
                               uint32_t* out_mesh) {
    if (!w || !out_mesh || !lod_distances) return set_error(RVE_ERR_INVALID_ARGUMENT, "null argument");
    RVE_REQUIRE(num_lods > 0, "Mesh has no LODs!");  // Debug::Assert at :1281
    MeshCollection m;
    m.radius = radius;
    for (uint32_t i = 0; i < num_lods; ++i)
        m.lods.push_back({index_counts ? index_counts[i] : 0u, index_starts ? index_starts[i] : 0u,
                          base_vertices ? base_vertices[i] : 0u});
    int rc = rve_buffer_create(w->ctx, num_lods * sizeof(float), &m.lodDistances);
    if (rc) return rc;
    rc = rve_buffer_upload(w->ctx, m.lodDistances, 0, lod_distances, num_lods * sizeof(float));
    if (rc) return rc;
    m.lodDistancesHost.assign(lod_distances, lod_distances + num_lods);
    w->meshes.push_back(m);
    *out_mesh = uint32_t(w->meshes.size() - 1);
    return 0;
}

int rvs_material_create(rvs_world* w, uint32_t kind, uint32_t* out_material) {
    if (!w || !out_material) return set_error(RVE_ERR_INVALID_ARGUMENT, "null argument");
    Material m;
    m.kind = kind;
    w->materials.push_back(m);
    *out_material = uint32_t(w->materials.size() - 1);
    return 0;
}

int rvs_material_add_command(rvs_world* w, uint32_t material, uint32_t mesh, rve_buffer entity_ids, uint32_t entity_count) {
    if (!w) return set_error(RVE_ERR_INVALID_ARGUMENT, "world is null");
    RVE_REQUIRE(material < w->materials.size() && mesh < w->meshes.size(), "bad material / mesh index");
    w->materials[material].commands.push_back({mesh, entity_ids, entity_count});
    return 0;
}

int rvs_material_add_command_set(rvs_world* w, uint32_t material, uint32_t mesh, rvs_sparse_set* entities) {
    if (!w || !entities) return set_error(RVE_ERR_INVALID_ARGUMENT, "null argument");
    RVE_REQUIRE(material < w->materials.size() && mesh < w->meshes.size(), "bad material / mesh index");
    Command c{mesh, RVE_BUFFER_INVALID, 0};
    c.set = entities;
    w->materials[material].commands.push_back(c);
    return 0;
}

int rvs_material_set_command_entities(rvs_world* w, uint32_t material, uint32_t command_index, rve_buffer entity_ids,
                                      uint32_t entity_count) {
    if (!w) return set_error(RVE_ERR_INVALID_ARGUMENT, "world is null");
    RVE_REQUIRE(material < w->materials.size() && command_index < w->materials[material].commands.size(), "bad command index");
    auto& c = w->materials[material].commands[command_index];
    c.entityIds = entity_ids;
    c.count = entity_count;
    c.sortedValid = false;  // also the way to say "the list's contents changed" for a caller-owned buffer
    return 0;
}

int rvs_material_outputs(rvs_world* w, uint32_t material, rve_buffer* culling_buffer, rve_buffer* indirect_buffer,
                         uint32_t* num_lods_total, uint32_t* num_entities_total) {
    if (!w) return set_error(RVE_ERR_INVALID_ARGUMENT, "world is null");
    RVE_REQUIRE(material < w->materials.size(), "bad material index");
    const Material& m = w->materials[material];
    if (culling_buffer) *culling_buffer = m.cullingBuffer;
    if (indirect_buffer) *indirect_buffer = m.indirectBuffer;
    if (num_lods_total) *num_lods_total = m.numLODs;
    if (num_entities_total) *num_entities_total = m.numEntities;
    return 0;
}

int rvs_sorted_list_stats(rvs_world* w, uint32_t* full_rebuilds, uint32_t* delta_updates) {
    if (!w) return set_error(RVE_ERR_INVALID_ARGUMENT, "null world");
    if (full_rebuilds) *full_rebuilds = w->sortFullRebuilds;
    if (delta_updates) *delta_updates = w->sortDeltaUpdates;
    return 0;
}

int rvs_last_cull_info(rvs_world* w, rve_buffer* cubo_buffer, uint32_t* num_cubos, uint32_t* total_entities) {
    if (!w) return set_error(RVE_ERR_INVALID_ARGUMENT, "world is null");
    if (cubo_buffer) *cubo_buffer = w->cuboBuffer;
    if (num_cubos) *num_cubos = w->lastNumCubos;
    if (total_entities) *total_entities = w->lastTotalEntities;
    return 0;
}

}  // extern "C"
