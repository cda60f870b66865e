// frame_internal.h — World::RenderData stand-in used by the shim (frame.cu, frame_cull.cu, frame_skinned.cu)
#pragma once
#include <cstdint>
#include <utility>
#include <vector>

#include "../../include/rve_shim.h"

namespace rve {
// change log of a sparse set (csrc/sync.cu): off until switched on; take_changes returns the net ids that entered / left the
// set since the last call and restarts the log - false when the log was off or overflowed (the consumer rebuilds from the list)
uint64_t sparse_set_version(rvs_sparse_set* s);  // changes with every Emplace / Erase that altered the set
// (`owner`: the one consumer the log is kept for - a second one is refused and keeps rebuilding)
bool sparse_set_track_changes(rvs_sparse_set* s, uint64_t owner);
bool sparse_set_take_changes(rvs_sparse_set* s, uint64_t owner, std::vector<uint32_t>* entered, std::vector<uint32_t>* left);
}  // namespace rve

namespace rvs {

struct LODMesh {
    uint32_t indexCount, indexStart, baseVertex;
};

// MeshCollectionStatic (include/RavEngine/MeshCollection.hpp:24-59)
struct MeshCollection {
    float radius = 0;
    std::vector<LODMesh> lods;
    rve_buffer lodDistances = RVE_BUFFER_INVALID;
    std::vector<float> lodDistancesHost;  // the same values, for the host-built cull tables
};

// MDIICommand::command (include/RavEngine/World.hpp:339-349)
struct Command {
    uint32_t mesh;
    rve_buffer entityIds;
    uint32_t count;
    rvs_sparse_set* set = nullptr;  // when set: entityIds / count follow the set (synced before every cull)
    // RVS_SORT_ENTITY_LISTS: ascending device copy of the list, rebuilt when the list changed; the cull reads it instead
    rve_buffer sortedIds = RVE_BUFFER_INVALID;
    size_t sortedBytes = 0;
    rve_buffer sortedOf = RVE_BUFFER_INVALID;
    uint32_t sortedCount = 0;
    bool sortedValid = false;
    // ... for a sparse-set-backed list the copy is maintained through a bitmap the command owns: after a small change only
    // the ids that entered / left the set are applied to it and the copy is re-expanded (csrc/sort_ids.cu, sort_lists_persistent)
    rve_buffer sortBitmap = RVE_BUFFER_INVALID;
    size_t sortBitmapBytes = 0;
    uint32_t bitmapIdLimit = 0;
    bool bitmapValid = false;
    uint64_t sortOwnerId = 0;  // this command's identity towards the set's change log
    uint64_t sortedVersion = ~0ull;  // version of the set the copy was made from
};

// MDICommandBase + MDIICommand (World.hpp:334-349)
struct Material {
    uint32_t kind = RVS_MAT_LIT_OPAQUE;
    std::vector<Command> commands;
    rve_buffer cullingBuffer = RVE_BUFFER_INVALID, indirectBuffer = RVE_BUFFER_INVALID;
    size_t cullingBytes = 0, indirectBytes = 0;
    uint32_t numLODs = 0, numEntities = 0;
};

struct SkinnedMesh {
    uint32_t numVerts, numIndices, indexStart, vertexStart;
    float radius;
    rve_buffer weights;
    rve_buffer lodDistances;  // single +inf entry (MeshCollection.cpp:24)
};

// MDIICommandSkinned::command (World.hpp:351-362)
struct SkinnedCommand {
    uint32_t mesh, numBones;
    rve_buffer entityIds;
    uint32_t count;
    const float* palettes;
};

struct SkinnedMaterial {
    uint32_t kind = RVS_MAT_LIT_OPAQUE;
    std::vector<SkinnedCommand> commands;
    rve_buffer cullingBuffer = RVE_BUFFER_INVALID, indirectBuffer = RVE_BUFFER_INVALID, cuboBuffer = RVE_BUFFER_INVALID;
    size_t cullingCount = 0, indirectCount = 0, cuboBytes = 0;
    uint32_t numEntities = 0;
};

}  // namespace rvs

struct rvs_world {
    rve_ctx* ctx = nullptr;
    uint32_t flags = 0;
    rve_buffer model = RVE_BUFFER_INVALID, layers = RVE_BUFFER_INVALID, flagsBuf = RVE_BUFFER_INVALID;
    std::vector<rvs::MeshCollection> meshes;
    std::vector<rvs::Material> materials;
    // renderData.cuboBuffer (World.hpp:447)
    rve_buffer cuboBuffer = RVE_BUFFER_INVALID;
    size_t cuboBytes = 0;
    uint32_t lastNumCubos = 0, lastTotalEntities = 0;
    uint32_t sortFullRebuilds = 0, sortDeltaUpdates = 0;  // rvs_sorted_list_stats
    // lightClusterBuffer (src/RenderEngine.cpp:839-845)
    rve_buffer clusterBuffer = RVE_BUFFER_INVALID;
    size_t clusterBytes = 0;
    // rvs_frame_encode: one cluster buffer per camera of the frame (buffer, bytes)
    std::vector<std::pair<rve_buffer, size_t>> frameClusters;
    uint32_t frameViews = 0;
    // skinned path
    std::vector<rvs::SkinnedMesh> skinnedMeshes;
    std::vector<rvs::SkinnedMaterial> skinnedMaterials;
    rve_buffer arena[5] = {RVE_BUFFER_INVALID, RVE_BUFFER_INVALID, RVE_BUFFER_INVALID, RVE_BUFFER_INVALID, RVE_BUFFER_INVALID};
    rve_buffer skinnedOut[5] = {RVE_BUFFER_INVALID, RVE_BUFFER_INVALID, RVE_BUFFER_INVALID, RVE_BUFFER_INVALID, RVE_BUFFER_INVALID};
    size_t skinnedOutCount = 0;  // vertices each shared skinned stream can hold
    rve_buffer poseBuffer = RVE_BUFFER_INVALID;
    size_t poseCount = 0;  // matrices
    // scratch host staging reused across frames (stagingBufferPool stand-in)
    std::vector<unsigned char> staging;
};

namespace rvs {
int realloc_buffer(rvs_world* w, rve_buffer* buf, size_t* cur_bytes, size_t count, size_t stride);
}
