// mesh_arena.cu — SURVEY 8f-4: the `.rvem` mesh / weights container and the shared vertex arena, i.e. the data formats
// on the input side of skinning and the source of every indexStart / baseVertex / vertexReadOffset of the path.
//
//   .rvem           include/RavEngine/Mesh.hpp:12-27 (16-byte header: 'rvem', numVertices, numIndicies, attribute bits),
//                   reader src/MeshAsset.cpp:23-79, writer tools/rvemc/rvemc.cpp:246-287: SoA streams in the order
//                   positions, normals, tangents, bitangents (vec3), uv0, lightmapUV (vec2) - each only when its bit is
//                   set - then u32 indices, then (skinned) VertexWeights[numVertices] (4 x {u32 joint, f32 influence},
//                   include/RavEngine/Skeleton.hpp:50-60, read by src/MeshAssetSkinned.cpp:38-56).
//   arena           src/RenderEngine_Allocate.cpp:12-100 (AllocateMesh), :101-145 (DeallocateMesh), :168-260 (growth):
//                   six SoA vertex buffers + one index buffer with element (not byte) offsets, two free lists.
//
// The list semantics are the reference's, quirks included, because they decide the offsets every draw command carries:
// placement looks at the FIRST free range only (findPlacement compares every candidate against *begin()), a range that
// fits exactly is erased, otherwise it shrinks from the front; growth asks for currentSize + request, rounds up to a
// power of 1.5 in float (closest_power_of<float>), re-packs the allocated ranges in list order and extends the free
// range with the largest start (stale holes stay in the free list); deallocation removes the first allocated node with
// start >= and count >= the request, widens EVERY bordering free range (both, when two border) and otherwise appends
// the requested range. One deliberate difference: on a growth the reference re-packs all six vertex buffers with the
// SAME node list, so buffers 2..6 are copied from already-updated starts (garbage after a hole); here every stream is
// copied from the true old starts.
#include <cmath>
#include <cstring>
#include <list>
#include <vector>

#include "internal.h"
#include "../../include/rve_shim.h"

using namespace rve;

namespace {

struct Range {
    uint32_t start = 0, count = 0;
};
using RangeList = std::list<Range>;

constexpr int kVertexStreams = 6;
constexpr uint32_t kStride[7] = {12, 12, 12, 12, 8, 8, 4};  // pos, normal, tangent, bitangent, uv0, lightmapUV, index

// closest_power_of<float>(x, 1.5f): pow(p, ceil(log(x) / log(p))) evaluated in float (include/RavEngine/Common3D.hpp:106-109)
uint32_t closest_power_of_1_5(uint32_t v) {
    const float x = float(v), p = 1.5f;
    return uint32_t(std::pow(p, std::ceil(std::log(x) / std::log(p))));
}

}  // namespace

struct rvs_arena {
    rve_ctx* ctx = nullptr;
    rve_buffer stream[7] = {RVE_BUFFER_INVALID, RVE_BUFFER_INVALID, RVE_BUFFER_INVALID, RVE_BUFFER_INVALID,
                            RVE_BUFFER_INVALID, RVE_BUFFER_INVALID, RVE_BUFFER_INVALID};
    uint32_t vertexSize = 0, indexSize = 0;
    RangeList vertexFree, indexFree, vertexAllocated, indexAllocated;
    struct Handle {
        RangeList::iterator vert, index;
        bool live = false;
    };
    std::vector<Handle> handles;  // MeshRange: iterators into the allocated lists (they follow a re-pack)
    uint32_t growths = 0;
};

namespace {

RangeList::iterator find_placement(uint32_t requested, RangeList& freeList) {
    auto best = freeList.begin();
    for (auto it = freeList.begin(); it != freeList.end(); ++it)
        if (requested <= best->count) return it;  // sic: the first range decides (src/RenderEngine_Allocate.cpp:23-31)
    return freeList.end();
}

void extend_last_range(RangeList& freeList, uint32_t oldSize, uint32_t newSize) {
    if (freeList.empty()) {
        freeList.push_back(Range{oldSize, newSize - oldSize});
        return;
    }
    auto last = freeList.begin();
    for (auto it = freeList.begin(); it != freeList.end(); ++it)
        if (it->start > last->start) last = it;
    last->count = newSize - last->start;
}

// ReallocateGeneric for one family (vertex: streams 0..5, index: stream 6)
int grow(rvs_arena* a, bool vertex, uint32_t minSize) {
    const uint32_t newSize = closest_power_of_1_5(minSize);
    uint32_t& cur = vertex ? a->vertexSize : a->indexSize;
    RangeList& allocated = vertex ? a->vertexAllocated : a->indexAllocated;
    RangeList& freeList = vertex ? a->vertexFree : a->indexFree;
    RVE_REQUIRE(newSize > cur, "arena growth did not enlarge the buffer (request overflows u32 elements)");
    const int s0 = vertex ? 0 : 6, s1 = vertex ? kVertexStreams : 7;
    std::vector<uint32_t> oldStart;
    for (const Range& r : allocated) oldStart.push_back(r.start);
    for (int s = s0; s < s1; ++s) {
        rve_buffer fresh = RVE_BUFFER_INVALID;
        int rc = rve_buffer_create(a->ctx, size_t(newSize) * kStride[s], &fresh);
        if (rc) return rc;
        uint32_t offset = 0;
        size_t k = 0;
        for (const Range& r : allocated) {
            if (r.count) {
                rc = rve_buffer_copy(a->ctx, a->stream[s], size_t(oldStart[k]) * kStride[s], fresh, size_t(offset) * kStride[s],
                                     size_t(r.count) * kStride[s]);
                if (rc) return rc;
            }
            offset += r.count;
            ++k;
        }
        rc = rve_ctx_synchronize(a->ctx);
        if (rc) return rc;
        rve_buffer_destroy(a->ctx, a->stream[s]);
        a->stream[s] = fresh;
    }
    uint32_t offset = 0;
    for (Range& r : allocated) {
        r.start = offset;
        offset += r.count;
    }
    extend_last_range(freeList, cur, newSize);
    cur = newSize;
    a->growths++;
    return 0;
}

int allocation_location(rvs_arena* a, bool vertex, uint32_t size, RangeList::iterator* out) {
    RangeList& freeList = vertex ? a->vertexFree : a->indexFree;
    for (int guard = 0; guard < 64; ++guard) {
        auto it = find_placement(size, freeList);
        if (it != freeList.end()) {
            *out = it;
            return 0;
        }
        const uint64_t want = uint64_t(vertex ? a->vertexSize : a->indexSize) + size;
        RVE_REQUIRE(want <= 0x7FFFFFFFull, "arena would exceed 2^31 elements");
        const int rc = grow(a, vertex, uint32_t(want));
        if (rc) return rc;
    }
    return set_error(RVE_ERR_UNSUPPORTED,
                     "the reference's placement loop does not terminate for this free list (its first range is a hole that never grows)");
}

RangeList::iterator consume_range(RangeList::iterator allocation, uint32_t allocatedSize, RangeList& freeList, RangeList& allocatedList) {
    Range r = *allocation;
    r.count = allocatedSize;
    allocatedList.push_back(r);
    if (allocation->count == allocatedSize) {
        freeList.erase(allocation);
    } else {
        allocation->start += allocatedSize;
        allocation->count -= allocatedSize;
    }
    return --allocatedList.end();
}

void deallocate(Range range, RangeList& allocatedList, RangeList& freeList, RangeList::iterator* erased) {
    Range found;
    *erased = allocatedList.end();
    for (auto it = allocatedList.begin(); it != allocatedList.end(); ++it)
        if (it->start >= range.start && it->count >= range.count) {
            found = *it;
            *erased = it;
            break;
        }
    bool overlap = false;
    for (Range& f : freeList) {
        if (f.start + f.count == found.start) {
            f.count += found.count;
            overlap = true;
        }
        if (found.start + found.count == f.start) {
            f.start -= found.count;
            f.count += found.count;
            overlap = true;
        }
    }
    if (!overlap) freeList.push_back(range);
}

}  // namespace

extern "C" {

int rvs_arena_create(rve_ctx* ctx, uint32_t initial_verts, uint32_t initial_indices, rvs_arena** out) {
    if (!ctx || !out) return set_error(RVE_ERR_INVALID_ARGUMENT, "null argument");
    auto* a = new rvs_arena;
    a->ctx = ctx;
    a->vertexSize = initial_verts ? initial_verts : 1024u;     // include/RavEngine/RenderEngine.hpp:87
    a->indexSize = initial_indices ? initial_indices : 1536u;
    a->vertexFree.push_back(Range{0, a->vertexSize});          // RenderEngine.hpp:335-336
    a->indexFree.push_back(Range{0, a->indexSize});
    for (int s = 0; s < 7; ++s) {
        const int rc = rve_buffer_create(ctx, size_t(s < kVertexStreams ? a->vertexSize : a->indexSize) * kStride[s], &a->stream[s]);
        if (rc) {
            rvs_arena_destroy(a);
            return rc;
        }
    }
    *out = a;
    return 0;
}

int rvs_arena_destroy(rvs_arena* a) {
    if (!a) return 0;
    for (int s = 0; s < 7; ++s)
        if (a->stream[s] != RVE_BUFFER_INVALID) rve_buffer_destroy(a->ctx, a->stream[s]);
    delete a;
    return 0;
}

int rvs_arena_allocate_mesh(rvs_arena* a, const rvs_mesh_view* mesh, uint32_t* out_range) {
    if (!a || !mesh || !out_range) return set_error(RVE_ERR_INVALID_ARGUMENT, "null argument");
    RangeList::iterator vAlloc, iAlloc;
    int rc = allocation_location(a, true, mesh->num_vertices, &vAlloc);
    if (rc) return rc;
    rc = allocation_location(a, false, mesh->num_indices, &iAlloc);
    if (rc) return rc;
    rvs_arena::Handle h;
    h.vert = consume_range(vAlloc, mesh->num_vertices, a->vertexFree, a->vertexAllocated);
    h.index = consume_range(iAlloc, mesh->num_indices, a->indexFree, a->indexAllocated);
    h.live = true;
    const void* src[7] = {mesh->positions, mesh->normals, mesh->tangents, mesh->bitangents, mesh->uv0, mesh->lightmap_uvs, mesh->indices};
    for (int s = 0; s < 7; ++s) {
        const uint32_t n = s < kVertexStreams ? mesh->num_vertices : mesh->num_indices;
        const uint32_t at = s < kVertexStreams ? h.vert->start : h.index->start;
        if (!src[s] || !n) continue;  // an absent attribute uploads an empty span (src/RenderEngine_Allocate.cpp:86-93)
        rc = rve_buffer_upload(a->ctx, a->stream[s], size_t(at) * kStride[s], src[s], size_t(n) * kStride[s]);
        if (rc) return rc;
    }
    a->handles.push_back(h);
    *out_range = uint32_t(a->handles.size() - 1);
    return 0;
}

int rvs_arena_deallocate_mesh(rvs_arena* a, uint32_t range) {
    if (!a) return set_error(RVE_ERR_INVALID_ARGUMENT, "arena is null");
    RVE_REQUIRE(range < a->handles.size() && a->handles[range].live, "unknown or already released mesh range");
    rvs_arena::Handle& h = a->handles[range];
    const Range v = *h.vert, i = *h.index;
    RangeList::iterator ev, ei;
    deallocate(v, a->vertexAllocated, a->vertexFree, &ev);
    deallocate(i, a->indexAllocated, a->indexFree, &ei);
    // the reference erases the first node that satisfies start >= and count >= (possibly another mesh's); every handle
    // that pointed at an erased node dies with it
    for (auto& o : a->handles)
        if (o.live && (o.vert == ev || o.index == ei)) o.live = false;
    if (ev != a->vertexAllocated.end()) a->vertexAllocated.erase(ev);
    if (ei != a->indexAllocated.end()) a->indexAllocated.erase(ei);
    h.live = false;
    return 0;
}

int rvs_arena_range(rvs_arena* a, uint32_t range, uint32_t* vertex_start, uint32_t* vertex_count, uint32_t* index_start,
                    uint32_t* index_count) {
    if (!a) return set_error(RVE_ERR_INVALID_ARGUMENT, "arena is null");
    RVE_REQUIRE(range < a->handles.size() && a->handles[range].live, "unknown or released mesh range");
    const rvs_arena::Handle& h = a->handles[range];
    if (vertex_start) *vertex_start = h.vert->start;
    if (vertex_count) *vertex_count = h.vert->count;
    if (index_start) *index_start = h.index->start;
    if (index_count) *index_count = h.index->count;
    return 0;
}

int rvs_arena_buffers(rvs_arena* a, rve_buffer out_streams[7], uint32_t* vertex_capacity, uint32_t* index_capacity) {
    if (!a) return set_error(RVE_ERR_INVALID_ARGUMENT, "arena is null");
    if (out_streams)
        for (int s = 0; s < 7; ++s) out_streams[s] = a->stream[s];
    if (vertex_capacity) *vertex_capacity = a->vertexSize;
    if (index_capacity) *index_capacity = a->indexSize;
    return 0;
}

uint32_t rvs_arena_list(rvs_arena* a, uint32_t which, uint32_t* out_pairs, uint32_t max_pairs) {
    if (!a) return 0;
    const RangeList& l = which == RVS_ARENA_VERTEX_FREE ? a->vertexFree : which == RVS_ARENA_INDEX_FREE ? a->indexFree
                       : which == RVS_ARENA_VERTEX_ALLOCATED ? a->vertexAllocated : a->indexAllocated;
    uint32_t n = 0;
    for (const Range& r : l) {
        if (out_pairs && n < max_pairs) {
            out_pairs[2 * n] = r.start;
            out_pairs[2 * n + 1] = r.count;
        }
        ++n;
    }
    return n;
}

int rvs_rvem_parse(const void* data, uint64_t size, rvs_rvem_info* out) {
    if (!data || !out) return set_error(RVE_ERR_INVALID_ARGUMENT, "null argument");
    RVE_REQUIRE(size >= 16, "not a mesh: shorter than the 16-byte .rvem header");
    const uint8_t* p = static_cast<const uint8_t*>(data);
    RVE_REQUIRE(std::memcmp(p, "rvem", 4) == 0, "Header does not match, data is not a mesh!");  // src/MeshAsset.cpp:30-32
    rvs_rvem_info info{};
    std::memcpy(&info.num_vertices, p + 4, 4);
    std::memcpy(&info.num_indices, p + 8, 4);
    info.attributes = p[12];
    uint64_t at = 16;
    const uint64_t none = ~0ull;
    uint64_t* off[6] = {&info.off_positions, &info.off_normals, &info.off_tangents, &info.off_bitangents, &info.off_uv0, &info.off_lightmap_uv};
    for (int s = 0; s < 6; ++s) {
        *off[s] = none;
        if (info.attributes & (RVS_RVEM_HAS_POSITIONS << s)) {
            *off[s] = at;
            at += uint64_t(info.num_vertices) * kStride[s];
        }
    }
    info.off_indices = at;
    at += uint64_t(info.num_indices) * 4;
    RVE_REQUIRE(at <= size, "truncated .rvem: the streams announced by the header run past the end of the data");
    info.mesh_bytes = at;
    info.off_weights = none;
    info.weights_bytes = 0;
    if (size - at >= 4) {  // src/MeshAssetSkinned.cpp:41-47: anything after the mesh is the weight table
        info.off_weights = at;
        info.weights_bytes = ((size - at) / 32) * 32;
    }
    *out = info;
    return 0;
}

int rvs_arena_load_rvem(rvs_arena* a, const void* data, uint64_t size, uint32_t skinned, uint32_t* out_range, rve_buffer* out_weights) {
    if (!a || !out_range) return set_error(RVE_ERR_INVALID_ARGUMENT, "null argument");
    rvs_rvem_info info;
    int rc = rvs_rvem_parse(data, size, &info);
    if (rc) return rc;
    if (skinned) {
        RVE_REQUIRE(out_weights, "a skinned load needs somewhere to return the weights buffer");
        RVE_REQUIRE(info.off_weights != ~0ull, "Mesh is probably not a skinned mesh");  // src/MeshAssetSkinned.cpp:41-43
        RVE_REQUIRE(info.weights_bytes / 32 == info.num_vertices,
                    "Skin does not have vertex weights for every vertex, input file is corrupt");  // :47
    }
    const uint8_t* p = static_cast<const uint8_t*>(data);
    const uint64_t none = ~0ull;
    auto at = [&](uint64_t off) -> const void* { return off == none ? nullptr : p + off; };
    rvs_mesh_view v{};
    v.positions = static_cast<const float*>(at(info.off_positions));
    v.normals = static_cast<const float*>(at(info.off_normals));
    v.tangents = static_cast<const float*>(at(info.off_tangents));
    v.bitangents = static_cast<const float*>(at(info.off_bitangents));
    v.uv0 = static_cast<const float*>(at(info.off_uv0));
    v.lightmap_uvs = static_cast<const float*>(at(info.off_lightmap_uv));
    v.indices = reinterpret_cast<const uint32_t*>(p + info.off_indices);
    v.num_vertices = info.num_vertices;
    v.num_indices = info.num_indices;
    rc = rvs_arena_allocate_mesh(a, &v, out_range);
    if (rc) return rc;
    if (skinned) {
        rve_buffer w = RVE_BUFFER_INVALID;
        rc = rve_buffer_create(a->ctx, info.weights_bytes ? info.weights_bytes : 32, &w);
        if (rc) return rc;
        if (info.weights_bytes) rc = rve_buffer_upload(a->ctx, w, 0, p + info.off_weights, info.weights_bytes);
        if (rc) return rc;
        *out_weights = w;
    }
    return 0;
}

}  // extern "C"
