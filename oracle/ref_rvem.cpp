// ref_rvem.cpp — TEST INFRASTRUCTURE. Drives the reference's OWN .rvem writer so that its bytes pin rvs_rvem_parse / the
// Python writer in ravengine_b200/mesh_arena.py. Nothing of the reference is copied into this repository:
//   * SerializedMeshDataHeader, MeshPart and VertexWeights come from the reference's headers where they lie
//     (include/RavEngine/Mesh.hpp:12-27,35-60, include/RavEngine/Skeleton.hpp:50-60; glm and parallel-hashmap from deps/);
//   * SerializeMeshPart itself (tools/rvemc/rvemc.cpp:233-292) lives in a .cpp that also pulls assimp / cxxopts / fmt /
//     simdjson / meshoptimizer, so oracle/Makefile cuts exactly that function out of the file at BUILD time into
//     oracle/_ref/rvemc_serialize.inc (git-ignored build output) and this file compiles it unchanged, with the two
//     names it needs from its surroundings declared below (SkinnedMeshPart as at rvemc.cpp:51-53, FATAL as at :25, and
//     fmt::format for its one error message).
// Built into oracle/_ref/libref_rvem.so only where /root/reference exists; tests/golden/make_golden_rvem.py turns its
// output files into tests/golden/rvem_*.bin.
#include <RavEngine/Mesh.hpp>
#include <RavEngine/Skeleton.hpp>
#include <RavEngine/CaseAnalysis.hpp>

#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <filesystem>
#include <fstream>
#include <iostream>
#include <string>
#include <variant>
#include <vector>

using namespace std;
using namespace RavEngine;

#define FATAL(reason) {std::cerr << "rvemc error: " << reason << std::endl; std::exit(1);}
namespace fmt {
inline std::string format(const char* f, const std::string& a) { return std::string(f) + " " + a; }
}  // namespace fmt

struct SkinnedMeshPart : public MeshPart {
    std::vector<VertexWeights> vertexWeights;
};

#include "_ref/rvemc_serialize.inc"  // void SerializeMeshPart(const std::filesystem::path&, const std::variant<MeshPart,SkinnedMeshPart>&)

// host arrays in, the reference's file out. Streams: positions/normals/tangents/bitangents 3 floats, uv0/lightmap 2 floats
// per vertex (lightmap may be null = absent), u32 indices, weights 8 words per vertex (skinned only).
extern "C" __attribute__((visibility("default"))) int ref_rvem_write(const char* path, uint32_t nVerts, uint32_t nIndices, const float* pos,
                                                                     const float* nrm, const float* tan, const float* bit, const float* uv0,
                                                                     const float* lightmap, const uint32_t* indices, const uint32_t* weights,
                                                                     int skinned) {
    static_assert(sizeof(VertexWeights) == 32, "VertexWeights layout");
    auto fill = [&](MeshPart& m) {
        m.positions.resize(nVerts);
        m.normals.resize(nVerts);
        m.tangents.resize(nVerts);
        m.bitangents.resize(nVerts);
        m.uv0.resize(nVerts);
        std::memcpy(m.positions.data(), pos, size_t(nVerts) * sizeof(m.positions[0]));
        std::memcpy(m.normals.data(), nrm, size_t(nVerts) * sizeof(m.normals[0]));
        std::memcpy(m.tangents.data(), tan, size_t(nVerts) * sizeof(m.tangents[0]));
        std::memcpy(m.bitangents.data(), bit, size_t(nVerts) * sizeof(m.bitangents[0]));
        std::memcpy(m.uv0.data(), uv0, size_t(nVerts) * sizeof(m.uv0[0]));
        if (lightmap) {
            m.lightmapUVs.resize(nVerts);
            std::memcpy(m.lightmapUVs.data(), lightmap, size_t(nVerts) * sizeof(m.lightmapUVs[0]));
        }
        m.indices.assign(indices, indices + nIndices);
    };
    if (sizeof(MeshPart{}.positions[0]) != 12 || sizeof(MeshPart{}.uv0[0]) != 8) return -1;
    if (skinned) {
        SkinnedMeshPart m;
        fill(m);
        m.vertexWeights.resize(nVerts);
        std::memcpy(m.vertexWeights.data(), weights, size_t(nVerts) * sizeof(VertexWeights));
        SerializeMeshPart(path, std::variant<MeshPart, SkinnedMeshPart>(std::move(m)));
    } else {
        MeshPart m;
        fill(m);
        SerializeMeshPart(path, std::variant<MeshPart, SkinnedMeshPart>(std::move(m)));
    }
    return 0;
}
