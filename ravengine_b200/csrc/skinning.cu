// skinning.cu — batched linear-blend skinning (shaders/skinning_cs.csh:95-138)
// and skinned draw-call preparation (8a-4, shaders/skinned_mesh_drawcall.csh:25-39).
//
// The shader maps x = object (fast axis), y = vertex, so neighbouring invocations write
// addresses numVertices apart and every (object, vertex) pair re-reads the vertex streams
// and four 64-B matrices from memory. Here a CTA owns (a group of objects) x (a run of
// vertices): the group's joint palettes are staged once into shared memory as 3x4 row
// blocks with a 48-B joint stride (conflict-free 128-bit reads for distinct joints mod 8),
// each thread keeps its vertices' inputs and weights in registers across the object loop,
// and the vertex index is the fast axis so the five output streams are written in
// contiguous runs.
#include "internal.h"

namespace rve {

constexpr int kSkinThreads = 256;
constexpr int kSkinMaxGroup = 8;

struct SkinParams {
    RveSkinningUBO u;
    float* __restrict__ outPos;
    float* __restrict__ outNrm;
    float* __restrict__ outTan;
    float* __restrict__ outBit;
    float* __restrict__ outUV;
    const float* __restrict__ inPos;
    const float* __restrict__ inNrm;
    const float* __restrict__ inTan;
    const float* __restrict__ inBit;
    const float* __restrict__ inUV;
    const float4* __restrict__ pose;  // mat4 = 4 x float4 columns
    const RveVertexJointBinding* __restrict__ weights;
    uint32_t groupBase;     // first object group of this launch (grid.y is capped at 65535 groups per launch)
    uint32_t group;         // objects per CTA
    uint32_t vertsPerCta;   // multiple of kSkinThreads
};

struct Vec3 {
    float x, y, z;
};

__device__ __forceinline__ Vec3 ld3(const float* __restrict__ p, uint64_t i) {
    return {__ldg(p + i * 3), __ldg(p + i * 3 + 1), __ldg(p + i * 3 + 2)};
}
__device__ __forceinline__ void st3(float* __restrict__ p, uint64_t i, Vec3 v) {
    p[i * 3] = v.x;
    p[i * 3 + 1] = v.y;
    p[i * 3 + 2] = v.z;
}

// rows[r] = (m[0][r], m[1][r], m[2][r], m[3][r]) of the blended matrix, r = 0..2
struct Blend {
    float4 r0, r1, r2;
};

__device__ __forceinline__ float4 fma4(float4 acc, float w, float4 m) {
    // totalmtx += weight * posed_mtx (skinning_cs.csh:115): one fma per component (arithmetic contract)
    return make_float4(__fmaf_rn(w, m.x, acc.x), __fmaf_rn(w, m.y, acc.y), __fmaf_rn(w, m.z, acc.z), __fmaf_rn(w, m.w, acc.w));
}

template <bool kSharedPalette>
__global__ void __launch_bounds__(kSkinThreads) skin_kernel(const SkinParams p) {
    extern __shared__ float4 s_pal[];  // [group][numBones][3] row blocks
    const uint32_t obj0 = (p.groupBase + blockIdx.y) * p.group;
    const uint32_t nObj = min(p.group, p.u.numObjects - obj0);
    const uint32_t v0 = blockIdx.x * p.vertsPerCta;

    if (kSharedPalette) {
        // stage palettes: thread handles (object, joint, column): coalesced 128-bit reads,
        // transposed scalar writes into the row-block layout
        float* s_f = reinterpret_cast<float*>(s_pal);
        const uint32_t per_obj = p.u.numBones * 4;
        for (uint32_t i = threadIdx.x; i < nObj * per_obj; i += blockDim.x) {
            const uint32_t o = i / per_obj, rem = i % per_obj;
            const uint32_t j = rem >> 2, c = rem & 3;
            const uint64_t bone = uint64_t(p.u.numBones) * (obj0 + o) + p.u.boneReadOffset + j;
            const float4 col = __ldg(p.pose + bone * 4 + c);
            float* dst = s_f + (uint64_t(o) * p.u.numBones + j) * 12;
            dst[0 + c] = col.x;
            dst[4 + c] = col.y;
            dst[8 + c] = col.z;
        }
        __syncthreads();
    }

    const uint32_t vEnd = min(v0 + p.vertsPerCta, p.u.numVertices);
    for (uint32_t vertID = v0 + threadIdx.x; vertID < vEnd; vertID += blockDim.x) {
        const uint64_t readOffset = uint64_t(vertID) + p.u.vertexReadOffset;
        const float4* wp = reinterpret_cast<const float4*>(p.weights + vertID);
        const float4 wa = __ldg(wp), wb = __ldg(wp + 1);
        const uint32_t j0 = __float_as_uint(wa.x), j1 = __float_as_uint(wa.z);
        const uint32_t j2 = __float_as_uint(wb.x), j3 = __float_as_uint(wb.z);
        const float w0 = wa.y, w1 = wa.w, w2 = wb.y, w3 = wb.w;
        const Vec3 pos = ld3(p.inPos, readOffset), nrm = ld3(p.inNrm, readOffset);
        const Vec3 tan = ld3(p.inTan, readOffset), bit = ld3(p.inBit, readOffset);
        const float2 uv = make_float2(__ldg(p.inUV + readOffset * 2), __ldg(p.inUV + readOffset * 2 + 1));

        for (uint32_t o = 0; o < nObj; ++o) {
            Blend M;
            M.r0 = M.r1 = M.r2 = make_float4(0.f, 0.f, 0.f, 0.f);
            const uint32_t js[4] = {j0, j1, j2, j3};
            const float ws[4] = {w0, w1, w2, w3};
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                float4 a, b, c;
                if (kSharedPalette) {
                    const float4* jb = s_pal + (uint64_t(o) * p.u.numBones + js[i]) * 3;
                    a = jb[0];
                    b = jb[1];
                    c = jb[2];
                } else {
                    const uint64_t bone = uint64_t(p.u.numBones) * (obj0 + o) + p.u.boneReadOffset + js[i];
                    const float4 c0 = __ldg(p.pose + bone * 4), c1 = __ldg(p.pose + bone * 4 + 1);
                    const float4 c2 = __ldg(p.pose + bone * 4 + 2), c3 = __ldg(p.pose + bone * 4 + 3);
                    a = make_float4(c0.x, c1.x, c2.x, c3.x);
                    b = make_float4(c0.y, c1.y, c2.y, c3.y);
                    c = make_float4(c0.z, c1.z, c2.z, c3.z);
                }
                M.r0 = fma4(M.r0, ws[i], a);
                M.r1 = fma4(M.r1, ws[i], b);
                M.r2 = fma4(M.r2, ws[i], c);
            }
            const uint64_t writeOffset = uint64_t(vertID) + uint64_t(obj0 + o) * p.u.numVertices + p.u.vertexWriteOffset;
            // (totalmtx * vec4(pos,1)).xyz: t = col0*x; fma(col1,y,t); fma(col2,z,t); fma(col3,1,t) = t + col3
            Vec3 op;
            op.x = dot3(pos.x, pos.y, pos.z, M.r0.x, M.r0.y, M.r0.z) + M.r0.w;
            op.y = dot3(pos.x, pos.y, pos.z, M.r1.x, M.r1.y, M.r1.z) + M.r1.w;
            op.z = dot3(pos.x, pos.y, pos.z, M.r2.x, M.r2.y, M.r2.z) + M.r2.w;
            st3(p.outPos, writeOffset, op);
            auto rot = [&](Vec3 v) -> Vec3 {
                return {dot3(v.x, v.y, v.z, M.r0.x, M.r0.y, M.r0.z), dot3(v.x, v.y, v.z, M.r1.x, M.r1.y, M.r1.z),
                        dot3(v.x, v.y, v.z, M.r2.x, M.r2.y, M.r2.z)};
            };
            st3(p.outNrm, writeOffset, rot(nrm));
            st3(p.outTan, writeOffset, rot(tan));
            st3(p.outBit, writeOffset, rot(bit));
            p.outUV[writeOffset * 2] = uv.x;
            p.outUV[writeOffset * 2 + 1] = uv.y;
        }
    }
}

__global__ void skinned_drawcall_kernel(const RveSkinningPrepareUBO u, RveIndirectIndexedCommand* __restrict__ commands) {
    const uint32_t objectID = blockIdx.x * blockDim.x + threadIdx.x;
    if (objectID >= u.nTotalObjects) return;
    uint32_t* c = reinterpret_cast<uint32_t*>(commands + (uint64_t(u.drawCallBufferOffset) + objectID));
    c[0] = u.nIndicesInThisMesh;
    c[1] = 0;
    c[2] = u.indexBufferOffset + u.nIndicesInThisMesh * objectID;
    c[3] = u.vertexBufferOffset + u.nVerticesInThisMesh * objectID;
    c[4] = u.baseInstanceOffset + objectID;
}

}  // namespace rve

using namespace rve;

extern "C" {

int rve_skin(rve_ctx* ctx, const RveSkinningUBO* ubo, const rve_buffer out_streams[5], const rve_buffer in_streams[5],
             rve_buffer pose, rve_buffer weights) {
    if (!ctx || !ubo || !out_streams || !in_streams) return set_error(RVE_ERR_INVALID_ARGUMENT, "null argument");
    RVE_PROFILE_FN_N("Enc Pose Skinned Meshes");
    DeviceGuard g(ctx->device);
    if (ubo->numObjects == 0 || ubo->numVertices == 0) return 0;
    BufferRec* o[5];
    BufferRec* in[5];
    static const char* names[5] = {"position", "normal", "tangent", "bitangent", "uv"};
    for (int i = 0; i < 5; ++i) {
        o[i] = get_buffer(ctx, out_streams[i], names[i]);
        in[i] = get_buffer(ctx, in_streams[i], names[i]);
        if (!o[i] || !in[i]) return RVE_ERR_BAD_HANDLE;
    }
    BufferRec* bp = get_buffer(ctx, pose, "pose");
    BufferRec* bw = get_buffer(ctx, weights, "weights");
    if (!bp || !bw) return RVE_ERR_BAD_HANDLE;
    const uint64_t outVerts = uint64_t(ubo->vertexWriteOffset) + uint64_t(ubo->numObjects) * ubo->numVertices;
    const uint64_t inVerts = uint64_t(ubo->vertexReadOffset) + ubo->numVertices;
    for (int i = 0; i < 5; ++i) {
        const size_t elem = (i == 4) ? 8 : 12;
        RVE_REQUIRE(outVerts * elem <= o[i]->bytes, "skinning output stream too small");
        RVE_REQUIRE(inVerts * elem <= in[i]->bytes, "skinning input stream too small");
    }
    RVE_REQUIRE((uint64_t(ubo->boneReadOffset) + uint64_t(ubo->numBones) * ubo->numObjects) * 64 <= bp->bytes,
                "pose buffer too small");
    RVE_REQUIRE(uint64_t(ubo->numVertices) * 32 <= bw->bytes, "weights buffer too small");
    RVE_REQUIRE((reinterpret_cast<uintptr_t>(bp->ptr) & 15) == 0 && (reinterpret_cast<uintptr_t>(bw->ptr) & 15) == 0,
                "pose / weights buffers must be 16-byte aligned");

    SkinParams p{};
    p.u = *ubo;
    p.outPos = static_cast<float*>(o[0]->ptr);
    p.outNrm = static_cast<float*>(o[1]->ptr);
    p.outTan = static_cast<float*>(o[2]->ptr);
    p.outBit = static_cast<float*>(o[3]->ptr);
    p.outUV = static_cast<float*>(o[4]->ptr);
    p.inPos = static_cast<const float*>(in[0]->ptr);
    p.inNrm = static_cast<const float*>(in[1]->ptr);
    p.inTan = static_cast<const float*>(in[2]->ptr);
    p.inBit = static_cast<const float*>(in[3]->ptr);
    p.inUV = static_cast<const float*>(in[4]->ptr);
    p.pose = static_cast<const float4*>(bp->ptr);
    p.weights = static_cast<const RveVertexJointBinding*>(bw->ptr);

    const size_t palBytes = size_t(ubo->numBones) * 48;
    const size_t smemBudget = 96 * 1024;
    uint32_t group = kSkinMaxGroup;
    while (group > 1 && palBytes * group > smemBudget) group >>= 1;
    const bool shared = palBytes * group <= 200 * 1024 && ubo->numBones > 0;
    if (group > ubo->numObjects) group = ubo->numObjects;
    p.group = group;
    // vertices per CTA: enough CTAs to fill the machine, few enough to amortise the palette staging
    uint32_t vertsPerCta = kSkinThreads * 4;
    p.vertsPerCta = vertsPerCta;
    const uint32_t groups = (ubo->numObjects + group - 1) / group;
    ProfScope prof(ctx, RVE_PROF_SKIN);
    const size_t smem = shared ? palBytes * group : 0;
    if (smem > 48 * 1024) RVE_CUDA(cudaFuncSetAttribute(skin_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
    for (uint32_t g0 = 0; g0 < groups; g0 += 65535u) {  // grid.y holds 65535 groups: more objects take further launches
        p.groupBase = g0;
        const dim3 grid((ubo->numVertices + vertsPerCta - 1) / vertsPerCta, groups - g0 < 65535u ? groups - g0 : 65535u);
        if (shared) skin_kernel<true><<<grid, kSkinThreads, smem, ctx->stream>>>(p);
        else skin_kernel<false><<<grid, kSkinThreads, 0, ctx->stream>>>(p);
        ctx->launches++;
        RVE_CUDA(cudaGetLastError());
    }
    return 0;
}

int rve_skinned_drawcall_prepare(rve_ctx* ctx, const RveSkinningPrepareUBO* ubo, rve_buffer indirect) {
    if (!ctx || !ubo) return set_error(RVE_ERR_INVALID_ARGUMENT, "null argument");
    DeviceGuard g(ctx->device);
    if (ubo->nTotalObjects == 0) return 0;
    BufferRec* b = get_buffer(ctx, indirect, "indirect");
    if (!b) return RVE_ERR_BAD_HANDLE;
    RVE_REQUIRE((uint64_t(ubo->drawCallBufferOffset) + ubo->nTotalObjects) * sizeof(RveIndirectIndexedCommand) <= b->bytes,
                "indirect buffer too small");
    const uint32_t threads = 256;
    skinned_drawcall_kernel<<<(ubo->nTotalObjects + threads - 1) / threads, threads, 0, ctx->stream>>>(
        *ubo, static_cast<RveIndirectIndexedCommand*>(b->ptr));
    ctx->launches++;
    RVE_CUDA(cudaGetLastError());
    return 0;
}

}  // extern "C"
