// clusters.cu — clustered light grid build (shaders/cluster_build_grid.csh:16-79)
// and light assignment (8a-6, shaders/cluster_assign_lights.csh:28-88).
//
// The reference bins with 27 CTAs x 128 threads, one thread per cluster walking all lights
// serially through a 512-B-stride struct array. Here the lights are first reduced once to
// 16-B view-space records (centre, radius^2), then one WARP owns a cluster: 32 lights are
// tested per step (lane = light), hits are appended in index order with ballot/popc, and the
// warp stops as soon as 16 lights are found — the result is, as in the shader, the <=16
// lowest-index intersecting lights in ascending order. Eight clusters per CTA share each
// light tile through shared memory.
#include <cmath>

#include "internal.h"

namespace rve {

constexpr int kAssignThreads = 256;
constexpr int kAssignWarps = kAssignThreads / 32;
constexpr int kLightTile = 1024;

__global__ void cluster_build_kernel(const RveGridBuildUBO u, const float* __restrict__ slices,
                                     RveCluster* __restrict__ clusters, uint32_t numClusters) {
    const uint32_t tileIndex = blockIdx.x * blockDim.x + threadIdx.x;
    if (tileIndex >= numClusters) return;
    const uint32_t gx = tileIndex % u.gridSize[0];
    const uint32_t gy = (tileIndex / u.gridSize[0]) % u.gridSize[1];
    const uint32_t gz = tileIndex / (u.gridSize[0] * u.gridSize[1]);
    // uvec2 / uvec2: integer division, then conversion (cluster_build_grid.csh:48)
    const float tileSizeX = float(u.screenDim[0] / u.gridSize[0]);
    const float tileSizeY = float(u.screenDim[1] / u.gridSize[1]);

    // screenToView (cluster_build_grid.csh:16-26): ndc = (s/dim*2-1, 0, 1); view = invProj*ndc; /= w
    auto screenToView = [&](float sx, float sy, float& vx, float& vy, float& vz) {
        const float nx = sx / float(u.screenDim[0]) * 2.0f - 1.0f;
        const float ny = sy / float(u.screenDim[1]) * 2.0f - 1.0f;
        // mat*vec of the arithmetic contract: t = col0*nx; fma(col1,ny,t); fma(col2,0,t); fma(col3,1,t)
        float v[4];
#pragma unroll
        for (int r = 0; r < 4; ++r)
            v[r] = __fmaf_rn(u.invProj[12 + r], 1.0f,
                             __fmaf_rn(u.invProj[8 + r], 0.0f, __fmaf_rn(u.invProj[4 + r], ny, u.invProj[0 + r] * nx)));
        const float iw = __frcp_rn(v[3]);  // viewCoord /= viewCoord.w == viewCoord * (1/w)
        vx = v[0] * iw;
        vy = v[1] * iw;
        vz = v[2] * iw;
    };
    float mnx, mny, mnz, mxx, mxy, mxz;
    screenToView(float(gx) * tileSizeX, float(gy) * tileSizeY, mnx, mny, mnz);
    screenToView(float(gx + 1) * tileSizeX, float(gy + 1) * tileSizeY, mxx, mxy, mxz);
    const float tileNear = slices[gz], tileFar = slices[gz + 1];

    // lineIntersectionWithZPlane(eye=0, end, zDistance): direction = end - 0; normal = (0,0,-1)
    //   t = (zDistance - dot(normal, 0)) / dot(normal, direction);  result = 0 + t*direction
    auto isect = [](float ex, float ey, float ez, float zDistance, float& ox, float& oy, float& oz) {
        const float dx = ex - 0.0f, dy = ey - 0.0f, dz = ez - 0.0f;
        const float dotNS = __fmaf_rn(-1.0f, 0.0f, __fmaf_rn(0.0f, 0.0f, 0.0f * 0.0f));
        const float dotND = __fmaf_rn(-1.0f, dz, __fmaf_rn(0.0f, dy, 0.0f * dx));
        const float t = (zDistance - dotNS) / dotND;
        ox = __fmaf_rn(dx, t, 0.0f);  // startPoint + t * direction
        oy = __fmaf_rn(dy, t, 0.0f);
        oz = __fmaf_rn(dz, t, 0.0f);
    };
    float a[3], b[3], c[3], d[3];
    isect(mnx, mny, mnz, tileNear, a[0], a[1], a[2]);  // minPointNear
    isect(mnx, mny, mnz, tileFar, b[0], b[1], b[2]);   // minPointFar
    isect(mxx, mxy, mxz, tileNear, c[0], c[1], c[2]);  // maxPointNear
    isect(mxx, mxy, mxz, tileFar, d[0], d[1], d[2]);   // maxPointFar
    RveCluster* cl = clusters + tileIndex;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        // GLSL min/max spelled out: these floats are compared byte for byte (sign of zero included)
        cl->minPoint[k] = (b[k] < a[k]) ? b[k] : a[k];
        cl->maxPoint[k] = (c[k] < d[k]) ? d[k] : c[k];
    }
}

// light -> (view-space centre, radius^2); stride/offsets select point or spot layout
__global__ void light_prepare_kernel(const unsigned char* __restrict__ lights, uint32_t count, uint32_t stride,
                                     uint32_t posOffset, uint32_t intensityOffset, bool originFromMatrix,
                                     const RveGridAssignUBO u, float4* __restrict__ out) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= count) return;
    const unsigned char* L = lights + uint64_t(i) * stride;
    const float* pp = reinterpret_cast<const float*>(L + posOffset);
    float px = __ldg(pp), py = __ldg(pp + 1), pz = __ldg(pp + 2);
    (void)originFromMatrix;  // (worldTransform * vec4(0,0,0,1)).xyz is column 3 itself (arithmetic contract)
    const float intensity = __ldg(reinterpret_cast<const float*>(L + intensityOffset));
    const float* m = u.viewMat;
    float4 r;
    r.x = __fmaf_rn(m[12], 1.0f, __fmaf_rn(m[8], pz, __fmaf_rn(m[4], py, m[0] * px)));
    r.y = __fmaf_rn(m[13], 1.0f, __fmaf_rn(m[9], pz, __fmaf_rn(m[5], py, m[1] * px)));
    r.z = __fmaf_rn(m[14], 1.0f, __fmaf_rn(m[10], pz, __fmaf_rn(m[6], py, m[2] * px)));
    const float radius = sqrtf(intensity / RVE_LIGHT_MIN_INFLUENCE);  // cluster_shared.glsl:41-44
    r.w = radius * radius;
    out[i] = r;
}

// One warp per cluster, eight clusters per CTA (x-neighbours of one y / z-slice: clusters are indexed x fastest); kinds
// processed one after the other (points, then spots). Per tile of kLightTile lights the CTA first rejects, all threads
// together, every light whose sphere misses the UNION of its eight AABBs - one y band and one depth slice of the grid,
// so the large majority - and compacts the rest IN INDEX ORDER (ballot / popc, a 32-entry scan) into shared memory;
// the warps then run the shader's exact test (cluster_assign_lights.csh:28-35) over the survivors only and append the
// first 16 in index order (:60-67). The union test uses the same arithmetic and is conservative in floating point
// (clamp is exact, rounding is monotonic: the distance to a superset box never rounds above the distance to the box),
// so the result is the shader's, bit for bit.
__device__ __forceinline__ bool sphere_hits_box(const float4 L, const float (&mn)[3], const float (&mx)[3]) {
    const float qx = gmin(gmax(L.x, mn[0]), mx[0]) - L.x;
    const float qy = gmin(gmax(L.y, mn[1]), mx[1]) - L.y;
    const float qz = gmin(gmax(L.z, mn[2]), mx[2]) - L.z;
    return dot3(qx, qy, qz, qx, qy, qz) <= L.w;
}

__global__ void __launch_bounds__(kAssignThreads)
cluster_assign_kernel(RveCluster* __restrict__ clusters, uint32_t numClusters, const float4* __restrict__ pointRec,
                      uint32_t pointCount, const float4* __restrict__ spotRec, uint32_t spotCount) {
    __shared__ float4 s_surv[kLightTile];      // lights of the tile that touch the CTA's union box, in index order
    __shared__ uint32_t s_survIdx[kLightTile];
    __shared__ float s_box[kAssignWarps][6];
    __shared__ uint32_t s_off[kAssignWarps * 4 + 1];
    constexpr uint32_t kSteps = kLightTile / (kAssignWarps * 32);  // prefilter steps per warp and tile
    static_assert(kSteps == 4, "s_off is sized for four prefilter steps per warp");
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t index = blockIdx.x * kAssignWarps + warp;
    const bool have = index < numClusters;
    float mn[3] = {0, 0, 0}, mx[3] = {0, 0, 0};
    if (have) {
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            mn[k] = clusters[index].minPoint[k];
            mx[k] = clusters[index].maxPoint[k];
        }
    }
    if (lane == 0) {
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            s_box[warp][k] = have ? mn[k] : __int_as_float(0x7f800000);       // +inf: ignored by the min
            s_box[warp][3 + k] = have ? mx[k] : -__int_as_float(0x7f800000);  // -inf
        }
    }
    __syncthreads();
    float umn[3], umx[3];  // union of the CTA's boxes
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        umn[k] = s_box[0][k];
        umx[k] = s_box[0][3 + k];
#pragma unroll
        for (int w = 1; w < kAssignWarps; ++w) {
            umn[k] = fminf(umn[k], s_box[w][k]);
            umx[k] = fmaxf(umx[k], s_box[w][3 + k]);
        }
    }
    for (int kind = 0; kind < 2; ++kind) {
        const float4* rec = kind ? spotRec : pointRec;
        const uint32_t count = kind ? spotCount : pointCount;
        const uint32_t cap = kind ? RVE_CLUSTER_MAX_SPOTS : RVE_CLUSTER_MAX_POINTS;
        uint32_t* outIdx = have ? (kind ? clusters[index].spotLightIndices : clusters[index].pointLightIndices) : nullptr;
        uint32_t found = 0;  // warp-uniform
        // the next tile's records are requested one tile ahead: at three CTAs per SM the tile loop is otherwise bound by the
        // latency of these loads (a view whose clusters never fill scans all 64 tiles)
        float4 Lnext[kSteps];
#pragma unroll
        for (uint32_t st = 0; st < kSteps; ++st) {
            const uint32_t i = (warp * kSteps + st) * 32 + lane;
            Lnext[st] = i < count ? __ldg(rec + i) : make_float4(0.f, 0.f, 0.f, -1.f);
        }
        for (uint32_t base = 0; base < count; base += kLightTile) {
            // stop when every cluster of the CTA is full (also: the survivor list of the previous tile is free again)
            if (__syncthreads_and(!have || found >= cap)) break;
            // ---- prefilter: warp w owns lights [base + 128 w, base + 128 (w + 1)) of the tile, 32 per step ----
            float4 L[kSteps];
            uint32_t ballots[kSteps];
#pragma unroll
            for (uint32_t st = 0; st < kSteps; ++st) L[st] = Lnext[st];
#pragma unroll
            for (uint32_t st = 0; st < kSteps; ++st) {
                const uint32_t i = base + kLightTile + (warp * kSteps + st) * 32 + lane;
                Lnext[st] = i < count ? __ldg(rec + i) : make_float4(0.f, 0.f, 0.f, -1.f);
            }
#pragma unroll
            for (uint32_t st = 0; st < kSteps; ++st) {
                const uint32_t i = base + (warp * kSteps + st) * 32 + lane;
                const bool hit = i < count && sphere_hits_box(L[st], umn, umx);
                ballots[st] = __ballot_sync(0xffffffffu, hit);
                if (lane == 0) s_off[warp * kSteps + st] = __popc(ballots[st]);
            }
            __syncthreads();
            if (warp == 0) {  // exclusive scan of the 32 (warp, step) counts
                const uint32_t v = s_off[lane];
                uint32_t x = v;
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) {
                    const uint32_t y = __shfl_up_sync(0xffffffffu, x, o);
                    if (lane >= uint32_t(o)) x += y;
                }
                s_off[lane] = x - v;
                if (lane == 31) s_off[32] = x;
            }
            __syncthreads();
#pragma unroll
            for (uint32_t st = 0; st < kSteps; ++st)
                if ((ballots[st] >> lane) & 1u) {
                    const uint32_t pos = s_off[warp * kSteps + st] + __popc(ballots[st] & ((1u << lane) - 1u));
                    s_surv[pos] = L[st];
                    s_survIdx[pos] = base + (warp * kSteps + st) * 32 + lane;
                }
            __syncthreads();
            // ---- exact test over the survivors, in index order ----
            const uint32_t n = s_off[32];
            if (have) {
                for (uint32_t j = 0; j < n && found < cap; j += 32) {
                    const bool hit = (j + lane < n) && sphere_hits_box(s_surv[j + lane], mn, mx);
                    const uint32_t ballot = __ballot_sync(0xffffffffu, hit);
                    if (hit) {
                        const uint32_t pos = found + __popc(ballot & ((1u << lane) - 1u));
                        if (pos < cap) outIdx[pos] = s_survIdx[j + lane];
                    }
                    found = min(cap, found + uint32_t(__popc(ballot)));
                }
            }
        }
        if (have && lane == 0) {
            if (kind) clusters[index].spotLightCount = found;
            else clusters[index].pointLightCount = found;
        }
        __syncthreads();
    }
}

}  // namespace rve

using namespace rve;

extern "C" {

int rve_cluster_build(rve_ctx* ctx, const RveGridBuildUBO* ubo, rve_buffer clusters) {
    if (!ctx || !ubo) return set_error(RVE_ERR_INVALID_ARGUMENT, "null argument");
    DeviceGuard g(ctx->device);
    const uint64_t n64 = uint64_t(ubo->gridSize[0]) * ubo->gridSize[1] * ubo->gridSize[2];
    if (n64 == 0) return 0;
    RVE_REQUIRE(n64 <= 0x7FFFFFFFull, "cluster grid too large");
    BufferRec* b = get_buffer(ctx, clusters, "clusters");
    if (!b) return RVE_ERR_BAD_HANDLE;
    RVE_REQUIRE(n64 * sizeof(RveCluster) <= b->bytes, "cluster buffer too small");
    // slice depths zFar*pow(zNear/zFar, k/gridZ) (cluster_build_grid.csh:58-61): evaluated once
    // with the host libm (GPU pow is vendor-specific; the oracle uses the same libm)
    const uint32_t gz = ubo->gridSize[2];
    std::vector<float> slices(gz + 1);
    for (uint32_t k = 0; k <= gz; ++k) slices[k] = ubo->zFar * std::pow(ubo->zNear / ubo->zFar, float(k) / float(gz));
    int rc = ensure_scratch(ctx, reinterpret_cast<void**>(&ctx->d_slices), &ctx->d_slices_cap, slices.size() * sizeof(float));
    if (rc) return rc;
    rc = staged_h2d(ctx, ctx->d_slices, slices.data(), slices.size() * sizeof(float));
    if (rc) return rc;
    const uint32_t n = uint32_t(n64), threads = 128;
    cluster_build_kernel<<<(n + threads - 1) / threads, threads, 0, ctx->stream>>>(*ubo, ctx->d_slices,
                                                                                 static_cast<RveCluster*>(b->ptr), n);
    ctx->launches++;
    RVE_CUDA(cudaGetLastError());
    return 0;
}

int rve_cluster_assign(rve_ctx* ctx, const RveGridAssignUBO* ubo, rve_buffer clusters, rve_buffer point_lights,
                       rve_buffer spot_lights, uint32_t num_clusters) {
    if (!ctx || !ubo) return set_error(RVE_ERR_INVALID_ARGUMENT, "null argument");
    DeviceGuard g(ctx->device);
    if (num_clusters == 0) return 0;
    BufferRec* b = get_buffer(ctx, clusters, "clusters");
    if (!b) return RVE_ERR_BAD_HANDLE;
    RVE_REQUIRE(uint64_t(num_clusters) * sizeof(RveCluster) <= b->bytes, "cluster buffer too small");
    const unsigned char* pl = nullptr;
    const unsigned char* sl = nullptr;
    if (ubo->pointLightCount) {
        BufferRec* p = get_buffer(ctx, point_lights, "point lights");
        if (!p) return RVE_ERR_BAD_HANDLE;
        RVE_REQUIRE(uint64_t(ubo->pointLightCount) * sizeof(RvePointLight) <= p->bytes, "point light buffer too small");
        pl = static_cast<const unsigned char*>(p->ptr);
    }
    if (ubo->spotLightCount) {
        BufferRec* s = get_buffer(ctx, spot_lights, "spot lights");
        if (!s) return RVE_ERR_BAD_HANDLE;
        RVE_REQUIRE(uint64_t(ubo->spotLightCount) * sizeof(RveSpotLight) <= s->bytes, "spot light buffer too small");
        sl = static_cast<const unsigned char*>(s->ptr);
    }
    const size_t total = size_t(ubo->pointLightCount) + ubo->spotLightCount;
    int rc = ensure_scratch(ctx, reinterpret_cast<void**>(&ctx->d_lights), &ctx->d_lights_cap, (total + 1) * sizeof(float4));
    if (rc) return rc;
    float4* prec = ctx->d_lights;
    float4* srec = ctx->d_lights + ubo->pointLightCount;
    const uint32_t threads = 256;
    if (ubo->pointLightCount) {
        light_prepare_kernel<<<(ubo->pointLightCount + threads - 1) / threads, threads, 0, ctx->stream>>>(
            pl, ubo->pointLightCount, uint32_t(sizeof(RvePointLight)), uint32_t(offsetof(RvePointLight, position)),
            uint32_t(offsetof(RvePointLight, intensity)), false, *ubo, prec);
        ctx->launches++;
    }
    if (ubo->spotLightCount) {
        light_prepare_kernel<<<(ubo->spotLightCount + threads - 1) / threads, threads, 0, ctx->stream>>>(
            sl, ubo->spotLightCount, uint32_t(sizeof(RveSpotLight)), uint32_t(offsetof(RveSpotLight, worldTransform) + 48),
            uint32_t(offsetof(RveSpotLight, intensity)), true, *ubo, srec);
        ctx->launches++;
    }
    ProfScope prof(ctx, RVE_PROF_CLUSTER_ASSIGN);
    cluster_assign_kernel<<<(num_clusters + kAssignWarps - 1) / kAssignWarps, kAssignThreads, 0, ctx->stream>>>(
        static_cast<RveCluster*>(b->ptr), num_clusters, prec, ubo->pointLightCount, srec, ubo->spotLightCount);
    ctx->launches++;
    RVE_CUDA(cudaGetLastError());
    return 0;
}

}  // extern "C"
