/*
 * oracle.cpp — CPU restatement of the reference's render-prep compute shaders.
 *
 * TEST INFRASTRUCTURE ONLY (see oracle.h). PARITY UNPINNED by reference tests:
 * the shader sources below are the only specification.
 *
 * Each function follows the GLSL source line ranges it cites. GLSL semantics
 * restated here (GLSL 4.60 spec, 8.3 "Common Functions", 5.10 "Vector and
 * Matrix Operations"):
 *   min(x,y)  = y < x ? y : x          max(x,y) = x < y ? y : x
 *   clamp(x,a,b) = min(max(x,a),b)
 *
 * ARITHMETIC CONTRACT (shared bit for bit with the CUDA kernels; DESIGN.md
 * "Arithmetic contract"). GLSL fixes neither the evaluation order of dot /
 * mat*vec nor whether a*b+c is fused (no `precise` qualifier in the shaders,
 * no NoContraction in the SPIR-V), and Vulkan only asks 2.5 ULP of a division,
 * so every GPU the reference runs on contracts to FMA and divides by
 * reciprocal-multiply. This restatement pins ONE such evaluation, the one a
 * shader compiler emits, with explicit fused multiply-adds:
 *   dot(a,b)  = fma(a.w,b.w, fma(a.z,b.z, fma(a.y,b.y, a.x*b.x)))   (x first)
 *   length(v) = sqrt(dot(v,v))                       (correctly rounded sqrt)
 *   mat*vec   = t=col0*v.x; t=fma(col1,v.y,t); t=fma(col2,v.z,t); t=fma(col3,v.w,t)
 *   v / s     = v * (1/s) for a vector v and scalar s (1/s correctly rounded)
 *   a*b + c   = fma(a,b,c) where the shader writes exactly that shape
 *               (u*W - 0.5 in the sampler, total += w*M in skinning)
 * Nothing else is fused: build with -ffp-contract=off so the compiler adds no
 * contraction of its own; fma() is the C99 correctly-rounded one (hardware
 * FMA3 when built with -mfma, libm otherwise - same bits either way).
 * Products with a literal 0/1 vector component are folded to their exact
 * value: (model*vec4(0,0,0,1)).xyz == model[3].xyz and
 * mat3(model)*vec3(0,0,r) == model[2].xyz*r (equal up to the sign of zero for
 * finite matrices; the sign of a zero never reaches a comparison).
 * Parity is claimed for inputs that produce no NaN.
 *
 * Two further definitional choices where the reference's result depends on GPU
 * hardware (SURVEY.md 7 "hard parts" 1-2), both documented in DESIGN.md:
 *   - ceil(log2(x)) (defaultcull.csh:231) is evaluated exactly from the float
 *     exponent (no libm);
 *   - the min-reduction sampler (src/RenderEngine.cpp:323-331: linear filter,
 *     nearest mip, clamp-to-edge, VK_SAMPLER_REDUCTION_MODE_MIN) takes the min
 *     over all four texels of the bilinear footprint floor(u*W-0.5)+{0,1},
 *     clamped, which is also what the shader's Metal/WGSL emulation branch
 *     aims for (defaultcull.csh:258-265, depthpyramid.csh:22-24).
 */
#include "oracle.h"

#include <atomic>
#include <cmath>
#include <cstring>
#include <functional>
#include <limits>
#include <thread>
#include <vector>

#if defined(ORC_USE_TASKFLOW)
#include <taskflow/algorithm/for_each.hpp>
#include <taskflow/taskflow.hpp>
#endif

namespace {

// ---- GLSL scalar semantics -------------------------------------------------
inline float gmin(float x, float y) { return (y < x) ? y : x; }
inline float gmax(float x, float y) { return (x < y) ? y : x; }

struct vec3 {
    float x, y, z;
};
struct vec4 {
    float x, y, z, w;
};

inline vec3 operator+(vec3 a, vec3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
inline vec3 operator-(vec3 a, vec3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
inline vec3 operator*(vec3 a, float s) { return {a.x * s, a.y * s, a.z * s}; }
inline vec4 operator*(vec4 a, float s) { return {a.x * s, a.y * s, a.z * s, a.w * s}; }
inline vec4 operator+(vec4 a, vec4 b) { return {a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w}; }
// ORC_VARIANT_* build the OTHER evaluations GLSL allows (tools/contract_sensitivity.py counts how many decisions of a
// scene depend on the choice; the product is only ever compared with the default build):
//   ORC_VARIANT_NO_FMA        a*b+c rounded twice everywhere        ORC_VARIANT_DOT_REVERSED  dot accumulates from the last component
//   ORC_VARIANT_TRUE_DIV      vector / scalar divides each component (no reciprocal-multiply)
#if defined(ORC_VARIANT_NO_FMA)
inline float fma_(float a, float b, float c) {
    volatile float p = a * b;  // rounded product
    return p + c;
}
#else
inline float fma_(float a, float b, float c) { return __builtin_fmaf(a, b, c); }
#endif
inline float rcp_(float s) { return 1.0f / s; }
#if defined(ORC_VARIANT_DOT_REVERSED)
inline float dot(vec3 a, vec3 b) { return fma_(a.x, b.x, fma_(a.y, b.y, a.z * b.z)); }
inline float dot(vec4 a, vec4 b) { return fma_(a.x, b.x, fma_(a.y, b.y, fma_(a.z, b.z, a.w * b.w))); }
#else
inline float dot(vec3 a, vec3 b) { return fma_(a.z, b.z, fma_(a.y, b.y, a.x * b.x)); }
inline float dot(vec4 a, vec4 b) { return fma_(a.w, b.w, fma_(a.z, b.z, fma_(a.y, b.y, a.x * b.x))); }
#endif
inline float length(vec3 v) { return std::sqrt(dot(v, v)); }
inline vec3 fma3(vec3 a, float s, vec3 c) { return {fma_(a.x, s, c.x), fma_(a.y, s, c.y), fma_(a.z, s, c.z)}; }
inline vec4 fma4(vec4 a, float s, vec4 c) { return {fma_(a.x, s, c.x), fma_(a.y, s, c.y), fma_(a.z, s, c.z), fma_(a.w, s, c.w)}; }
// vector / scalar = vector * (1/scalar)
#if defined(ORC_VARIANT_TRUE_DIV)
inline vec3 vdiv(vec3 v, float s) { return {v.x / s, v.y / s, v.z / s}; }
inline vec4 vdiv(vec4 v, float s) { return {v.x / s, v.y / s, v.z / s, v.w / s}; }
#else
inline vec3 vdiv(vec3 v, float s) { return v * rcp_(s); }
inline vec4 vdiv(vec4 v, float s) { return v * rcp_(s); }
#endif
inline vec3 vmin(vec3 a, vec3 b) { return {gmin(a.x, b.x), gmin(a.y, b.y), gmin(a.z, b.z)}; }
inline vec3 vmax(vec3 a, vec3 b) { return {gmax(a.x, b.x), gmax(a.y, b.y), gmax(a.z, b.z)}; }

// column-major mat4: m[col*4+row]
struct mat4 {
    float m[16];
    vec4 col(int c) const { return {m[c * 4 + 0], m[c * 4 + 1], m[c * 4 + 2], m[c * 4 + 3]}; }
    vec3 col3(int c) const { return {m[c * 4 + 0], m[c * 4 + 1], m[c * 4 + 2]}; }
};
inline mat4 load_mat4(const float* p) {
    mat4 r;
    std::memcpy(r.m, p, 64);
    return r;
}
inline vec4 mul(const mat4& M, vec4 v) {
    return fma4(M.col(3), v.w, fma4(M.col(2), v.z, fma4(M.col(1), v.y, M.col(0) * v.x)));
}
// mat3(M) * v
inline vec3 mul3(const mat4& M, vec3 v) { return fma3(M.col3(2), v.z, fma3(M.col3(1), v.y, M.col3(0) * v.x)); }

// ---- parallel-for over chunks ------------------------------------------------
void parallel_chunks(uint64_t num_chunks, int nthreads, const std::function<void(uint64_t)>& fn) {
    if (nthreads <= 1 || num_chunks <= 1) {
        for (uint64_t c = 0; c < num_chunks; ++c) fn(c);
        return;
    }
#if defined(ORC_USE_TASKFLOW)
    // the reference's executor: tf::Executor + for_each_index over workgroup-sized chunks
    static thread_local std::unique_ptr<tf::Executor> exec;
    static thread_local int exec_threads = 0;
    if (!exec || exec_threads != nthreads) {
        exec = std::make_unique<tf::Executor>(nthreads);
        exec_threads = nthreads;
    }
    tf::Taskflow flow;
    // coarse blocks of chunks so the task count stays bounded
    const uint64_t blocks = std::min<uint64_t>(num_chunks, uint64_t(nthreads) * 64);
    flow.for_each_index(uint64_t(0), blocks, uint64_t(1), [&](uint64_t b) {
        const uint64_t lo = num_chunks * b / blocks, hi = num_chunks * (b + 1) / blocks;
        for (uint64_t c = lo; c < hi; ++c) fn(c);
    });
    exec->run(flow).wait();
#else
    std::atomic<uint64_t> next{0};
    const uint64_t grain = std::max<uint64_t>(1, num_chunks / (uint64_t(nthreads) * 64));
    auto worker = [&]() {
        for (;;) {
            uint64_t lo = next.fetch_add(grain, std::memory_order_relaxed);
            if (lo >= num_chunks) break;
            uint64_t hi = std::min(num_chunks, lo + grain);
            for (uint64_t c = lo; c < hi; ++c) fn(c);
        }
    };
    std::vector<std::thread> pool;
    pool.reserve(nthreads - 1);
    for (int t = 1; t < nthreads; ++t) pool.emplace_back(worker);
    worker();
    for (auto& t : pool) t.join();
#endif
}

// ---- pyramid -----------------------------------------------------------------
// floor(x) as a clamped texel index pair (clamp-to-edge); NaN / huge handled in float first
inline void footprint(float x, uint32_t n, uint32_t& i0, uint32_t& i1) {
    const float fx = std::floor(x);
    if (!(fx >= 0.0f)) {  // negative or NaN: both taps clamp to texel 0 (i0 = -1 -> 0, i0+1 = 0)
        i0 = 0;
        i1 = 0;
        return;
    }
    if (fx >= float(n - 1)) {
        i0 = n - 1;
        i1 = n - 1;
        return;
    }
    i0 = uint32_t(fx);
    i1 = i0 + 1;
}

inline float sample_min(const float* img, uint32_t w, uint32_t h, float u, float v) {
    const float x = fma_(u, float(w), -0.5f);
    const float y = fma_(v, float(h), -0.5f);
    uint32_t x0, x1, y0, y1;
    footprint(x, w, x0, x1);
    footprint(y, h, y0, y1);
    const float a = img[uint64_t(y0) * w + x0];
    const float b = img[uint64_t(y0) * w + x1];
    const float c = img[uint64_t(y1) * w + x0];
    const float d = img[uint64_t(y1) * w + x1];
    return gmin(gmin(a, b), gmin(c, d));
}

inline uint64_t level_offset(uint32_t dim, uint32_t level) {
    const uint64_t d0 = dim, dl = dim >> level;
    return 4 * (d0 * d0 - dl * dl) / 3;
}

// exact ceil(log2(v)) in float, from the exponent (no libm)
inline float ceil_log2_exact(float v) {
    if (std::isnan(v) || v < 0.0f) return std::numeric_limits<float>::quiet_NaN();
    if (v == 0.0f) return -std::numeric_limits<float>::infinity();
    if (std::isinf(v)) return v;
    int e;
    const float m = std::frexp(v, &e);  // v = m * 2^e, m in [0.5,1)
    return (m == 0.5f) ? float(e - 1) : float(e);
}

// ---- cull --------------------------------------------------------------------
struct ClipBoundingBoxResult {
    float minX, maxX, minY, maxY, referenceZ;
};

// defaultcull.csh:108-145
ClipBoundingBoxResult projectWorldSpaceSphere(vec3 center, float r, const mat4& viewProj) {
    const vec3 boxCorners[8] = {
        center + vec3{r, -r, -r}, center + vec3{-r, -r, -r}, center + vec3{-r, r, -r}, center + vec3{r, r, -r},
        center + vec3{r, -r, r},  center + vec3{-r, -r, r},  center + vec3{-r, r, r},  center + vec3{r, r, r},
    };
    ClipBoundingBoxResult result{1, 0, 1, 0, 0};
    for (int i = 0; i < 8; i++) {
        vec4 p = mul(viewProj, vec4{boxCorners[i].x, boxCorners[i].y, boxCorners[i].z, 1.0f});
        p = vdiv(p, p.w);  // projected /= projected.w
        result.minX = gmin(p.x, result.minX);
        result.maxX = gmax(p.x, result.maxX);
        result.minY = gmin(p.y, result.minY);
        result.maxY = gmax(p.y, result.maxY);
        result.referenceZ = gmax(p.z, result.referenceZ);
    }
    return result;
}

// defaultcull.csh:81-92
void MakeFrustumPlanes(const mat4& vp, vec4 planes[6]) {
    auto M = [&](int i, int r) { return vp.m[i * 4 + r]; };
    float* p[6] = {&planes[0].x, &planes[1].x, &planes[2].x, &planes[3].x, &planes[4].x, &planes[5].x};
    for (int i = 0; i < 4; ++i) p[0][i] = M(i, 3) + M(i, 0);
    for (int i = 0; i < 4; ++i) p[1][i] = M(i, 3) - M(i, 0);
    for (int i = 0; i < 4; ++i) p[2][i] = M(i, 3) + M(i, 1);
    for (int i = 0; i < 4; ++i) p[3][i] = M(i, 3) - M(i, 1);
    for (int i = 0; i < 4; ++i) p[4][i] = M(i, 3) + M(i, 2);
    for (int i = 0; i < 4; ++i) p[5][i] = M(i, 3) - M(i, 2);
    for (int i = 0; i < 6; ++i) {
        const float len = length(vec3{planes[i].x, planes[i].y, planes[i].z});
        planes[i] = vdiv(planes[i], len);
    }
}

inline float DistanceToPlane(vec4 plane, vec3 p) { return dot(vec4{p.x, p.y, p.z, 1.0f}, plane); }

// defaultcull.csh:71-78
inline float CullSphere(const vec4 planes[6], vec3 c, float r) {
    const float dist01 = gmin(DistanceToPlane(planes[0], c), DistanceToPlane(planes[1], c));
    const float dist23 = gmin(DistanceToPlane(planes[2], c), DistanceToPlane(planes[3], c));
    const float dist45 = gmin(DistanceToPlane(planes[4], c), DistanceToPlane(planes[5], c));
    return gmin(gmin(dist01, dist23), dist45) + r;
}

struct CullArgs {
    const RveCullingUBO* ubo;
    const RveCullingUBOinstance* cubos;
    const float* model;
    const uint32_t* layers;
    const uint16_t* flags;
    const float* pyr;
    uint32_t dim, levels;
    void* const* bindless;
    bool atomic_append;
};

// one invocation of defaultcull.csh main() (:148-313)
void cull_invocation(const CullArgs& a, uint32_t gid) {
    const RveCullingUBO& ubo = *a.ubo;
    uint32_t cuboInstance = ~0u;
    uint32_t runningTotal = 0;
    for (uint32_t i = 0; i < ubo.numCubos; i++) {
        runningTotal += a.cubos[i].numObjects;
        if (gid < runningTotal) {
            cuboInstance = i;
            break;
        }
    }
    if (cuboInstance == ~0u) return;
    const RveCullingUBOinstance cubo = a.cubos[cuboInstance];
    const uint32_t countBase = runningTotal - cubo.numObjects;
    const uint32_t currentEntity = gid - countBase;
    const uint32_t entityID =
        static_cast<const uint32_t*>(a.bindless[cubo.entityIDInputBufferBindlessHandle])[currentEntity];

    if ((ubo.cameraRenderLayers & a.layers[entityID]) == 0) return;

    const uint16_t attributeBitmask = a.flags[entityID];
    const bool skipFrustumCulling = !(attributeBitmask & 1);
    const bool skipOcclusionCulling = !(attributeBitmask & (1 << 1));
    const bool castsShadows = (attributeBitmask & (1 << 2)) != 0;

    const uint32_t isSingleInstanceMode = (ubo.singleInstanceModeAndShadowMode & 1) ? 1u : 0u;
    const bool isShadowMode = (ubo.singleInstanceModeAndShadowMode & (1 << 1)) != 0;
    const bool shouldConsider = !isShadowMode || (isShadowMode && castsShadows);
    if (!shouldConsider) return;

    const mat4 model = load_mat4(a.model + uint64_t(entityID) * 16);
    const mat4 viewProj = load_mat4(ubo.viewProj);

    vec4 planes[6];
    MakeFrustumPlanes(viewProj, planes);

    // modelNoTranslate * vec3(0,0,r) etc. folded to column * r, (model * vec4(0,0,0,1)).xyz to
    // column 3 (exact values of those products, see the arithmetic contract above)
    vec3 radvecs[3] = {model.col3(2) * cubo.radius, model.col3(1) * cubo.radius, model.col3(0) * cubo.radius};
    float radius = 0;
    for (int i = 0; i < 3; i++) radius = gmax(radius, length(radvecs[i]));
    const vec3 center = model.col3(3);

    bool isOnCamera = skipFrustumCulling ? true : (CullSphere(planes, center, radius) > 0);

    if (isOnCamera && !skipOcclusionCulling) {
        const ClipBoundingBoxResult projected = projectWorldSpaceSphere(center, radius, viewProj);
        const float mipDim = float(a.dim);
        const float bbwidth = (projected.maxX - projected.minX) * mipDim;
        const float bbheight = (projected.maxY - projected.minY) * mipDim;
        if (projected.referenceZ <= 1) {
            float miplevel = ceil_log2_exact(gmax(bbwidth, bbheight)) - 1;
            miplevel = gmax(0.0f, miplevel);
            const uint32_t maxLevel = a.levels;
            float minDepth = 1;
            if (miplevel > float(maxLevel - 1)) {
                minDepth = 0;
            } else {
                float ndc[4][2] = {{projected.minX, projected.maxY},
                                   {projected.maxX, projected.maxY},
                                   {projected.maxX, projected.minY},
                                   {projected.minX, projected.minY}};
                const uint32_t level = uint32_t(miplevel);
                const uint32_t d = a.dim >> level;
                const float* img = a.pyr + level_offset(a.dim, level);
                for (int i = 0; i < 4; i++) {
                    float u = (ndc[i][0] + 1) * 0.5f;
                    float v = (ndc[i][1] + 1) * 0.5f;
                    v = 1 - v;
                    const float depth = sample_min(img, d, d, u, v);
                    minDepth = gmin(minDepth, depth);
                }
            }
            isOnCamera = isOnCamera && projected.referenceZ >= minDepth;
        }
    }

    if (isOnCamera) {
        uint32_t lodID = 0;
        if (cubo.numLODs > 1) {
            const vec3 worldSpacePos = center;
            const vec3 C = worldSpacePos - vec3{ubo.camPos[0], ubo.camPos[1], ubo.camPos[2]};
            const float distFromCamera = dot(C, C);
            const float* lodDist = static_cast<const float*>(a.bindless[cubo.lodDistanceBufferBindlessHandle]);
            float currentBest = lodDist[0];
            for (uint32_t i = 0; i < cubo.numLODs; i++) {
                float m = lodDist[i];
                m *= m;
                if (distFromCamera > m && m > currentBest) {
                    lodID = i;
                    currentBest = m;
                }
            }
        }
        const uint32_t indirectBufferIdx = cubo.indirectBufferOffset + lodID + isSingleInstanceMode * currentEntity;
        auto* indirect =
            static_cast<RveIndirectIndexedCommand*>(a.bindless[cubo.indirectOutputBufferBindlessHandle]);
        uint32_t idx;
        if (a.atomic_append) {
            idx = __atomic_fetch_add(&indirect[indirectBufferIdx].instanceCount, 1u, __ATOMIC_RELAXED);
        } else {
            idx = indirect[indirectBufferIdx].instanceCount++;
        }
        const uint32_t idxLODOffset = cubo.numObjects * lodID + cubo.cullingBufferOffset;
        const uint32_t entityIDidx = idx + idxLODOffset + currentEntity * isSingleInstanceMode;
        static_cast<uint32_t*>(a.bindless[cubo.idOutputBufferBindlessHandle])[entityIDidx] = entityID;
    }
}

// ---- rng shared with the device generator ---------------------------------------
inline uint64_t splitmix64(uint64_t x) {
    x += 0x9E3779B97F4A7C15ull;
    uint64_t z = x;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
inline float u01(uint64_t seed, uint64_t idx, uint32_t k) {
    const uint64_t h = splitmix64(seed ^ splitmix64(idx * 16 + k));
    return float(uint32_t(h >> 40)) * (1.0f / 16777216.0f);
}

}  // namespace

extern "C" {

void orc_pyramid_dims(uint32_t width, uint32_t* dim, uint32_t* num_levels) {
    // src/DepthPyramid.cpp:10-17 — width = min(width,1024); dim = 2^floor(log2(width)); levels = 1+floor(log2(dim))
    uint32_t w = width < 1024 ? width : 1024;
    uint32_t d = 1, l = 0;
    while (d * 2 <= w) {
        d *= 2;
        l++;
    }
    *dim = d;
    *num_levels = 1 + l;
}

uint64_t orc_pyramid_texel_count(uint32_t dim) { return (4ull * dim * dim - 1) / 3; }
uint64_t orc_pyramid_level_offset(uint32_t dim, uint32_t level) { return level_offset(dim, level); }

float orc_sample_min(const float* image, uint32_t w, uint32_t h, float u, float v) {
    return sample_min(image, w, h, u, v);
}
float orc_ceil_log2(float v) { return ceil_log2_exact(v); }

void orc_pyramid_build(const float* depth, uint32_t width, uint32_t height, float* pyramid, uint32_t dim,
                       uint32_t num_levels) {
    // mip 0 — depthpyramidcopy.fsh:10-20: texcoord = gl_FragCoord.xy / targetDim, gl_FragCoord = pixel centre
    for (uint32_t y = 0; y < dim; ++y)
        for (uint32_t x = 0; x < dim; ++x) {
            const float u = (float(x) + 0.5f) / float(dim);
            const float v = (float(y) + 0.5f) / float(dim);
            pyramid[uint64_t(y) * dim + x] = sample_min(depth, width, height, u, v);
        }
    // mips 1.. — depthpyramid.csh:9-28: uv = (pos + 0.5) / dimOut sampled from the previous level
    for (uint32_t l = 1; l < num_levels; ++l) {
        const uint32_t din = dim >> (l - 1), dout = dim >> l;
        const float* src = pyramid + level_offset(dim, l - 1);
        float* dst = pyramid + level_offset(dim, l);
        for (uint32_t y = 0; y < dout; ++y)
            for (uint32_t x = 0; x < dout; ++x) {
                const float u = (float(x) + 0.5f) / float(dout);
                const float v = (float(y) + 0.5f) / float(dout);
                dst[uint64_t(y) * dout + x] = sample_min(src, din, din, u, v);
            }
    }
}

// the same build with rows dealt to worker threads (the CPU baseline uses every host thread the engine would)
void orc_pyramid_build_mt(const float* depth, uint32_t width, uint32_t height, float* pyramid, uint32_t dim, uint32_t num_levels,
                          int nthreads) {
    parallel_chunks(dim, nthreads, [&](uint64_t y) {
        for (uint32_t x = 0; x < dim; ++x) {
            const float u = (float(x) + 0.5f) / float(dim);
            const float v = (float(y) + 0.5f) / float(dim);
            pyramid[uint64_t(y) * dim + x] = sample_min(depth, width, height, u, v);
        }
    });
    for (uint32_t l = 1; l < num_levels; ++l) {
        const uint32_t din = dim >> (l - 1), dout = dim >> l;
        const float* src = pyramid + level_offset(dim, l - 1);
        float* dst = pyramid + level_offset(dim, l);
        parallel_chunks(dout, dout >= 64 ? nthreads : 1, [&](uint64_t y) {
            for (uint32_t x = 0; x < dout; ++x) {
                const float u = (float(x) + 0.5f) / float(dout);
                const float v = (float(y) + 0.5f) / float(dout);
                dst[uint64_t(y) * dout + x] = sample_min(src, din, din, u, v);
            }
        });
    }
}

void orc_cull(const RveCullingUBO* ubo, const RveCullingUBOinstance* cubos, const float* model_matrices,
              const uint32_t* render_layers, const uint16_t* per_object_flags, const float* pyramid,
              uint32_t pyr_dim, uint32_t pyr_levels, void* const* bindless, uint32_t num_groups_x, int nthreads) {
    CullArgs a{ubo,     cubos,      model_matrices, render_layers, per_object_flags,
               pyramid, pyr_dim,    pyr_levels,     bindless,      nthreads > 1};
    // local_size_x = 64 (defaultcull.csh:147): one chunk = one workgroup
    parallel_chunks(num_groups_x, nthreads, [&](uint64_t group) {
        for (uint32_t t = 0; t < 64; ++t) cull_invocation(a, uint32_t(group * 64 + t));
    });
}

void orc_skin(const RveSkinningUBO* ubo, float* out_pos, float* out_nrm, float* out_tan, float* out_bitan,
              float* out_uv, const float* in_pos, const float* in_nrm, const float* in_tan, const float* in_bitan,
              const float* in_uv, const float* pose, const RveVertexJointBinding* weights, int nthreads) {
    const RveSkinningUBO u = *ubo;
    // skinning_cs.csh:95-138; invocation (x = object, y = vertex)
    parallel_chunks(u.numObjects, nthreads, [&](uint64_t objID) {
        for (uint32_t vertID = 0; vertID < u.numVertices; ++vertID) {
            const uint32_t bone_begin = u.numBones * uint32_t(objID) + u.boneReadOffset;
            float total[16];
            for (int k = 0; k < 16; ++k) total[k] = 0.0f;
            const RveVertexJointBinding wv = weights[vertID];
            for (int i = 0; i < 4; ++i) {
                const uint32_t joint_idx = wv.weights[i].joint_index;
                const float weight = wv.weights[i].influence;
                const float* posed = pose + uint64_t(bone_begin + joint_idx) * 16;
                for (int k = 0; k < 16; ++k) total[k] = fma_(weight, posed[k], total[k]);  // totalmtx += weight * posed_mtx
            }
            const mat4 M = load_mat4(total);
            const uint64_t writeOffset = uint64_t(vertID) + uint64_t(objID) * u.numVertices + u.vertexWriteOffset;
            const uint64_t readOffset = uint64_t(vertID) + u.vertexReadOffset;
            out_uv[writeOffset * 2 + 0] = in_uv[readOffset * 2 + 0];
            out_uv[writeOffset * 2 + 1] = in_uv[readOffset * 2 + 1];
            const vec4 p = mul(M, vec4{in_pos[readOffset * 3], in_pos[readOffset * 3 + 1], in_pos[readOffset * 3 + 2], 1.0f});
            out_pos[writeOffset * 3 + 0] = p.x;
            out_pos[writeOffset * 3 + 1] = p.y;
            out_pos[writeOffset * 3 + 2] = p.z;
            const float* ins[3] = {in_nrm, in_tan, in_bitan};
            float* outs[3] = {out_nrm, out_tan, out_bitan};
            for (int s = 0; s < 3; ++s) {
                const vec3 r = mul3(M, vec3{ins[s][readOffset * 3], ins[s][readOffset * 3 + 1], ins[s][readOffset * 3 + 2]});
                outs[s][writeOffset * 3 + 0] = r.x;
                outs[s][writeOffset * 3 + 1] = r.y;
                outs[s][writeOffset * 3 + 2] = r.z;
            }
        }
    });
}

void orc_skin_f64(const RveSkinningUBO* ubo, double* out_pos, double* out_nrm, double* out_tan, double* out_bitan,
                  const float* in_pos, const float* in_nrm, const float* in_tan, const float* in_bitan,
                  const float* pose, const RveVertexJointBinding* weights) {
    const RveSkinningUBO u = *ubo;
    for (uint32_t objID = 0; objID < u.numObjects; ++objID)
        for (uint32_t vertID = 0; vertID < u.numVertices; ++vertID) {
            const uint32_t bone_begin = u.numBones * objID + u.boneReadOffset;
            double M[16] = {0};
            for (int i = 0; i < 4; ++i) {
                const float* posed = pose + uint64_t(bone_begin + weights[vertID].weights[i].joint_index) * 16;
                for (int k = 0; k < 16; ++k) M[k] += double(weights[vertID].weights[i].influence) * double(posed[k]);
            }
            const uint64_t wo = uint64_t(vertID) + uint64_t(objID) * u.numVertices + u.vertexWriteOffset;
            const uint64_t ro = uint64_t(vertID) + u.vertexReadOffset;
            const float* ins[4] = {in_pos, in_nrm, in_tan, in_bitan};
            double* outs[4] = {out_pos, out_nrm, out_tan, out_bitan};
            for (int s = 0; s < 4; ++s) {
                const double x = ins[s][ro * 3], y = ins[s][ro * 3 + 1], z = ins[s][ro * 3 + 2];
                const double w = (s == 0) ? 1.0 : 0.0;
                for (int r = 0; r < 3; ++r)
                    outs[s][wo * 3 + r] = M[0 * 4 + r] * x + M[1 * 4 + r] * y + M[2 * 4 + r] * z + M[3 * 4 + r] * w;
            }
        }
}

void orc_skinned_drawcall_prepare(const RveSkinningPrepareUBO* ubo, RveIndirectIndexedCommand* commands) {
    // skinned_mesh_drawcall.csh:25-39
    for (uint32_t objectID = 0; objectID < ubo->nTotalObjects; ++objectID) {
        commands[ubo->drawCallBufferOffset + objectID] = RveIndirectIndexedCommand{
            ubo->nIndicesInThisMesh, 0, ubo->indexBufferOffset + ubo->nIndicesInThisMesh * objectID,
            ubo->vertexBufferOffset + ubo->nVerticesInThisMesh * objectID, ubo->baseInstanceOffset + objectID};
    }
}

void orc_cluster_build(const RveGridBuildUBO* ubo_p, RveCluster* clusters) {
    const RveGridBuildUBO ubo = *ubo_p;
    const mat4 invProj = load_mat4(ubo.invProj);
    // cluster_build_grid.csh:16-26
    auto screenToView = [&](float sx, float sy) -> vec3 {
        const vec4 ndc{sx / float(ubo.screenDim[0]) * 2.0f - 1.0f, sy / float(ubo.screenDim[1]) * 2.0f - 1.0f, 0.0f, 1.0f};
        vec4 v = mul(invProj, ndc);
        v = vdiv(v, v.w);  // view / view.w
        return {v.x, v.y, v.z};
    };
    // cluster_build_grid.csh:29-38
    auto lineIntersectionWithZPlane = [](vec3 startPoint, vec3 endPoint, float zDistance) -> vec3 {
        const vec3 direction = endPoint - startPoint;
        const vec3 normal{0.0f, 0.0f, -1.0f};
        const float t = (zDistance - dot(normal, startPoint)) / dot(normal, direction);
        return fma3(direction, t, startPoint);  // startPoint + t * direction
    };
    const vec3 eyePos{0, 0, 0};
    for (uint32_t gz = 0; gz < ubo.gridSize[2]; ++gz)
        for (uint32_t gy = 0; gy < ubo.gridSize[1]; ++gy)
            for (uint32_t gx = 0; gx < ubo.gridSize[0]; ++gx) {
                const uint32_t tileIndex = gx + (gy * ubo.gridSize[0]) + (gz * ubo.gridSize[0] * ubo.gridSize[1]);
                // vec2 tileSize = ubo.screenDimensions / ubo.gridSize.xy: uvec2 / uvec2 is an INTEGER
                // division in GLSL, converted to float afterwards (cluster_build_grid.csh:48)
                const float tileSizeX = float(ubo.screenDim[0] / ubo.gridSize[0]);
                const float tileSizeY = float(ubo.screenDim[1] / ubo.gridSize[1]);
                const vec3 minPoint_viewSpace = screenToView(float(gx) * tileSizeX, float(gy) * tileSizeY);
                const vec3 maxPoint_viewSpace = screenToView(float(gx + 1) * tileSizeX, float(gy + 1) * tileSizeY);
                // pow() differs between GPU vendors and libm; the slice depths are evaluated
                // with the host libm powf on both sides (DESIGN.md "definitional choices")
                const float tileNear = ubo.zFar * std::pow(ubo.zNear / ubo.zFar, float(gz) / float(ubo.gridSize[2]));
                const float tileFar = ubo.zFar * std::pow(ubo.zNear / ubo.zFar, float(gz + 1) / float(ubo.gridSize[2]));
                const vec3 minPointNear = lineIntersectionWithZPlane(eyePos, minPoint_viewSpace, tileNear);
                const vec3 minPointFar = lineIntersectionWithZPlane(eyePos, minPoint_viewSpace, tileFar);
                const vec3 maxPointNear = lineIntersectionWithZPlane(eyePos, maxPoint_viewSpace, tileNear);
                const vec3 maxPointFar = lineIntersectionWithZPlane(eyePos, maxPoint_viewSpace, tileFar);
                const vec3 mn = vmin(minPointNear, minPointFar);
                const vec3 mx = vmax(maxPointNear, maxPointFar);
                clusters[tileIndex].minPoint[0] = mn.x;
                clusters[tileIndex].minPoint[1] = mn.y;
                clusters[tileIndex].minPoint[2] = mn.z;
                clusters[tileIndex].maxPoint[0] = mx.x;
                clusters[tileIndex].maxPoint[1] = mx.y;
                clusters[tileIndex].maxPoint[2] = mx.z;
            }
}

void orc_cluster_assign(const RveGridAssignUBO* ubo_p, RveCluster* clusters, const RvePointLight* points,
                        const RveSpotLight* spots, uint32_t num_clusters, int nthreads) {
    const RveGridAssignUBO ubo = *ubo_p;
    const mat4 viewMatrix = load_mat4(ubo.viewMat);
    // cluster_assign_lights.csh:28-35
    auto sphereAABBIntersection = [](vec3 center, float radius, vec3 aabbMin, vec3 aabbMax) {
        const vec3 closestPoint = vmin(vmax(center, aabbMin), aabbMax);
        const vec3 d = closestPoint - center;
        const float distanceSquared = dot(d, d);
        return distanceSquared <= radius * radius;
    };
    // cluster_shared.glsl:41-44
    auto getPointLightRadius = [](float intensity) { return std::sqrt(intensity / RVE_LIGHT_MIN_INFLUENCE); };
    const uint64_t chunks = (uint64_t(num_clusters) + 127) / 128;  // LOCAL_SIZE 128
    parallel_chunks(chunks, nthreads, [&](uint64_t chunk) {
        for (uint32_t index = uint32_t(chunk * 128); index < num_clusters && index < (chunk + 1) * 128; ++index) {
            RveCluster cluster = clusters[index];
            cluster.pointLightCount = 0;
            cluster.spotLightCount = 0;
            const vec3 aabbMin{cluster.minPoint[0], cluster.minPoint[1], cluster.minPoint[2]};
            const vec3 aabbMax{cluster.maxPoint[0], cluster.maxPoint[1], cluster.maxPoint[2]};
            for (uint32_t i = 0; i < ubo.pointLightCount; ++i) {
                const vec4 c4 = mul(viewMatrix, vec4{points[i].position[0], points[i].position[1], points[i].position[2], 1.0f});
                const float radius = getPointLightRadius(points[i].intensity);
                if (sphereAABBIntersection({c4.x, c4.y, c4.z}, radius, aabbMin, aabbMax) &&
                    cluster.pointLightCount < RVE_CLUSTER_MAX_POINTS) {
                    cluster.pointLightIndices[cluster.pointLightCount] = i;
                    cluster.pointLightCount++;
                }
            }
            for (uint32_t i = 0; i < ubo.spotLightCount; ++i) {
                const mat4 wt = load_mat4(spots[i].worldTransform);
                const vec3 wpos = wt.col3(3);  // (worldTransform * vec4(0,0,0,1)).xyz, folded (arithmetic contract)
                const vec4 c4 = mul(viewMatrix, vec4{wpos.x, wpos.y, wpos.z, 1.0f});
                const float radius = getPointLightRadius(spots[i].intensity);
                // the shader has no capacity check here (cluster_assign_lights.csh:80-83; writes
                // past spotLightIndices[16] are out of bounds in the reference). We define the
                // result as capped at CLUSTER_MAX_SPOTS, like the point-light loop.
                if (sphereAABBIntersection({c4.x, c4.y, c4.z}, radius, aabbMin, aabbMax) &&
                    cluster.spotLightCount < RVE_CLUSTER_MAX_SPOTS) {
                    cluster.spotLightIndices[cluster.spotLightCount] = i;
                    cluster.spotLightCount++;
                }
            }
            clusters[index] = cluster;
        }
    });
}

void orc_gen_instances(const RveSceneGenParams* pp, uint64_t first_index, uint32_t count, float* model,
                       uint32_t* layers, uint16_t* flags, uint32_t* ids) {
    const RveSceneGenParams p = *pp;
    for (uint32_t i = 0; i < count; ++i) {
        const uint64_t g = first_index + i;
        const uint32_t id = p.id_base + i;
        const float px = (2.0f * u01(p.seed, g, 0) - 1.0f) * p.extent_xy;
        const float pz = (2.0f * u01(p.seed, g, 1) - 1.0f) * p.extent_xy;
        const float py = (2.0f * u01(p.seed, g, 2) - 1.0f) * p.extent_z;
        float qx = 2.0f * u01(p.seed, g, 3) - 1.0f;
        float qy = 2.0f * u01(p.seed, g, 4) - 1.0f;
        float qz = 2.0f * u01(p.seed, g, 5) - 1.0f;
        float qw = 2.0f * u01(p.seed, g, 6) - 1.0f;
        const float n2 = ((qx * qx + qy * qy) + qz * qz) + qw * qw;
        if (n2 < 1e-6f) {
            qx = qy = qz = 0.0f;
            qw = 1.0f;
        } else {
            const float n = std::sqrt(n2);
            qx = qx / n;
            qy = qy / n;
            qz = qz / n;
            qw = qw / n;
        }
        const float s = p.scale_min + (p.scale_max - p.scale_min) * u01(p.seed, g, 7);
        const float xx = qx * qx, yy = qy * qy, zz = qz * qz;
        const float xy = qx * qy, xz = qx * qz, yz = qy * qz;
        const float wx = qw * qx, wy = qw * qy, wz = qw * qz;
        float* m = model + uint64_t(id) * 16;
        m[0] = (1.0f - 2.0f * (yy + zz)) * s;
        m[1] = (2.0f * (xy + wz)) * s;
        m[2] = (2.0f * (xz - wy)) * s;
        m[3] = 0.0f;
        m[4] = (2.0f * (xy - wz)) * s;
        m[5] = (1.0f - 2.0f * (xx + zz)) * s;
        m[6] = (2.0f * (yz + wx)) * s;
        m[7] = 0.0f;
        m[8] = (2.0f * (xz + wy)) * s;
        m[9] = (2.0f * (yz - wx)) * s;
        m[10] = (1.0f - 2.0f * (xx + yy)) * s;
        m[11] = 0.0f;
        m[12] = px;
        m[13] = py;
        m[14] = pz;
        m[15] = 1.0f;
        layers[id] = (u01(p.seed, g, 8) < p.frac_layer_zero) ? 0u : 0xFFFFFFFFu;
        flags[id] = (u01(p.seed, g, 9) < p.frac_no_cull_bits) ? uint16_t(0xFFFC) : uint16_t(0xFFFF);
        ids[i] = id;
    }
}

int orc_default_workers(void) {
    // include/RavEngine/App.hpp:112-118
    const int hc = int(std::thread::hardware_concurrency());
    return std::max(hc - 2, 2);
}

int orc_uses_taskflow(void) {
#if defined(ORC_USE_TASKFLOW)
    return 1;
#else
    return 0;
#endif
}

}  // extern "C"
