// csm.cu — cascaded-shadow-map matrices: the per (camera, directional light, cascade) frustum fit of
// src/RenderEngine_Draw.cpp:149-293, i.e. the inputs of the shadow culls. The reference computes them serially on
// the CPU with glm and uploads them; here one thread per (camera, light, cascade) writes them on the device, so a
// frame's cascades can feed rve_cull_views without a host round trip.
//
// f32 arithmetic in glm's own operation order (deps/glm 0.9.9: perspectiveRH_ZO / orthoRH_ZO with swapped near-far
// = reverse Z, include/RavEngine/mathtypes.hpp:141-149; compute_inverse<4,4>; lookAtRH; mat*mat and mat*vec as
// column sums), no contraction (-fmad=false). tan() is the one library call: the caller passes tan(fovy/2).
// Quirks preserved: the cascade frustum uses 1.5x the camera's fov (:219) and the INTEGER ratio width/height as
// aspect (:219); the output slot is passIndex + lightIndex with passIndex advancing by ONE per camera (:278), so
// slots collide when there are several lights and several cameras; cascadeDistances is per light, last camera wins.
#include <cmath>

#include "internal.h"

namespace rve {

struct V4 {
    float x, y, z, w;
};
struct M4 {
    V4 c[4];  // columns
};

__device__ __forceinline__ V4 mulvs(V4 a, float s) { return V4{a.x * s, a.y * s, a.z * s, a.w * s}; }
__device__ __forceinline__ V4 addv(V4 a, V4 b) { return V4{a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w}; }
// glm mat4 * vec4 (type_mat4x4.inl): (m0*v.x + m1*v.y) + (m2*v.z + m3*v.w)
__device__ __forceinline__ V4 mulmv(const M4& m, V4 v) {
    return addv(addv(mulvs(m.c[0], v.x), mulvs(m.c[1], v.y)), addv(mulvs(m.c[2], v.z), mulvs(m.c[3], v.w)));
}
// glm mat4 * mat4: column j = ((A0*B[j].x + A1*B[j].y) + A2*B[j].z) + A3*B[j].w
__device__ __forceinline__ M4 mulmm(const M4& a, const M4& b) {
    M4 r;
#pragma unroll
    for (int j = 0; j < 4; ++j)
        r.c[j] = addv(addv(addv(mulvs(a.c[0], b.c[j].x), mulvs(a.c[1], b.c[j].y)), mulvs(a.c[2], b.c[j].z)), mulvs(a.c[3], b.c[j].w));
    return r;
}
__device__ __forceinline__ float el(const M4& m, int c, int r) { return (&m.c[c].x)[r]; }

// glm::inverse, compute_inverse<4,4> (detail/func_matrix.inl:294-352)
__device__ M4 inverse4(const M4& m) {
#define E(c, r) el(m, c, r)
    const float c00 = E(2, 2) * E(3, 3) - E(3, 2) * E(2, 3), c02 = E(1, 2) * E(3, 3) - E(3, 2) * E(1, 3), c03 = E(1, 2) * E(2, 3) - E(2, 2) * E(1, 3);
    const float c04 = E(2, 1) * E(3, 3) - E(3, 1) * E(2, 3), c06 = E(1, 1) * E(3, 3) - E(3, 1) * E(1, 3), c07 = E(1, 1) * E(2, 3) - E(2, 1) * E(1, 3);
    const float c08 = E(2, 1) * E(3, 2) - E(3, 1) * E(2, 2), c10 = E(1, 1) * E(3, 2) - E(3, 1) * E(1, 2), c11 = E(1, 1) * E(2, 2) - E(2, 1) * E(1, 2);
    const float c12 = E(2, 0) * E(3, 3) - E(3, 0) * E(2, 3), c14 = E(1, 0) * E(3, 3) - E(3, 0) * E(1, 3), c15 = E(1, 0) * E(2, 3) - E(2, 0) * E(1, 3);
    const float c16 = E(2, 0) * E(3, 2) - E(3, 0) * E(2, 2), c18 = E(1, 0) * E(3, 2) - E(3, 0) * E(1, 2), c19 = E(1, 0) * E(2, 2) - E(2, 0) * E(1, 2);
    const float c20 = E(2, 0) * E(3, 1) - E(3, 0) * E(2, 1), c22 = E(1, 0) * E(3, 1) - E(3, 0) * E(1, 1), c23 = E(1, 0) * E(2, 1) - E(2, 0) * E(1, 1);
    const V4 f0{c00, c00, c02, c03}, f1{c04, c04, c06, c07}, f2{c08, c08, c10, c11}, f3{c12, c12, c14, c15}, f4{c16, c16, c18, c19},
        f5{c20, c20, c22, c23};
    const V4 v0{E(1, 0), E(0, 0), E(0, 0), E(0, 0)}, v1{E(1, 1), E(0, 1), E(0, 1), E(0, 1)}, v2{E(1, 2), E(0, 2), E(0, 2), E(0, 2)},
        v3{E(1, 3), E(0, 3), E(0, 3), E(0, 3)};
#undef E
    auto mulvv = [](V4 a, V4 b) { return V4{a.x * b.x, a.y * b.y, a.z * b.z, a.w * b.w}; };
    auto subv = [](V4 a, V4 b) { return V4{a.x - b.x, a.y - b.y, a.z - b.z, a.w - b.w}; };
    const V4 i0 = addv(subv(mulvv(v1, f0), mulvv(v2, f1)), mulvv(v3, f2));
    const V4 i1 = addv(subv(mulvv(v0, f0), mulvv(v2, f3)), mulvv(v3, f4));
    const V4 i2 = addv(subv(mulvv(v0, f1), mulvv(v1, f3)), mulvv(v3, f5));
    const V4 i3 = addv(subv(mulvv(v0, f2), mulvv(v1, f4)), mulvv(v2, f5));
    const V4 sa{1.f, -1.f, 1.f, -1.f}, sb{-1.f, 1.f, -1.f, 1.f};
    M4 inv{{mulvv(i0, sa), mulvv(i1, sb), mulvv(i2, sa), mulvv(i3, sb)}};
    const V4 row0{inv.c[0].x, inv.c[1].x, inv.c[2].x, inv.c[3].x};
    const V4 d0 = mulvv(m.c[0], row0);
    const float det = (d0.x + d0.y) + (d0.z + d0.w);
    const float ood = 1.0f / det;
    for (int j = 0; j < 4; ++j) inv.c[j] = mulvs(inv.c[j], ood);
    return inv;
}

struct V3 {
    float x, y, z;
};
__device__ __forceinline__ float dot3v(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }  // glm: (x*x + y*y) + z*z
__device__ __forceinline__ V3 cross3(V3 x, V3 y) { return V3{x.y * y.z - y.y * x.z, x.z * y.x - y.z * x.x, x.x * y.y - y.x * x.y}; }
__device__ __forceinline__ V3 normalize3(V3 v) {
    const float s = 1.0f / sqrtf(dot3v(v, v));  // inversesqrt
    return V3{v.x * s, v.y * s, v.z * s};
}

__global__ void csm_cascades_kernel(const RveCsmCamera* __restrict__ cams, uint32_t nCams, const RveCsmLight* __restrict__ lights,
                                    uint32_t nLights, float* __restrict__ outViewProj, float* __restrict__ outView,
                                    float* __restrict__ outProj, float* __restrict__ outDist) {
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t index = t & 3u, li = (t >> 2) % nLights, ci = (t >> 2) / nLights;
    if (ci >= nCams) return;
    const RveCsmLight L = lights[li];
    if (!L.castsShadows || index >= L.numCascades) return;
    // The reference loop is serial and slot passIndex + i is shared by every (camera, light) pair with the same sum:
    // the pair visited last (the largest camera index) owns the slot's cascade. Decide it up front, no write races.
    const uint32_t slot = ci + li;  // passIndex + i, passIndex advancing by one per camera (:280, :289)
    for (uint32_t c2 = ci + 1; c2 < nCams && c2 <= slot; ++c2) {
        const RveCsmLight& o = lights[slot - c2];
        if (o.castsShadows && index < o.numCascades) return;
    }
    const RveCsmCamera C = cams[ci];
    // near / far of the cascade (:206-216): glm::mix(a, b, t) = a*(1-t) + b*t
    float nearC = C.zNear, farC = C.zFar;
    if (index > 0) nearC = C.zNear * (1.0f - L.shadowCascades[index - 1]) + C.zFar * L.shadowCascades[index - 1];
    const uint32_t numCascades = L.numCascades < 4u ? L.numCascades : 4u;
    if (index < numCascades - 1) farC = C.zNear * (1.0f - L.shadowCascades[index]) + C.zFar * L.shadowCascades[index];
    farC *= L.maxShadowDistanceFactor;
    // RMath::perspectiveProjection = perspectiveRH_ZO(fovy, aspect, zFar, zNear) (:219)
    M4 proj{};
    {
        const float zn = farC, zf = nearC;  // swapped
        proj.c[0].x = 1.0f / (C.aspect * C.tanHalfFovy);
        proj.c[1].y = 1.0f / C.tanHalfFovy;
        proj.c[2].z = zf / (zn - zf);
        proj.c[2].w = -1.0f;
        proj.c[3].z = -(zf * zn) / (zf - zn);
    }
    M4 view;
    for (int j = 0; j < 4; ++j) view.c[j] = V4{C.viewOnly[4 * j], C.viewOnly[4 * j + 1], C.viewOnly[4 * j + 2], C.viewOnly[4 * j + 3]};
    const M4 inv = inverse4(mulmm(proj, view));
    // frustum corners in world space (:182-203) and their centre
    V4 corners[8];
    V3 center{0.f, 0.f, 0.f};
    int k = 0;
    for (int x = 0; x < 2; ++x)
        for (int y = 0; y < 2; ++y)
            for (int z = 0; z < 2; ++z) {
                const V4 pt = mulmv(inv, V4{2.0f * float(x) - 1.0f, 2.0f * float(y) - 1.0f, float(z), 1.0f});
                corners[k++] = V4{pt.x / pt.w, pt.y / pt.w, pt.z / pt.w, pt.w / pt.w};
            }
    for (int i = 0; i < 8; ++i) {
        center.x += corners[i].x;
        center.y += corners[i].y;
        center.z += corners[i].z;
    }
    center = V3{center.x / 8.0f, center.y / 8.0f, center.z / 8.0f};
    // lookAtRH(center + dir, center, up) (:231-235)
    const V3 eye{center.x + L.direction[0], center.y + L.direction[1], center.z + L.direction[2]};
    const V3 f = normalize3(V3{center.x - eye.x, center.y - eye.y, center.z - eye.z});
    const V3 s = normalize3(cross3(f, V3{0.f, 1.f, 0.f}));
    const V3 u = cross3(s, f);
    M4 lv{};
    lv.c[0] = V4{s.x, u.x, -f.x, 0.f};
    lv.c[1] = V4{s.y, u.y, -f.y, 0.f};
    lv.c[2] = V4{s.z, u.z, -f.z, 0.f};
    lv.c[3] = V4{-dot3v(s, eye), -dot3v(u, eye), dot3v(f, eye), 1.0f};
    float minX = 3.402823466e38f, maxX = -3.402823466e38f, minY = minX, maxY = maxX, minZ = minX, maxZ = maxX;
    for (int i = 0; i < 8; ++i) {
        const V4 trf = mulmv(lv, corners[i]);
        // std::min(a, b) = (b < a) ? b : a and std::max(a, b) = (a < b) ? b : a, signed zeros included
        minX = trf.x < minX ? trf.x : minX;
        maxX = maxX < trf.x ? trf.x : maxX;
        minY = trf.y < minY ? trf.y : minY;
        maxY = maxY < trf.y ? trf.y : maxY;
        minZ = trf.z < minZ ? trf.z : minZ;
        maxZ = maxZ < trf.z ? trf.z : maxZ;
    }
    const float zMult = 10.0f;  // :254-270
    minZ = minZ < 0 ? minZ * zMult : minZ / zMult;
    maxZ = maxZ < 0 ? maxZ / zMult : maxZ * zMult;
    // RMath::orthoProjection = orthoRH_ZO(l, r, b, t, zFar, zNear) with (zNear, zFar) = (minZ, maxZ)  (:275)
    M4 lp{};
    {
        const float zn = maxZ, zf = minZ;
        lp.c[0].x = 2.0f / (maxX - minX);
        lp.c[1].y = 2.0f / (maxY - minY);
        lp.c[2].z = -1.0f / (zf - zn);
        lp.c[3] = V4{-(maxX + minX) / (maxX - minX), -(maxY + minY) / (maxY - minY), -zn / (zf - zn), 1.0f};
    }
    const M4 lvp = mulmm(lp, lv);
    auto store = [&](float* dst, const M4& m) {
        if (!dst) return;
        float* o = dst + (size_t(slot) * 4 + index) * 16;
        for (int j = 0; j < 4; ++j) {
            o[4 * j] = m.c[j].x;
            o[4 * j + 1] = m.c[j].y;
            o[4 * j + 2] = m.c[j].z;
            o[4 * j + 3] = m.c[j].w;
        }
    };
    store(outViewProj, lvp);
    store(outView, lv);
    store(outProj, lp);
    if (outDist && ci == nCams - 1) outDist[li * 4 + index] = farC;  // light.cascadeDistances[index] = far: the last camera's value stays
}

}  // namespace rve

using namespace rve;

// Host half of the camera record: what src/RenderEngine_Draw.cpp:220 and perspectiveRH_ZO derive from the camera's
// fov and target size. fov * 1.5 is a double product narrowed to float (deg_to_rad takes decimalType = float),
// glm::radians multiplies by the float constant, tan is libm's tanf — the same call the reference makes.
extern "C" void rve_csm_camera_params(float fov_degrees, uint32_t target_width, uint32_t target_height, float* tan_half_fovy,
                                      float* aspect) {
    const float widened = float(double(fov_degrees) * 1.5);
    const float fovy = widened * 0.01745329251994329576923690768489f;
    if (tan_half_fovy) *tan_half_fovy = std::tan(fovy / 2.0f);
    if (aspect) *aspect = target_height ? float(target_width / target_height) : 0.0f;
}

extern "C" int rve_csm_cascades(rve_ctx* ctx, rve_buffer cameras, uint32_t num_cameras, rve_buffer lights, uint32_t num_lights,
                                rve_buffer out_view_proj, rve_buffer out_view, rve_buffer out_proj, rve_buffer out_distances) {
    if (!ctx) return set_error(RVE_ERR_INVALID_ARGUMENT, "null context");
    if (!num_cameras || !num_lights) return 0;
    DeviceGuard g(ctx->device);
    BufferRec* bc = get_buffer(ctx, cameras, "csm cameras");
    BufferRec* bl = get_buffer(ctx, lights, "csm lights");
    BufferRec* bo = get_buffer(ctx, out_view_proj, "csm lightViewProj");
    if (!bc || !bl || !bo) return RVE_ERR_BAD_HANDLE;
    const size_t slots = size_t(num_cameras) * num_lights;  // numVaryingElts (:151-156)
    RVE_REQUIRE(bc->bytes >= size_t(num_cameras) * sizeof(RveCsmCamera) && bl->bytes >= size_t(num_lights) * sizeof(RveCsmLight),
                "camera / light buffers too small");
    RVE_REQUIRE(bo->bytes >= slots * 4 * 64, "lightViewProj buffer smaller than cameras x lights x 4 matrices");
    float *pv = nullptr, *pp = nullptr, *pd = nullptr;
    if (out_view != RVE_BUFFER_INVALID) {
        BufferRec* b = get_buffer(ctx, out_view, "csm lightview");
        if (!b) return RVE_ERR_BAD_HANDLE;
        RVE_REQUIRE(b->bytes >= slots * 4 * 64, "lightview buffer too small");
        pv = static_cast<float*>(b->ptr);
    }
    if (out_proj != RVE_BUFFER_INVALID) {
        BufferRec* b = get_buffer(ctx, out_proj, "csm lightProj");
        if (!b) return RVE_ERR_BAD_HANDLE;
        RVE_REQUIRE(b->bytes >= slots * 4 * 64, "lightProj buffer too small");
        pp = static_cast<float*>(b->ptr);
    }
    if (out_distances != RVE_BUFFER_INVALID) {
        BufferRec* b = get_buffer(ctx, out_distances, "csm cascadeDistances");
        if (!b) return RVE_ERR_BAD_HANDLE;
        RVE_REQUIRE(b->bytes >= size_t(num_lights) * 16, "cascadeDistances buffer too small");
        pd = static_cast<float*>(b->ptr);
    }
    const uint32_t threads = num_cameras * num_lights * 4;
    csm_cascades_kernel<<<(threads + 127) / 128, 128, 0, ctx->stream>>>(static_cast<const RveCsmCamera*>(bc->ptr), num_cameras,
                                                                       static_cast<const RveCsmLight*>(bl->ptr), num_lights,
                                                                       static_cast<float*>(bo->ptr), pv, pp, pd);
    ctx->launches++;
    RVE_CUDA(cudaGetLastError());
    return 0;
}
