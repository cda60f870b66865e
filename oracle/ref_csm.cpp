// ref_csm.cpp — TEST INFRASTRUCTURE. The cascaded-shadow-map fit of src/RenderEngine_Draw.cpp:149-293 driven through
// the reference's OWN vendored glm 0.9.9.8 (headers used where they lie under /root/reference/deps/glm, nothing is
// copied): every matrix operation below is a glm call, so its outputs pin oracle/csm.py (the numpy restatement of
// glm's arithmetic) and, through it, the CUDA kernel. Built by oracle/Makefile into oracle/_ref/libref_csm.so only
// where /root/reference exists; tests/golden/make_golden_csm.py turns its outputs into tests/golden/csm.npz.
#define GLM_FORCE_DEPTH_ZERO_TO_ONE 1  // src/RenderEngine_Draw.cpp:2
#include <algorithm>
#include <array>
#include <cstdint>
#include <cstring>
#include <glm/glm.hpp>
#include <glm/gtc/matrix_transform.hpp>
#include <limits>

struct RefCsmCamera {  // RenderTargetCollection.hpp:45-54 (camData), the fields the loop reads
    float viewOnly[16];
    float zNear, zFar;
    float fovDegrees;
    uint32_t targetWidth, targetHeight;
};
struct RefCsmLight {  // DirectionalLight shadow settings + the upload record's direction
    float direction[3];
    uint32_t castsShadows;
    uint32_t numCascades;
    float shadowCascades[4];
    float maxShadowDistanceFactor;
};

namespace {
// include/RavEngine/mathtypes.hpp:141-149: reverse Z = near and far swapped
glm::mat4 reverse_perspective(float fovy, float aspect, float zn, float zf) { return glm::perspectiveRH_ZO(fovy, aspect, zf, zn); }
glm::mat4 reverse_ortho(float l, float r, float b, float t, float zn, float zf) { return glm::orthoRH_ZO(l, r, b, t, zf, zn); }
}  // namespace

extern "C" __attribute__((visibility("default"))) void ref_csm(const RefCsmCamera* cams, uint32_t nCams, const RefCsmLight* lights,
                                                               uint32_t nLights, float* viewProj, float* view, float* proj, float* dist,
                                                               float* tanHalfFovy) {
    for (uint32_t c = 0; c < nCams; ++c) {
        const RefCsmCamera& cam = cams[c];
        glm::mat4 camView;
        std::memcpy(&camView, cam.viewOnly, 64);
        // Common3D.hpp:117 deg_to_rad(float) = glm::radians; the argument fov * 1.5 is a double narrowed to float
        const float fovy = glm::radians(float(cam.fovDegrees * 1.5));
        if (tanHalfFovy) tanHalfFovy[c] = std::tan(fovy / 2.0f);  // what perspectiveRH_ZO computes inside
        const float aspect = float(cam.targetWidth / cam.targetHeight);
        for (uint32_t l = 0; l < nLights; ++l) {
            const RefCsmLight& lt = lights[l];
            if (!lt.castsShadows) continue;
            const uint32_t slot = c + l;  // passIndex + i
            const uint8_t clamped = std::min<uint8_t>(uint8_t(lt.numCascades), 4);
            for (uint32_t k = 0; k < lt.numCascades; ++k) {
                float zn = cam.zNear, zf = cam.zFar;
                if (k > 0) zn = glm::mix(cam.zNear, cam.zFar, lt.shadowCascades[k - 1]);
                if (int(k) < int(clamped) - 1) zf = glm::mix(cam.zNear, cam.zFar, lt.shadowCascades[k]);
                zf *= lt.maxShadowDistanceFactor;
                const glm::mat4 P = reverse_perspective(fovy, aspect, zn, zf);
                const glm::mat4 inv = glm::inverse(P * camView);
                std::array<glm::vec4, 8> corner;
                unsigned n = 0;
                for (unsigned x = 0; x < 2; ++x)
                    for (unsigned y = 0; y < 2; ++y)
                        for (unsigned z = 0; z < 2; ++z) {
                            const glm::vec4 pt = inv * glm::vec4(2.0f * x - 1.0f, 2.0f * y - 1.0f, z, 1.0f);
                            corner[n++] = pt / pt.w;
                        }
                glm::vec3 mid(0, 0, 0);
                for (const glm::vec4& v : corner) mid += glm::vec3(v);
                mid /= corner.size();
                const glm::vec3 dir(lt.direction[0], lt.direction[1], lt.direction[2]);
                const glm::mat4 LV = glm::lookAt(mid + dir, mid, glm::vec3(0.0f, 1.0f, 0.0f));
                float lo[3], hi[3];
                for (int a = 0; a < 3; ++a) {
                    lo[a] = std::numeric_limits<float>::max();
                    hi[a] = std::numeric_limits<float>::lowest();
                }
                for (const glm::vec4& v : corner) {
                    const glm::vec4 q = LV * v;
                    for (int a = 0; a < 3; ++a) {
                        lo[a] = std::min(lo[a], q[a]);
                        hi[a] = std::max(hi[a], q[a]);
                    }
                }
                constexpr float zMult = 10.0f;
                if (lo[2] < 0) lo[2] *= zMult; else lo[2] /= zMult;
                if (hi[2] < 0) hi[2] /= zMult; else hi[2] *= zMult;
                const glm::mat4 LP = reverse_ortho(lo[0], hi[0], lo[1], hi[1], lo[2], hi[2]);
                const glm::mat4 LVP = LP * LV;
                const size_t at = (size_t(slot) * 4 + k) * 16;
                std::memcpy(viewProj + at, &LVP, 64);
                std::memcpy(view + at, &LV, 64);
                std::memcpy(proj + at, &LP, 64);
                dist[l * 4 + k] = zf;
            }
        }
    }
}
