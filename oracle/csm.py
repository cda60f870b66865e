"""CPU oracle of the cascaded-shadow-map fit — TEST INFRASTRUCTURE ONLY.

A numpy-float32 restatement of src/RenderEngine_Draw.cpp:149-293 with glm 0.9.9.8's arithmetic written out
operation by operation (every intermediate is an np.float32, nothing is fused): perspectiveRH_ZO / orthoRH_ZO
(glm/ext/matrix_clip_space.inl) with near and far swapped as include/RavEngine/mathtypes.hpp:141-149 does,
compute_inverse<4,4> (glm/detail/func_matrix.inl), lookAtRH (glm/ext/matrix_transform.inl), mat4*mat4 and
mat4*vec4 (glm/detail/type_mat4x4.inl). PARITY PINNED: tests/test_csm_oracle.py checks it bit for bit against
tests/golden/csm.npz, which oracle/_ref/libref_csm.so (the same loop through the reference's own glm headers,
oracle/ref_csm.cpp) produced. Small cases only: pure-Python loops.
"""
from __future__ import annotations

import ctypes
import ctypes.util

import numpy as np

_libm = ctypes.CDLL(ctypes.util.find_library("m") or "libm.so.6")
_libm.tanf.restype = ctypes.c_float
_libm.tanf.argtypes = [ctypes.c_float]

F = np.float32
ONE, ZERO = F(1.0), F(0.0)


def _col_scale(col, s):
    return [c * s for c in col]


def _col_add(a, b):
    return [x + y for x, y in zip(a, b)]


def mat_vec(m, v):
    """glm mat4 * vec4: (m0*v0 + m1*v1) + (m2*v2 + m3*v3). m is a list of 4 columns."""
    return _col_add(_col_add(_col_scale(m[0], v[0]), _col_scale(m[1], v[1])), _col_add(_col_scale(m[2], v[2]), _col_scale(m[3], v[3])))


def mat_mat(a, b):
    """glm mat4 * mat4: column j = ((a0*b[j][0] + a1*b[j][1]) + a2*b[j][2]) + a3*b[j][3]."""
    return [_col_add(_col_add(_col_add(_col_scale(a[0], b[j][0]), _col_scale(a[1], b[j][1])), _col_scale(a[2], b[j][2])), _col_scale(a[3], b[j][3]))
            for j in range(4)]


def inverse(m):
    """glm::inverse of a mat4 (cofactor expansion, func_matrix.inl compute_inverse<4,4>)."""
    c00 = m[2][2] * m[3][3] - m[3][2] * m[2][3]
    c02 = m[1][2] * m[3][3] - m[3][2] * m[1][3]
    c03 = m[1][2] * m[2][3] - m[2][2] * m[1][3]
    c04 = m[2][1] * m[3][3] - m[3][1] * m[2][3]
    c06 = m[1][1] * m[3][3] - m[3][1] * m[1][3]
    c07 = m[1][1] * m[2][3] - m[2][1] * m[1][3]
    c08 = m[2][1] * m[3][2] - m[3][1] * m[2][2]
    c10 = m[1][1] * m[3][2] - m[3][1] * m[1][2]
    c11 = m[1][1] * m[2][2] - m[2][1] * m[1][2]
    c12 = m[2][0] * m[3][3] - m[3][0] * m[2][3]
    c14 = m[1][0] * m[3][3] - m[3][0] * m[1][3]
    c15 = m[1][0] * m[2][3] - m[2][0] * m[1][3]
    c16 = m[2][0] * m[3][2] - m[3][0] * m[2][2]
    c18 = m[1][0] * m[3][2] - m[3][0] * m[1][2]
    c19 = m[1][0] * m[2][2] - m[2][0] * m[1][2]
    c20 = m[2][0] * m[3][1] - m[3][0] * m[2][1]
    c22 = m[1][0] * m[3][1] - m[3][0] * m[1][1]
    c23 = m[1][0] * m[2][1] - m[2][0] * m[1][1]
    f0, f1, f2 = [c00, c00, c02, c03], [c04, c04, c06, c07], [c08, c08, c10, c11]
    f3, f4, f5 = [c12, c12, c14, c15], [c16, c16, c18, c19], [c20, c20, c22, c23]
    v0 = [m[1][0], m[0][0], m[0][0], m[0][0]]
    v1 = [m[1][1], m[0][1], m[0][1], m[0][1]]
    v2 = [m[1][2], m[0][2], m[0][2], m[0][2]]
    v3 = [m[1][3], m[0][3], m[0][3], m[0][3]]

    def comb(a, fa, b, fb, c, fc):  # a*fa - b*fb + c*fc, left to right
        return [(a[i] * fa[i] - b[i] * fb[i]) + c[i] * fc[i] for i in range(4)]

    i0, i1 = comb(v1, f0, v2, f1, v3, f2), comb(v0, f0, v2, f3, v3, f4)
    i2, i3 = comb(v0, f1, v1, f3, v3, f5), comb(v0, f2, v1, f4, v2, f5)
    sa, sb = [F(1), F(-1), F(1), F(-1)], [F(-1), F(1), F(-1), F(1)]
    inv = [[i0[i] * sa[i] for i in range(4)], [i1[i] * sb[i] for i in range(4)], [i2[i] * sa[i] for i in range(4)],
           [i3[i] * sb[i] for i in range(4)]]
    d = [m[0][i] * inv[i][0] for i in range(4)]
    det = (d[0] + d[1]) + (d[2] + d[3])
    ood = ONE / det
    return [_col_scale(col, ood) for col in inv]


def _dot3(a, b):
    return (a[0] * b[0] + a[1] * b[1]) + a[2] * b[2]


def _cross(x, y):
    return [x[1] * y[2] - y[1] * x[2], x[2] * y[0] - y[2] * x[0], x[0] * y[1] - y[0] * x[1]]


def _normalize(v):
    s = ONE / np.sqrt(_dot3(v, v))  # glm::inversesqrt
    return [c * s for c in v]


def look_at(eye, center, up):
    """glm::lookAtRH."""
    f = _normalize([center[i] - eye[i] for i in range(3)])
    s = _normalize(_cross(f, up))
    u = _cross(s, f)
    return [[s[0], u[0], -f[0], ZERO], [s[1], u[1], -f[1], ZERO], [s[2], u[2], -f[2], ZERO],
            [-_dot3(s, eye), -_dot3(u, eye), _dot3(f, eye), ONE]]


def reverse_perspective(tan_half, aspect, z_near, z_far):
    """RMath::perspectiveProjection(fovy, aspect, zNear, zFar) = perspectiveRH_ZO(fovy, aspect, zFar, zNear)."""
    zn, zf = z_far, z_near
    m = [[ZERO] * 4 for _ in range(4)]
    m[0][0] = ONE / (aspect * tan_half)
    m[1][1] = ONE / tan_half
    m[2][2] = zf / (zn - zf)
    m[2][3] = -ONE
    m[3][2] = -(zf * zn) / (zf - zn)
    return m


def reverse_ortho(left, right, bottom, top, z_near, z_far):
    """RMath::orthoProjection(l, r, b, t, zNear, zFar) = orthoRH_ZO(l, r, b, t, zFar, zNear)."""
    zn, zf = z_far, z_near
    m = [[ZERO] * 4 for _ in range(4)]
    m[0][0] = F(2) / (right - left)
    m[1][1] = F(2) / (top - bottom)
    m[2][2] = -ONE / (zf - zn)
    m[3][0] = -(right + left) / (right - left)
    m[3][1] = -(top + bottom) / (top - bottom)
    m[3][2] = -zn / (zf - zn)
    m[3][3] = ONE
    return m


def camera_params(fov_degrees: float, width: int, height: int) -> tuple[np.float32, np.float32]:
    """(tanHalfFovy, aspect) as src/RenderEngine_Draw.cpp:220 + perspectiveRH_ZO derive them: fov * 1.5 in double,
    narrowed to float, glm::radians, tan(fovy / 2) in float; aspect = float(width / height), integer division."""
    fovy = F(F(float(F(fov_degrees)) * 1.5) * F(0.01745329251994329576923690768489))
    return F(_libm.tanf(float(fovy / F(2)))), F(width // height)  # std::tan(float) = libm tanf (numpy's differs by an ulp)


def csm_cascades(cams, lights):
    """cams: list of dicts(viewOnly (16,) column-major, zNear, zFar, tanHalfFovy, aspect); lights: list of
    dicts(direction (3,), castsShadows, numCascades, shadowCascades (4,), maxShadowDistanceFactor).
    Returns (viewProj, view, proj) of shape (nCams*nLights, 4, 16) and distances (nLights, 4); untouched entries 0."""
    n_slots = len(cams) * len(lights)
    out_vp = np.zeros((n_slots, 4, 16), np.float32)
    out_v = np.zeros_like(out_vp)
    out_p = np.zeros_like(out_vp)
    dist = np.zeros((len(lights), 4), np.float32)
    for ci, cam in enumerate(cams):
        view = [[F(cam["viewOnly"][4 * j + r]) for r in range(4)] for j in range(4)]
        z0, z1 = F(cam["zNear"]), F(cam["zFar"])
        tan_half, aspect = F(cam["tanHalfFovy"]), F(cam["aspect"])
        for li, lt in enumerate(lights):
            if not lt["castsShadows"]:
                continue
            slot = ci + li  # passIndex + i (:280); passIndex advances once per camera (:289)
            casc = [F(x) for x in lt["shadowCascades"]]
            n_casc = min(int(lt["numCascades"]), 4)
            for k in range(int(lt["numCascades"])):
                near, far = z0, z1
                if k > 0:  # glm::mix(x, y, a) = x * (1 - a) + y * a
                    near = z0 * (ONE - casc[k - 1]) + z1 * casc[k - 1]
                if k < n_casc - 1:
                    far = z0 * (ONE - casc[k]) + z1 * casc[k]
                far = far * F(lt["maxShadowDistanceFactor"])
                proj = reverse_perspective(tan_half, aspect, near, far)
                inv = inverse(mat_mat(proj, view))
                corners = []
                for x in range(2):
                    for y in range(2):
                        for z in range(2):
                            pt = mat_vec(inv, [F(2) * F(x) - ONE, F(2) * F(y) - ONE, F(z), ONE])
                            corners.append([c / pt[3] for c in pt])
                center = [ZERO, ZERO, ZERO]
                for v in corners:
                    center = [center[i] + v[i] for i in range(3)]
                center = [c / F(8) for c in center]
                d = [F(x) for x in lt["direction"]]
                lv = look_at([center[i] + d[i] for i in range(3)], center, [ZERO, ONE, ZERO])
                big = np.finfo(np.float32).max
                lo, hi = [big] * 3, [-big] * 3
                for v in corners:
                    q = mat_vec(lv, v)
                    for a in range(3):
                        lo[a] = min(lo[a], q[a])
                        hi[a] = max(hi[a], q[a])
                z_mult = F(10)
                lo[2] = lo[2] * z_mult if lo[2] < 0 else lo[2] / z_mult
                hi[2] = hi[2] / z_mult if hi[2] < 0 else hi[2] * z_mult
                lp = reverse_ortho(lo[0], hi[0], lo[1], hi[1], lo[2], hi[2])
                lvp = mat_mat(lp, lv)
                out_vp[slot, k] = np.array(lvp, np.float32).reshape(16)
                out_v[slot, k] = np.array(lv, np.float32).reshape(16)
                out_p[slot, k] = np.array(lp, np.float32).reshape(16)
                dist[li, k] = far
    return out_vp, out_v, out_p, dist
