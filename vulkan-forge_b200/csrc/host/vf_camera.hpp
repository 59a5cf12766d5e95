// vf_camera.hpp -- host-side camera math of the drop-in module.
// Mirrors src/camera.rs (validators, exact error strings, gl_to_wgpu as written) on top of a
// restatement of glam 0.24.2's Mat4::look_at_rh / perspective_rh_gl (Cargo.lock:297-298; the
// crate is not vendored in the reference).  32 floats per frame: stays on the host, above the
// C ABI.  Compile with -ffp-contract=off (Rust never contracts).
#pragma once
#include <cmath>
#include <stdexcept>

namespace vfcam {

struct V3 { float x, y, z; };
struct M4 { float m[16]; };   // column-major: m[4*c + r]

inline V3 sub(V3 a, V3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
inline float dot(V3 a, V3 b) { return (a.x * b.x) + (a.y * b.y) + (a.z * b.z); }
inline V3 cross(V3 a, V3 b) {
    return {a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y};
}
inline V3 scale(V3 a, float s) { return {a.x * s, a.y * s, a.z * s}; }
inline float length(V3 a) { return std::sqrt(dot(a, a)); }
inline V3 normalize(V3 a) { return scale(a, 1.0f / length(a)); }
inline V3 normalize_or_zero(V3 a) {
    const float r = 1.0f / length(a);
    if (std::isfinite(r) && r > 0.0f) return scale(a, r);
    return {0.0f, 0.0f, 0.0f};
}
inline bool finite3(V3 a) { return std::isfinite(a.x) && std::isfinite(a.y) && std::isfinite(a.z); }

// src/camera.rs:24-30
constexpr const char* ERROR_FOVY = "fovy_deg must be finite and in (0, 180)";
constexpr const char* ERROR_NEAR = "znear must be finite and > 0";
constexpr const char* ERROR_FAR = "zfar must be finite and > znear";
constexpr const char* ERROR_ASPECT = "aspect must be finite and > 0";
constexpr const char* ERROR_VECFINITE = "eye/target/up components must be finite";
constexpr const char* ERROR_UPCOLINEAR = "up vector must not be colinear with view direction";
constexpr const char* ERROR_CLIP = "clip_space must be 'wgpu' or 'gl'";

// every validator raises RuntimeError, like PyRuntimeError in the reference (src/camera.rs:33-91)
inline void validate_vec3_finite(V3 v) { if (!finite3(v)) throw std::runtime_error(ERROR_VECFINITE); }
inline void validate_fovy(float f) {
    if (!std::isfinite(f) || f <= 0.0f || f >= 180.0f) throw std::runtime_error(ERROR_FOVY);
}
inline void validate_near(float n) { if (!std::isfinite(n) || n <= 0.0f) throw std::runtime_error(ERROR_NEAR); }
inline void validate_far(float f, float n) { if (!std::isfinite(f) || f <= n) throw std::runtime_error(ERROR_FAR); }
inline void validate_aspect(float a) { if (!std::isfinite(a) || a <= 0.0f) throw std::runtime_error(ERROR_ASPECT); }
inline void validate_up_not_colinear(V3 eye, V3 target, V3 up) {
    const V3 c = cross(normalize_or_zero(sub(target, eye)), normalize_or_zero(up));
    if (dot(c, c) < 1e-6f) throw std::runtime_error(ERROR_UPCOLINEAR);
}
// src/camera.rs:224-240 (order of the checks is part of the contract)
inline void validate_camera_params(V3 eye, V3 target, V3 up, float fovy_deg, float znear, float zfar) {
    validate_vec3_finite(eye);
    validate_vec3_finite(target);
    validate_vec3_finite(up);
    validate_up_not_colinear(eye, target, up);
    validate_fovy(fovy_deg);
    validate_near(znear);
    validate_far(zfar, znear);
}

// glam Mat4::look_at_rh(eye, center, up) = look_to_rh(eye, center - eye, up)
inline M4 look_at_rh(V3 eye, V3 center, V3 up) {
    const V3 f = normalize(sub(center, eye));
    const V3 s = normalize(cross(f, up));
    const V3 u = cross(s, f);
    M4 r;
    r.m[0] = s.x; r.m[1] = u.x; r.m[2] = -f.x; r.m[3] = 0.0f;
    r.m[4] = s.y; r.m[5] = u.y; r.m[6] = -f.y; r.m[7] = 0.0f;
    r.m[8] = s.z; r.m[9] = u.z; r.m[10] = -f.z; r.m[11] = 0.0f;
    r.m[12] = -dot(eye, s); r.m[13] = -dot(eye, u); r.m[14] = dot(eye, f); r.m[15] = 1.0f;
    return r;
}

// glam Mat4::perspective_rh_gl
inline M4 perspective_rh_gl(float fovy, float aspect, float zn, float zf) {
    const float inv_length = 1.0f / (zn - zf);
    const float f = 1.0f / std::tan(0.5f * fovy);
    M4 r{};
    r.m[0] = f / aspect;
    r.m[5] = f;
    r.m[10] = (zn + zf) * inv_length;
    r.m[11] = -1.0f;
    r.m[14] = (2.0f * zn * zf) * inv_length;
    return r;
}

// glam Mat4 * Mat4: each result column = a.x_axis*b.x + a.y_axis*b.y + a.z_axis*b.z + a.w_axis*b.w
inline M4 mul(const M4& a, const M4& b) {
    M4 r;
    for (int c = 0; c < 4; ++c)
        for (int k = 0; k < 4; ++k)
            r.m[4 * c + k] = ((a.m[k] * b.m[4 * c] + a.m[4 + k] * b.m[4 * c + 1]) + a.m[8 + k] * b.m[4 * c + 2]) +
                             a.m[12 + k] * b.m[4 * c + 3];
    return r;
}

// src/camera.rs:14-21: the literal is consumed column by column, so this is the transpose of
// the intended depth remap (z' = 0.5 z, w' = 0.5 z + w).  Reproduced as shipped.
inline M4 gl_to_wgpu() {
    return M4{{1.0f, 0.0f, 0.0f, 0.0f, 0.0f, 1.0f, 0.0f, 0.0f, 0.0f, 0.0f, 0.5f, 0.5f, 0.0f, 0.0f, 0.0f, 1.0f}};
}
// src/camera.rs:218-221
inline M4 perspective_wgpu(float fovy, float aspect, float zn, float zf) {
    return mul(gl_to_wgpu(), perspective_rh_gl(fovy, aspect, zn, zf));
}
// f32::to_radians
inline float to_radians(float deg) { return deg * (3.14159265358979323846f / 180.0f); }

}  // namespace vfcam
