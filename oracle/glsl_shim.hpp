// glsl_shim.hpp -- TEST INFRASTRUCTURE (part of oracle/; never linked into the product).
//
// A minimal GLSL-on-the-CPU vocabulary so that the reference's shaders
// (data/shader/nrc-train/nrc-train.comp, nrc-forward/nrc-forward.frag,
// nrc-step/nrc-step.comp, mrhe-step/mrhe-step.comp) can be restated function
// for function in C++.  Every operation is a plain IEEE binary32 operation
// (build with -ffp-contract=off); the evaluation order of the compound
// builtins (dot, length, normalize, mat4*vec4) is fixed HERE and is part of the
// parity definition: GLSL leaves it to the driver, so one order had to be
// chosen.  Transcendentals come from include/nrc_math.h (the normative
// definitions shared with the CUDA kernels).
#pragma once
#include <cstdint>
#include "../include/nrc_math.h"

namespace glsl {

typedef uint32_t uint;

struct vec2 {
    float x, y;
    vec2() : x(0), y(0) {}
    explicit vec2(float s) : x(s), y(s) {}
    vec2(float x_, float y_) : x(x_), y(y_) {}
};

struct vec3 {
    float x, y, z;
    vec3() : x(0), y(0), z(0) {}
    explicit vec3(float s) : x(s), y(s), z(s) {}
    vec3(float x_, float y_, float z_) : x(x_), y(y_), z(z_) {}
    vec2 xz() const { return vec2(x, z); }
    float operator[](int i) const { return i == 0 ? x : (i == 1 ? y : z); }
};

struct vec4 {
    float x, y, z, w;
    vec4() : x(0), y(0), z(0), w(0) {}
    explicit vec4(float s) : x(s), y(s), z(s), w(s) {}
    vec4(float x_, float y_, float z_, float w_) : x(x_), y(y_), z(z_), w(w_) {}
    vec4(const vec3& v, float w_) : x(v.x), y(v.y), z(v.z), w(w_) {}
    vec4(const vec2& v, float z_, float w_) : x(v.x), y(v.y), z(z_), w(w_) {}
    vec3 xyz() const { return vec3(x, y, z); }
};

struct uvec3 {
    uint x, y, z;
    uvec3() : x(0), y(0), z(0) {}
    uvec3(uint x_, uint y_, uint z_) : x(x_), y(y_), z(z_) {}
    // uvec3(vec3): float -> uint conversion; negative/NaN saturate to 0.
    explicit uvec3(const vec3& v) : x(nrc_f2u(v.x)), y(nrc_f2u(v.y)), z(nrc_f2u(v.z)) {}
};

inline vec3 to_vec3(const uvec3& u) { return vec3((float)u.x, (float)u.y, (float)u.z); }

// ---- vec2 ----
inline vec2 operator+(vec2 a, vec2 b) { return vec2(a.x + b.x, a.y + b.y); }
inline vec2 operator-(vec2 a, vec2 b) { return vec2(a.x - b.x, a.y - b.y); }
inline vec2 operator*(vec2 a, vec2 b) { return vec2(a.x * b.x, a.y * b.y); }
inline vec2 operator*(vec2 a, float s) { return vec2(a.x * s, a.y * s); }
inline vec2 operator*(float s, vec2 a) { return vec2(s * a.x, s * a.y); }
inline vec2 operator+(vec2 a, float s) { return vec2(a.x + s, a.y + s); }
inline vec2 operator-(vec2 a) { return vec2(-a.x, -a.y); }

// ---- vec3 ----
inline vec3 operator+(vec3 a, vec3 b) { return vec3(a.x + b.x, a.y + b.y, a.z + b.z); }
inline vec3 operator-(vec3 a, vec3 b) { return vec3(a.x - b.x, a.y - b.y, a.z - b.z); }
inline vec3 operator*(vec3 a, vec3 b) { return vec3(a.x * b.x, a.y * b.y, a.z * b.z); }
inline vec3 operator/(vec3 a, vec3 b) { return vec3(a.x / b.x, a.y / b.y, a.z / b.z); }
inline vec3 operator*(vec3 a, float s) { return vec3(a.x * s, a.y * s, a.z * s); }
inline vec3 operator*(float s, vec3 a) { return vec3(s * a.x, s * a.y, s * a.z); }
inline vec3 operator/(vec3 a, float s) { return vec3(a.x / s, a.y / s, a.z / s); }
inline vec3 operator-(vec3 a) { return vec3(-a.x, -a.y, -a.z); }
inline vec3& operator+=(vec3& a, vec3 b) { a = a + b; return a; }
inline vec3& operator*=(vec3& a, float s) { a = a * s; return a; }
inline vec3& operator/=(vec3& a, float s) { a = a / s; return a; }

// ---- vec4 ----
inline vec4 operator+(vec4 a, vec4 b) { return vec4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w); }
inline vec4 operator*(vec4 a, float s) { return vec4(a.x * s, a.y * s, a.z * s, a.w * s); }
inline vec4 operator/(vec4 a, float s) { return vec4(a.x / s, a.y / s, a.z / s, a.w / s); }
inline vec4& operator+=(vec4& a, vec4 b) { a = a + b; return a; }
inline vec4& operator/=(vec4& a, float s) { a = a / s; return a; }

// ---- builtins (evaluation order fixed here) ----
inline float dot(vec2 a, vec2 b) { return a.x * b.x + a.y * b.y; }
inline float dot(vec3 a, vec3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
inline float length(vec2 v) { return NRC_SQRT(dot(v, v)); }
inline float length(vec3 v) { return NRC_SQRT(dot(v, v)); }
inline float distance(vec3 a, vec3 b) { return length(a - b); }
// normalize(v) = v * (1 / length(v))
inline vec3 normalize(vec3 v) { float inv = 1.0f / length(v); return v * inv; }
inline vec3 abs(vec3 v) { return vec3(nrc_abs(v.x), nrc_abs(v.y), nrc_abs(v.z)); }
inline float abs(float v) { return nrc_abs(v); }
inline float max(float a, float b) { return nrc_max(a, b); }
inline float min(float a, float b) { return nrc_min(a, b); }
inline vec3 max(vec3 v, float s) { return vec3(nrc_max(v.x, s), nrc_max(v.y, s), nrc_max(v.z, s)); }
inline vec3 min(vec3 a, vec3 b) { return vec3(nrc_min(a.x, b.x), nrc_min(a.y, b.y), nrc_min(a.z, b.z)); }
inline vec3 floor(vec3 v) { return vec3(nrc_floor(v.x), nrc_floor(v.y), nrc_floor(v.z)); }
inline float floor(float v) { return nrc_floor(v); }
inline float fract(float v) { return nrc_fract(v); }
inline float sin(float v) { return nrc_sin(v); }
inline float cos(float v) { return nrc_cos(v); }
inline float acos(float v) { return nrc_acos(v); }
inline float asin(float v) { return nrc_asin(v); }
inline float atan(float y, float x) { return nrc_atan2(y, x); }
inline float exp(float v) { return nrc_exp(v); }
inline float sqrt(float v) { return NRC_SQRT(v); }
inline float sign(float v) { return nrc_sign(v); }
inline bool isnan(float v) { return v != v; }
inline bool isinf(float v) { return nrc_abs(v) == __builtin_inff(); }

// column-major 4x4, constructor takes 16 scalars column by column like GLSL.
struct mat4 {
    float c[4][4];  // c[col][row]
    mat4() { for (int i = 0; i < 4; i++) for (int j = 0; j < 4; j++) c[i][j] = 0.0f; }
    mat4(float a0, float a1, float a2, float a3, float b0, float b1, float b2, float b3,
         float c0, float c1, float c2, float c3, float d0, float d1, float d2, float d3)
    {
        c[0][0] = a0; c[0][1] = a1; c[0][2] = a2; c[0][3] = a3;
        c[1][0] = b0; c[1][1] = b1; c[1][2] = b2; c[1][3] = b3;
        c[2][0] = c0; c[2][1] = c1; c[2][2] = c2; c[2][3] = c3;
        c[3][0] = d0; c[3][1] = d1; c[3][2] = d2; c[3][3] = d3;
    }
};

// (M * v)[row] = ((M[0][row]*v.x + M[1][row]*v.y) + M[2][row]*v.z) + M[3][row]*v.w
inline vec4 operator*(const mat4& m, vec4 v)
{
    vec4 r;
    r.x = ((m.c[0][0] * v.x + m.c[1][0] * v.y) + m.c[2][0] * v.z) + m.c[3][0] * v.w;
    r.y = ((m.c[0][1] * v.x + m.c[1][1] * v.y) + m.c[2][1] * v.z) + m.c[3][1] * v.w;
    r.z = ((m.c[0][2] * v.x + m.c[1][2] * v.y) + m.c[2][2] * v.z) + m.c[3][2] * v.w;
    r.w = ((m.c[0][3] * v.x + m.c[1][3] * v.y) + m.c[2][3] * v.z) + m.c[3][3] * v.w;
    return r;
}

// ---- samplers (Vulkan semantics: texel centres at (i+0.5)/N, linear filter) ----

// sampler3D over VK_FORMAT_R8G8B8A8_UNORM, CLAMP_TO_BORDER, opaque black
// (src/main.cpp:124-128, src/Texture3D.cpp:25-40, 147-155).  Memory order
// index = 4*(i + W*j + W*H*k), x fastest.  Only .x is ever read by the shaders.
struct sampler3D {
    int w = 0, h = 0, d = 0;
    const uint8_t* rgba = nullptr;
    float texel_r(long i, long j, long k) const
    {
        if (i < 0 || j < 0 || k < 0 || i >= w || j >= h || k >= d) return 0.0f;  // border (0,0,0,1)
        return (float)rgba[4 * ((size_t)i + (size_t)w * (size_t)j + (size_t)w * (size_t)h * (size_t)k)] / 255.0f;
    }
};

// Returns .x of texture(tex, uvw).  Coordinates that are NaN or further than one
// texel outside the image see only border texels and return 0.
inline float texture_r(const sampler3D& t, vec3 uvw)
{
    float u = uvw.x * (float)t.w - 0.5f;
    float v = uvw.y * (float)t.h - 0.5f;
    float s = uvw.z * (float)t.d - 0.5f;
    if (!(u >= -1.0f && u < (float)t.w)) return 0.0f;
    if (!(v >= -1.0f && v < (float)t.h)) return 0.0f;
    if (!(s >= -1.0f && s < (float)t.d)) return 0.0f;
    float fu = nrc_floor(u), fv = nrc_floor(v), fs = nrc_floor(s);
    float a = u - fu, b = v - fv, c = s - fs;
    long i0 = (long)fu, j0 = (long)fv, k0 = (long)fs;
    float acc = 0.0f;
    for (int dk = 0; dk < 2; dk++)
        for (int dj = 0; dj < 2; dj++)
            for (int di = 0; di < 2; di++) {
                float wgt = (di ? a : 1.0f - a) * (dj ? b : 1.0f - b) * (dk ? c : 1.0f - c);
                acc += wgt * t.texel_r(i0 + di, j0 + dj, k0 + dk);
            }
    return acc;
}

// sampler2D over float texels with `ch` channels, CLAMP_TO_EDGE, linear
// (src/HdrEnvMap.cpp:126-148, 286-297).
struct sampler2D {
    int w = 0, h = 0, ch = 4;
    const float* data = nullptr;
    float texel(long i, long j, int c) const
    {
        if (i < 0) i = 0; if (j < 0) j = 0;
        if (i >= w) i = w - 1; if (j >= h) j = h - 1;
        return data[((size_t)j * (size_t)w + (size_t)i) * (size_t)ch + (size_t)c];
    }
};

inline vec4 texture(const sampler2D& t, vec2 uv)
{
    float u = uv.x * (float)t.w - 0.5f;
    float v = uv.y * (float)t.h - 0.5f;
    if (!(u == u) || !(v == v)) return vec4(0.0f);
    // clamp far-out coordinates first so the integer conversion is defined
    u = nrc_min(nrc_max(u, -1.0f), (float)t.w);
    v = nrc_min(nrc_max(v, -1.0f), (float)t.h);
    float fu = nrc_floor(u), fv = nrc_floor(v);
    float a = u - fu, b = v - fv;
    long i0 = (long)fu, j0 = (long)fv;
    float out[4] = {0, 0, 0, 0};
    for (int c = 0; c < t.ch && c < 4; c++) {
        float t00 = t.texel(i0, j0, c), t10 = t.texel(i0 + 1, j0, c);
        float t01 = t.texel(i0, j0 + 1, c), t11 = t.texel(i0 + 1, j0 + 1, c);
        out[c] = ((1.0f - a) * (1.0f - b) * t00 + a * (1.0f - b) * t10) + ((1.0f - a) * b * t01 + a * b * t11);
    }
    return vec4(out[0], out[1], out[2], out[3]);
}

struct sampler1D {
    int w = 0;
    const float* data = nullptr;
};

inline vec4 texture(const sampler1D& t, float uc)
{
    float u = uc * (float)t.w - 0.5f;
    if (!(u == u)) return vec4(0.0f);
    u = nrc_min(nrc_max(u, -1.0f), (float)t.w);
    float fu = nrc_floor(u);
    float a = u - fu;
    long i0 = (long)fu, i1 = i0 + 1;
    if (i0 < 0) i0 = 0; if (i1 < 0) i1 = 0;
    if (i0 >= t.w) i0 = t.w - 1; if (i1 >= t.w) i1 = t.w - 1;
    float r = (1.0f - a) * t.data[i0] + a * t.data[i1];
    return vec4(r, 0.0f, 0.0f, 1.0f);
}

}  // namespace glsl
