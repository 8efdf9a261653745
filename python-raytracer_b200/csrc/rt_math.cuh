// rt_math.cuh - float3 helpers shared by every kernel of the render path.
//
// All arithmetic is fp32 (BASELINE.json north_star); operation ORDER follows the reference's
// PyGLM expressions (cited where used) so results differ from it only by fp32-vs-fp64 scalar
// rounding and FMA contraction.  RT_HD functions also compile with a plain host compiler
// (tests/hostsim builds the same core for CPU-side debugging; never used by the product).
#pragma once
#include <cmath>
#include <cstdint>
#include <vector_types.h>

#if defined(__CUDACC__)
#define RT_HD __host__ __device__ __forceinline__
#define RT_HD_NOINLINE __host__ __device__ __noinline__
#else
#define RT_HD inline
#define RT_HD_NOINLINE inline
#endif

#if defined(__CUDA_ARCH__)
#define RT_LDG(p) __ldg(p)
#else
#define RT_LDG(p) (*(p))
#endif

#define RT_INF __builtin_huge_valf()

RT_HD int as_int(float f) { union { float f; int i; } u; u.f = f; return u.i; }
RT_HD float as_float(int i) { union { float f; int i; } u; u.i = i; return u.f; }
RT_HD float3 f3(float x, float y, float z) { float3 r; r.x = x; r.y = y; r.z = z; return r; }
RT_HD float3 f3(const float4& v) { return f3(v.x, v.y, v.z); }
RT_HD float4 f4(float x, float y, float z, float w) { float4 r; r.x = x; r.y = y; r.z = z; r.w = w; return r; }
RT_HD float3 operator+(float3 a, float3 b) { return f3(a.x + b.x, a.y + b.y, a.z + b.z); }
RT_HD float3 operator-(float3 a, float3 b) { return f3(a.x - b.x, a.y - b.y, a.z - b.z); }
RT_HD float3 operator*(float3 a, float3 b) { return f3(a.x * b.x, a.y * b.y, a.z * b.z); }
RT_HD float3 operator*(float3 a, float s) { return f3(a.x * s, a.y * s, a.z * s); }
RT_HD float3 operator*(float s, float3 a) { return f3(a.x * s, a.y * s, a.z * s); }
RT_HD float3 operator-(float3 a) { return f3(-a.x, -a.y, -a.z); }
RT_HD float dot(float3 a, float3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }   // GLM: (x+y)+z
RT_HD float3 cross(float3 a, float3 b) {
    return f3(a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y);
}
// GLM normalize: v * (1 / sqrt(dot(v, v))) with IEEE sqrt and division (nvcc default -prec-div/-prec-sqrt)
RT_HD float3 normalize(float3 a) { float inv = 1.0f / sqrtf(dot(a, a)); return a * inv; }
RT_HD float clamp01(float x) { return fmaxf(0.0f, fminf(1.0f, x)); }
RT_HD float3 clamp01(float3 a) { return f3(clamp01(a.x), clamp01(a.y), clamp01(a.z)); }
RT_HD float clampf(float x, float lo, float hi) { return fmaxf(lo, fminf(hi, x)); }
RT_HD float3 clamp3(float3 x, float3 lo, float3 hi) { return f3(clampf(x.x, lo.x, hi.x), clampf(x.y, lo.y, hi.y), clampf(x.z, lo.z, hi.z)); }
RT_HD float comp(float3 a, int k) { return k == 0 ? a.x : (k == 1 ? a.y : a.z); }

// GLSL reflect / refract in GLM's evaluation order (scene.py:99,196)
RT_HD float3 reflect3(float3 i, float3 n) { float d = dot(n, i); return i - (n * d) * 2.0f; }
RT_HD float3 refract3(float3 i, float3 n, float eta, bool& tir) {
    float d = dot(n, i);
    float k = 1.0f - (eta * eta) * (1.0f - d * d);
    tir = !(k >= 0.0f);
    if (tir) return f3(0, 0, 0);
    return i * eta - n * (eta * d + sqrtf(k));
}

// column-major 4x4 (m[col*4+row]) times (p,1) / (d,0), GLM order (m0*x + m1*y) + (m2*z + m3*w).  The matrix is read as
// its four columns (128-bit loads; DevXform is 16-byte aligned), not as 12-16 scalar loads.
#if defined(__CUDA_ARCH__)
#define RT_COL(m, k) __ldg(reinterpret_cast<const float4*>(m) + (k))
#else
inline float4 rt_col_host(const float* m, int k) { float4 r; __builtin_memcpy(&r, m + 4 * k, sizeof(r)); return r; }   // caller buffers: any alignment
#define RT_COL(m, k) rt_col_host(m, k)
#endif
RT_HD float3 xform_point(const float* __restrict__ m, float3 p) {
    const float4 c0 = RT_COL(m, 0), c1 = RT_COL(m, 1), c2 = RT_COL(m, 2), c3 = RT_COL(m, 3);
    return f3((c0.x * p.x + c1.x * p.y) + (c2.x * p.z + c3.x),
              (c0.y * p.x + c1.y * p.y) + (c2.y * p.z + c3.y),
              (c0.z * p.x + c1.z * p.y) + (c2.z * p.z + c3.z));
}
RT_HD float3 xform_dir(const float* __restrict__ m, float3 d) {
    const float4 c0 = RT_COL(m, 0), c1 = RT_COL(m, 1), c2 = RT_COL(m, 2);
    return f3((c0.x * d.x + c1.x * d.y) + c2.x * d.z,
              (c0.y * d.x + c1.y * d.y) + c2.y * d.z,
              (c0.z * d.x + c1.z * d.y) + c2.z * d.z);
}
// hierarchy.py:76: normalize(transpose(Minv) * vec4(n,0)).xyz -- the vec4 is normalised INCLUDING
// its w = Minv[:,3] . n, so normals under translated nodes come out shorter than unit
RT_HD float3 xform_normal(const float* __restrict__ mi, float3 n) {
    const float4 c0 = RT_COL(mi, 0), c1 = RT_COL(mi, 1), c2 = RT_COL(mi, 2), c3 = RT_COL(mi, 3);
    float x = (c0.x * n.x + c0.y * n.y) + c0.z * n.z;
    float y = (c1.x * n.x + c1.y * n.y) + c1.z * n.z;
    float z = (c2.x * n.x + c2.y * n.y) + c2.z * n.z;
    float w = (c3.x * n.x + c3.y * n.y) + c3.z * n.z;
    float inv = 1.0f / sqrtf((x * x + y * y) + (z * z + w * w));
    return f3(x * inv, y * inv, z * inv);
}

// Philox4x32-10 (Salmon et al. 2011); counter = (idx_lo, idx_hi, 0, 0), key = seed
RT_HD void philox4x32_10(uint32_t& c0, uint32_t& c1, uint32_t& c2, uint32_t& c3, uint32_t k0, uint32_t k1) {
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
    for (int r = 0; r < 10; r++) {
        uint64_t p0 = (uint64_t)0xD2511F53u * c0;
        uint64_t p1 = (uint64_t)0xCD9E8D57u * c2;
        uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0;
        uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1;
        c1 = (uint32_t)p1; c3 = (uint32_t)p0; c0 = n0; c2 = n2;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
}
