// rt_ingest.cuh - mesh ingestion on the device: what Mesh.__init__ / _compute_normals do per mesh at load time
// (mesh.py:17-70): (v + translate) * scale, area-weighted vertex normals, extent and gate radius.
//
// Bit-exact with the host-side numpy path of python-raytracer_b200/geometry.py (which is itself pinned to the
// reference through the golden fixtures): every product and sum is an explicitly rounded fp32 operation
// (__fmul_rn / __fadd_rn / __fsub_rn are never contracted into FMAs), and each vertex accumulates its faces'
// weighted normals in FACE ORDER, like the reference's `normals[face[k]] += weighted_normal` loop (mesh.py:55-68):
// the (vertex, corner) pairs are stable-sorted by vertex with the LBVH builder's radix sort, then one thread per
// vertex sums its run front to back.  fp32 addition is not associative, so atomics would not reproduce the result.
#pragma once
#include "rt_bvh.cuh"

#if defined(__CUDACC__)
namespace ing {

__device__ __forceinline__ float3 sub_rn(float3 a, float3 b) { return f3(__fsub_rn(a.x, b.x), __fsub_rn(a.y, b.y), __fsub_rn(a.z, b.z)); }
__device__ __forceinline__ float3 add_rn(float3 a, float3 b) { return f3(__fadd_rn(a.x, b.x), __fadd_rn(a.y, b.y), __fadd_rn(a.z, b.z)); }
__device__ __forceinline__ float3 mul_rn(float3 a, float s) { return f3(__fmul_rn(a.x, s), __fmul_rn(a.y, s), __fmul_rn(a.z, s)); }
__device__ __forceinline__ float dot_rn(float3 a, float3 b) {     // GLM / _vecmath.dot_rows: (x + y) + z
    return __fadd_rn(__fadd_rn(__fmul_rn(a.x, b.x), __fmul_rn(a.y, b.y)), __fmul_rn(a.z, b.z));
}
__device__ __forceinline__ float3 cross_rn(float3 a, float3 b) {  // _vecmath.cross_rows
    return f3(__fsub_rn(__fmul_rn(a.y, b.z), __fmul_rn(b.y, a.z)),
              __fsub_rn(__fmul_rn(a.z, b.x), __fmul_rn(b.z, a.x)),
              __fsub_rn(__fmul_rn(a.x, b.y), __fmul_rn(b.x, a.y)));
}
__device__ __forceinline__ float3 normalize_rn(float3 a) {        // a * (1 / sqrt(dot(a, a))), IEEE sqrt and division
    const float inv = __fdiv_rn(1.0f, __fsqrt_rn(dot_rn(a, a)));
    return mul_rn(a, inv);
}
__device__ __forceinline__ float3 vert3(const float* __restrict__ v, int i) { return f3(v[3 * i], v[3 * i + 1], v[3 * i + 2]); }

// floats as order-preserving unsigned keys, for atomicMin / atomicMax
__device__ __forceinline__ unsigned f2ord(float f) { unsigned u = __float_as_uint(f); return (u & 0x80000000u) ? ~u : (u | 0x80000000u); }
__host__ __device__ inline float ord2f(unsigned u) {
    u = (u & 0x80000000u) ? (u & 0x7fffffffu) : ~u;
    union { unsigned u; float f; } c; c.u = u; return c.f;
}

// mesh.py:24: (vec3(v) + translate) * scale -- the OBJ's doubles are rounded to fp32 first, like glm.vec3(...) does.
// Also folds the extent (mesh.py:29-34) into six atomics per block.
__global__ void __launch_bounds__(256) k_transform(const double* __restrict__ raw, int n, float3 t, float scale, float* __restrict__ out,
                                                   unsigned* __restrict__ ext /* min xyz, max xyz as ordered keys */) {
    __shared__ unsigned red[6][8];
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    unsigned k[6] = {0xffffffffu, 0xffffffffu, 0xffffffffu, 0u, 0u, 0u};
    if (i < n) {
        const float3 v = mul_rn(add_rn(f3((float)raw[3 * i], (float)raw[3 * i + 1], (float)raw[3 * i + 2]), t), scale);
        out[3 * i] = v.x; out[3 * i + 1] = v.y; out[3 * i + 2] = v.z;
        k[0] = k[3] = f2ord(v.x); k[1] = k[4] = f2ord(v.y); k[2] = k[5] = f2ord(v.z);
    }
    for (int c = 0; c < 6; c++) {
        unsigned x = c < 3 ? __reduce_min_sync(0xffffffffu, k[c]) : __reduce_max_sync(0xffffffffu, k[c]);
        if ((threadIdx.x & 31) == 0) red[c][threadIdx.x >> 5] = x;
    }
    __syncthreads();
    if (threadIdx.x < 6) {
        const int c = threadIdx.x;
        unsigned x = red[c][0];
        for (int w = 1; w < 8; w++) x = c < 3 ? min(x, red[c][w]) : max(x, red[c][w]);
        if (c < 3) atomicMin(ext + c, x); else atomicMax(ext + c, x);
    }
}

// mesh.py:43: max over the vertices of length(v - center); sqrt is monotone, so the max is taken on the rounded lengths
__global__ void __launch_bounds__(256) k_max_dist(const float* __restrict__ verts, int n, float3 center, unsigned* __restrict__ out) {
    __shared__ unsigned red[8];
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    unsigned k = 0u;
    if (i < n) { const float3 d = sub_rn(vert3(verts, i), center); k = f2ord(__fsqrt_rn(dot_rn(d, d))); }
    k = __reduce_max_sync(0xffffffffu, k);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = k;
    __syncthreads();
    if (threadIdx.x == 0) { for (int w = 1; w < 8; w++) k = max(k, red[w]); atomicMax(out, k); }
}

// mesh.py:56-64: per face normalize(cross(e1, e2)) * (length(cross(e1, e2)) / 2); also emits the three
// (vertex, corner) sort pairs of the face
__global__ void __launch_bounds__(256) k_face_normals(const float* __restrict__ verts, const int* __restrict__ faces, int n_faces,
                                                      float4* __restrict__ weighted, uint64_t* __restrict__ keys, int* __restrict__ vals) {
    const int f = blockIdx.x * blockDim.x + threadIdx.x;
    if (f >= n_faces) return;
    const int i0 = faces[3 * f], i1 = faces[3 * f + 1], i2 = faces[3 * f + 2];
    const float3 v0 = vert3(verts, i0);
    const float3 c = cross_rn(sub_rn(vert3(verts, i1), v0), sub_rn(vert3(verts, i2), v0));
    const float3 nrm = normalize_rn(c);
    const float area = (float)((double)__fsqrt_rn(dot_rn(c, c)) / 2.0);     // length in fp32, halved as a Python float
    const float3 w = mul_rn(nrm, area);
    weighted[f] = f4(w.x, w.y, w.z, 0.0f);
    keys[3 * f] = (uint64_t)(unsigned)i0; keys[3 * f + 1] = (uint64_t)(unsigned)i1; keys[3 * f + 2] = (uint64_t)(unsigned)i2;
    vals[3 * f] = 3 * f; vals[3 * f + 1] = 3 * f + 1; vals[3 * f + 2] = 3 * f + 2;
}

// mesh.py:66-70: one thread per vertex walks its run of the sorted pairs (corner order = face order) and normalises
__global__ void __launch_bounds__(256) k_vertex_normals(const uint64_t* __restrict__ keys, const int* __restrict__ vals, int n_pairs,
                                                        const float4* __restrict__ weighted, int n_verts, float* __restrict__ out) {
    const int v = blockIdx.x * blockDim.x + threadIdx.x;
    if (v >= n_verts) return;
    int lo = 0, hi = n_pairs;                 // lower bound of v in the sorted keys
    while (lo < hi) { const int mid = (lo + hi) >> 1; if (keys[mid] < (uint64_t)v) lo = mid + 1; else hi = mid; }
    float3 acc = f3(0.0f, 0.0f, 0.0f);
    for (int i = lo; i < n_pairs && keys[i] == (uint64_t)v; i++) acc = add_rn(acc, f3(weighted[vals[i] / 3]));
    const float3 nrm = normalize_rn(acc);     // a vertex no face uses: 0 * inf = NaN, as in the reference
    out[3 * v] = nrm.x; out[3 * v + 1] = nrm.y; out[3 * v + 2] = nrm.z;
}

}  // namespace ing
#endif
