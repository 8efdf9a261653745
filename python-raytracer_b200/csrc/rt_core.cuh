// rt_core.cuh - per-ray device functions of the render path: primitive intersection, LBVH
// traversal, CSG/transform nodes, closest hit, any-hit shadow test, Blinn-Phong shading.
//
// Written from scratch for the GPU; each function cites the reference lines whose RESULTS it
// must reproduce (file:line under the reference's provided/).  fp32 throughout.
#pragma once
#include "rt_types.cuh"

// Kernel variants (the HAS_NODES template argument of everything below; non-zero = the scene has CSG / transform nodes):
//   0  lean: no nodes, everything inlined          1  nodes, and NO mesh anywhere in the scene          2  nodes and meshes
// Variant 1 exists because the general kernels are instruction-fetch bound (ncu: 80 % of the GPC fetch bandwidth for k_hit<1,1>):
// without the LBVH walks, triangle tests and smooth-normal code they shrink from 116 / 107 KB to 64 / 75 KB, and the bundled CSG
// scenes (no mesh) render 24 % faster.
#define RT_V_LEAN 0
#define RT_V_NODES 1
#define RT_V_NODES_MESH 2
#define RT_EPS 1e-4f                 // geometry/__init__.py:12
#define RT_SHADOW_EPS 1e-4f          // geometry/__init__.py:39
#define RT_SHADOW_EPS_SPHERE 1e-3f   // simple_geometry.py:13

#if defined(__CUDA_ARCH__)
#define RT_RSQRT(x) rsqrtf(x)
#else
#define RT_RSQRT(x) (1.0f / sqrtf(x))
#endif

RT_HD float4 ld4(const float4* p) { return RT_LDG(p); }

struct TraceCtx {
    const SceneView& S;
    float time;          // scene.current_time of this ray (scene.py:68), carried per ray
    Tally& tally;
    // BVH traversal stack of this thread, RT_STACK_DEPTH entries in shared memory.  On the device it is held as ONE 32-bit
    // shared-window address (entry k at saddr + k * RT_STACK_PITCH, accessed with ld/st.shared): through a generic pointer the
    // compiler rebuilt the window base (S2R CgaCtaId, MOV, LEA) and the index product at every push and pop.
#if defined(__CUDA_ARCH__)
    unsigned saddr;
#else
    int* stack;          // entry k at stack[k * stack_stride]
    int stack_stride;
#endif
    int sbase;           // entries below sbase belong to the object-BVH walk in progress (top_next)
    int* spill;          // entries RT_STACK_DEPTH.. of the stack: RT_STACK_SPILL ints of thread-local memory owned by the kernel,
                         // touched only by unusually deep walks (kept OUT of this struct so its scalars stay in registers)
    // General kernels: S above is the kernel's by-value parameter (constant bank) for everything that is inlined into the kernel;
    // the out-of-line node functions get a context whose S is the one copy in global memory (global()), because a reference to
    // the parameter that reaches a real call makes the compiler copy the whole parameter into every thread's stack frame.  So: a
    // context bound to the parameter must never be passed to an RT_HD_NOINLINE function - pass cx.global().  Sg == nullptr (lean
    // kernels, host twin): there is no such function on the path / no such distinction.
    const SceneView* Sg;
#if defined(__CUDA_ARCH__)
    __device__ __forceinline__ TraceCtx(const SceneView& S_, float time_, Tally& tally_, int* stack_, int /*stride: RT_BLOCK*/, int* spill_, const SceneView* Sg_ = nullptr)
        : S(S_), time(time_), tally(tally_), sbase(0), spill(spill_), Sg(Sg_) {
        // the volatile move keeps the address in a register: left to itself the compiler re-derives it (two S2R, LEA, ...) at every use
        asm volatile("mov.u32 %0, %1;" : "=r"(saddr) : "r"(stack_ ? (unsigned)__cvta_generic_to_shared(stack_) : 0u));
    }
    __device__ __forceinline__ TraceCtx(const SceneView& S_, const TraceCtx& o) : S(S_), time(o.time), tally(o.tally), saddr(o.saddr), sbase(o.sbase), spill(o.spill), Sg(o.Sg) {}
    __device__ __forceinline__ TraceCtx global() const { return TraceCtx(*Sg, *this); }      // (only the general kernels call it: Sg is set)
#else
    TraceCtx(const SceneView& S_, float time_, Tally& tally_, int* stack_, int stride_, int* spill_, const SceneView* Sg_ = nullptr)
        : S(S_), time(time_), tally(tally_), stack(stack_), stack_stride(stride_), sbase(0), spill(spill_), Sg(Sg_) {}
    TraceCtx global() const { return *this; }
#endif
};
#define RT_STACK_PITCH (4 * RT_BLOCK)    // bytes between consecutive entries of one thread's stack (entries of a warp's lanes are adjacent words)
#if defined(__CUDA_ARCH__)
__device__ __forceinline__ void stk_st(const TraceCtx& cx, int k, int v) { asm volatile("st.shared.b32 [%0], %1;" :: "r"(cx.saddr + (unsigned)k * RT_STACK_PITCH), "r"(v)); }
__device__ __forceinline__ int stk_ld(const TraceCtx& cx, int k) { int v; asm volatile("ld.shared.b32 %0, [%1];" : "=r"(v) : "r"(cx.saddr + (unsigned)k * RT_STACK_PITCH)); return v; }
#else
inline void stk_st(const TraceCtx& cx, int k, int v) { cx.stack[k * cx.stack_stride] = v; }
inline int stk_ld(const TraceCtx& cx, int k) { return cx.stack[k * cx.stack_stride]; }
#endif
RT_HD void stk_push(TraceCtx& cx, int& sp, int v) {
    if (sp < RT_STACK_DEPTH) stk_st(cx, sp, v);
    else if (sp < RT_STACK_DEPTH + RT_STACK_SPILL) cx.spill[sp - RT_STACK_DEPTH] = v;
    else { cx.tally.err |= 1u << RT_ERR_LIMIT; return; }                 // deeper than both: entry dropped, render reports RT_ERR_LIMIT
    sp++;
}
RT_HD int stk_pop(TraceCtx& cx, int& sp) {
    --sp;
    return sp < RT_STACK_DEPTH ? stk_ld(cx, sp) : cx.spill[sp - RT_STACK_DEPTH];
}

// `center + speed * current_time` (simple_geometry.py:21-24,106-109,189-194)
RT_HD float3 moved(const DevObject* o, float3 base, float time) { return base + f3(ld4(&o->speed)) * time; }
// the same through a trace context: scenes without any moving object skip the 16-byte speed load of every primitive test
// (base + 0 * time == base for every finite time, so the result is identical)
RT_HD float3 moved(const TraceCtx& cx, const DevObject* o, float3 base) { return cx.S.any_motion ? moved(o, base, cx.time) : base; }

// Moving meshes (EXTENSION, off unless SceneView::mesh_motion: the reference stores a mesh's `speed` and never uses it,
// mesh.py:72-153).  A mesh that translates by speed * time is met in ITS rest frame: the ray origin moves by -speed * time,
// direction and hit times stay, so the one LBVH built over the rest pose serves every motion time (no swept node boxes,
// no refit); positions go back with +shift, normals are unchanged.
RT_HD float3 mesh_shift(const TraceCtx& cx, int oi) {
    return cx.S.mesh_motion ? f3(ld4(&cx.S.objects[oi].speed)) * cx.time : f3(0.0f, 0.0f, 0.0f);
}

// ------------------------------------------------------------------------------------------------
// Sphere (simple_geometry.py:20-80).  A = d.d is ray-constant and passed in.
RT_HD bool sphere_roots(float3 center, float radius, const Ray& r, float A, float& t1, float& t2) {
    float3 oc = r.o - center;
    float B = 2.0f * dot(r.d, oc);
    float Cc = dot(oc, oc) - radius * radius;
    float disc = B * B - 4.0f * A * Cc;
    if (disc < 0.0f) return false;
    float sq = sqrtf(disc);
    float inv2a = 1.0f / (2.0f * A);
    t1 = (-B - sq) * inv2a;
    t2 = (-B + sq) * inv2a;
    return true;
}

// Plane (simple_geometry.py:105-131): |d.n| > eps, t = (point - o).n / d.n (caller applies the t window)
RT_HD bool plane_t(float3 point, float3 n, const Ray& r, float& t) {
    float denom = dot(r.d, n);
    if (!(fabsf(denom) > RT_EPS)) return false;
    t = dot(point - r.o, n) / denom;
    return true;
}

// Checkerboard material choice (simple_geometry.py:133-148)
RT_HD int plane_material(const DevObject* o, float3 p, float time) {
    if (o->n_mats == 1) return o->mat0;
    if (o->n_mats == 0) return -2;
    float3 position = moved(o, f3(ld4(&o->a)), time);
    float3 n = f3(ld4(&o->b));
    float3 pp = p - n * dot(p - position, n);
    float x = dot(pp - position, f3(ld4(&o->ax_w)));
    float z = dot(pp - position, f3(ld4(&o->ax_h)));
    double s = floor((double)position.x - (double)x) + floor((double)position.z - (double)z);
    double parity = s - 2.0 * floor(s * 0.5);       // Python (dx + dz) % 2
    return parity == 0.0 ? o->mat0 : o->mat1;
}

// Box slab test with labelled entry interval (simple_geometry.py:196-226, helperclasses.py:62-66).
// Returns false on the reference's early `return []` (zero direction component with the origin outside).
RT_HD bool box_slabs(float3 mn, float3 mx, const Ray& r, float& entry, float& exit_, int& label) {
    float st[3], en[3];
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
    for (int k = 0; k < 3; k++) {
        float o = comp(r.o, k), d = comp(r.d, k), lo = comp(mn, k), hi = comp(mx, k);
        if (d == 0.0f) {
            if (!(lo < o && o < hi)) return false;
            st[k] = -RT_INF; en[k] = RT_INF;
        } else {
            float t1 = (lo - o) / d, t2 = (hi - o) / d;
            st[k] = fminf(t1, t2); en[k] = fmaxf(t1, t2);
        }
    }
    entry = st[0]; label = 0;                           // max(..., key=start): first maximal wins
    if (st[1] > entry) { entry = st[1]; label = 1; }
    if (st[2] > entry) { entry = st[2]; label = 2; }
    exit_ = en[0];                                      // min(..., key=end)
    if (en[1] < exit_) exit_ = en[1];
    if (en[2] < exit_) exit_ = en[2];
    return true;
}
RT_HD float3 box_normal(const Ray& r, int label) {     // simple_geometry.py:231-242
    float d = comp(r.d, label);
    float s = d < 0.0f ? 1.0f : (d > 0.0f ? -1.0f : 0.0f);
    return f3(label == 0 ? s : 0.0f, label == 1 ? s : 0.0f, label == 2 ? s : 0.0f);
}

// ------------------------------------------------------------------------------------------------
// How the mesh functions below read a mesh header (DevMesh).  MeshG: through a pointer into the header array in global memory
// (meshes under CSG / transform nodes, the general kernels).  MeshC: field by field out of SceneConst in the kernel parameter bank
// (the k-th top-level mesh, lean kernels): every access is a constant-bank operand or an LDC, nothing is held in registers across
// the walk and nothing goes through the L1 data pipe.  (A POINTER to the parameter copy would make the compiler move the whole
// SceneView to local memory; hence the accessors.)
struct MeshG {
    const DevMesh* m;
    RT_HD const float4* tri() const { return m->tri; }
    RT_HD const BvhNode4* nodes() const { return m->nodes; }
    RT_HD const int4* face_vidx() const { return m->face_vidx; }
    RT_HD const float4* vnormals() const { return m->vnormals; }
    RT_HD int n_tris() const { return m->n_tris; }
    RT_HD int flat_shaded() const { return m->flat_shaded; }
    RT_HD int gate() const { return m->gate; }
    RT_HD float gate_radius() const { return m->gate_radius; }
    RT_HD float4 gate_min() const { return ld4(&m->gate_min); }
    RT_HD float4 gate_max() const { return ld4(&m->gate_max); }
    RT_HD float4 gate_center() const { return ld4(&m->gate_center); }
};
struct MeshC {
    const SceneView& S; int k;
    RT_HD const float4* tri() const { return S.c.sep_mesh[k].tri; }
    RT_HD const BvhNode4* nodes() const { return S.c.sep_mesh[k].nodes; }
    RT_HD const int4* face_vidx() const { return S.c.sep_mesh[k].face_vidx; }
    RT_HD const float4* vnormals() const { return S.c.sep_mesh[k].vnormals; }
    RT_HD int n_tris() const { return S.c.sep_mesh[k].n_tris; }
    RT_HD int flat_shaded() const { return S.c.sep_mesh[k].flat_shaded; }
    RT_HD int gate() const { return S.c.sep_mesh[k].gate; }
    RT_HD float gate_radius() const { return S.c.sep_mesh[k].gate_radius; }
    RT_HD float4 gate_min() const { return S.c.sep_mesh[k].gate_min; }
    RT_HD float4 gate_max() const { return S.c.sep_mesh[k].gate_max; }
    RT_HD float4 gate_center() const { return S.c.sep_mesh[k].gate_center; }
};

// ------------------------------------------------------------------------------------------------
// Mesh gate (bounding_volumes.py:18-37,49-83; SURVEY Q11: a ray starting inside the AABB never hits the mesh)
template <class M>
RT_HD bool mesh_gate(const M& m, const Ray& r) {
    if (m.gate() == RT_GATE_SPHERE) {
        float t1, t2;
        if (!sphere_roots(f3(m.gate_center()), m.gate_radius(), r, dot(r.d, r.d), t1, t2)) return false;
        return t1 > 0.0f || t2 > 0.0f;
    }
    float en, ex; int label;
    if (!box_slabs(f3(m.gate_min()), f3(m.gate_max()), r, en, ex, label)) return false;
    return !(en > ex || en < 0.0f);
}

// Triangle tests.  Closest-hit form (mesh.py:82-102): plane hit with the NORMALISED face normal,
// |d.n| < 1e-4 rejected, t < 0 rejected, three inclusive edge tests.
RT_HD bool tri_hit(float3 v0, float3 v1, float3 v2, const Ray& r, float& t) {
    float3 c = cross(v1 - v0, v2 - v0);
    float3 n = c * RT_RSQRT(dot(c, c));
    float denom = dot(r.d, n);
    if (fabsf(denom) < RT_EPS) return false;
    t = dot(v0 - r.o, n) / denom;
    if (t < 0.0f) return false;
    float3 p = r.o + r.d * t;
    float b0 = dot(cross(v1 - v0, p - v0), n);
    float b1 = dot(cross(v2 - v1, p - v1), n);
    float b2 = dot(cross(v0 - v2, p - v2), n);
    return b0 >= 0.0f && b1 >= 0.0f && b2 >= 0.0f;
}
// Shadow form (mesh.py:125-151): UN-normalised normal with the same 1e-4 cut-off, t < shadow_epsilon skipped,
// no upper bound on t.
RT_HD bool tri_shadow(float3 v0, float3 v1, float3 v2, const Ray& r) {
    float3 n = cross(v1 - v0, v2 - v0);
    float denom = dot(r.d, n);
    if (fabsf(denom) < RT_EPS) return false;
    float t = dot(v0 - r.o, n) / denom;
    if (t < RT_SHADOW_EPS) return false;
    float3 p = r.o + r.d * t;
    return dot(cross(v1 - v0, p - v0), n) >= 0.0f && dot(cross(v2 - v1, p - v1), n) >= 0.0f &&
           dot(cross(v0 - v2, p - v2), n) >= 0.0f;
}

// ------------------------------------------------------------------------------------------------
// 4-wide BVH walk shared by the per-mesh LBVH and the object BVH.
// Slab tests in FMA form t = plane * (1/d) + (-o/d).  A direction component smaller than 2^-60 gets a reciprocal
// of +-2^60 instead of inf, so the two products stay finite (inf - inf would poison the interval).
// nx/ny/nz: byte offset inside a BvhNode4 of the plane set the ray ENTERS through on each axis (lo planes for a
// non-negative direction component, hi planes otherwise); the exit planes are at (nx ^ 48), (80 - ny), (112 - nz).
// Picking near/far by address removes the per-box min/max of the classic slab test.
struct RayInv { float3 inv, oinv; int nx, ny, nz; };
// the reciprocal only scales conservative (padded) box tests: the 1-ulp MUFU.RCP is enough on the device
#if defined(__CUDA_ARCH__)
// one MUFU.RCP: the callers keep |x| >= 2^-60, so none of the range handling of __fdividef / IEEE division is needed
__device__ __forceinline__ float rt_rcp_approx(float x) { float r; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
#define RT_RCP(x) rt_rcp_approx(x)
#else
#define RT_RCP(x) (1.0f / (x))
#endif
RT_HD float safe_rcp(float d) { return fabsf(d) >= 8.673617e-19f ? RT_RCP(d) : copysignf(1.1529215e18f, d); }
RT_HD RayInv ray_inv(const Ray& r) {
    RayInv ri;
    ri.inv = f3(safe_rcp(r.d.x), safe_rcp(r.d.y), safe_rcp(r.d.z));
    ri.oinv = f3(-r.o.x * ri.inv.x, -r.o.y * ri.inv.y, -r.o.z * ri.inv.z);
    ri.nx = ri.inv.x < 0.0f ? 48 : 0; ri.ny = ri.inv.y < 0.0f ? 64 : 16; ri.nz = ri.inv.z < 0.0f ? 80 : 32;
    return ri;
}
// the same ray with another origin (moving meshes: the ray in the mesh's rest frame): only the origin terms change
RT_HD RayInv ray_inv_moved(const RayInv& ri, float3 o) {
    RayInv m = ri;
    m.oinv = f3(-o.x * ri.inv.x, -o.y * ri.inv.y, -o.z * ri.inv.z);
    return m;
}
// Mesh gate for rays whose reciprocal direction is at hand (top-level meshes): same decisions as mesh_gate up to the
// last ulp of the slab distances ((lo - o) * (1/d) instead of (lo - o) / d; six IEEE divisions saved per ray)
template <class M>
RT_HD bool mesh_gate_rcp(const M& m, const Ray& r, const RayInv& ri) {
    if (m.gate() == RT_GATE_SPHERE) return mesh_gate(m, r);
    const float3 mn = f3(m.gate_min()), mx = f3(m.gate_max());
    float entry = -RT_INF, exit_ = RT_INF;
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
    for (int k = 0; k < 3; k++) {
        const float o = comp(r.o, k), d = comp(r.d, k), lo = comp(mn, k), hi = comp(mx, k), iv = comp(ri.inv, k);
        if (d == 0.0f) { if (!(lo < o && o < hi)) return false; }
        else {
            const float t1 = (lo - o) * iv, t2 = (hi - o) * iv;
            entry = fmaxf(entry, fminf(t1, t2)); exit_ = fminf(exit_, fmaxf(t1, t2));
        }
    }
    return !(entry > exit_ || entry < 0.0f);
}
// does the segment [0, tmax] of the ray touch the scene's bounded part (SceneView::world_lo/hi)?  Only ever says "cannot": the box is padded
RT_HD bool touches_world(const SceneView& S, const RayInv& ri, float tmax) {
    const float ax = fmaf(S.world_lo.x, ri.inv.x, ri.oinv.x), bx = fmaf(S.world_hi.x, ri.inv.x, ri.oinv.x);
    const float ay = fmaf(S.world_lo.y, ri.inv.y, ri.oinv.y), by = fmaf(S.world_hi.y, ri.inv.y, ri.oinv.y);
    const float az = fmaf(S.world_lo.z, ri.inv.z, ri.oinv.z), bz = fmaf(S.world_hi.z, ri.inv.z, ri.oinv.z);
    const float near_ = fmaxf(fmaxf(fminf(ax, bx), fminf(ay, by)), fmaxf(fminf(az, bz), 0.0f));
    const float far_ = fminf(fminf(fmaxf(ax, bx), fmaxf(ay, by)), fminf(fmaxf(az, bz), tmax));
    return near_ <= far_;
}
RT_HD unsigned as_uint(float f) { return (unsigned)as_int(f); }
RT_HD int link_of(const int4& l, unsigned slot) { return (slot & 2u) ? ((slot & 1u) ? l.w : l.z) : ((slot & 1u) ? l.y : l.x); }
RT_HD void key_swap(unsigned& a, unsigned& b) { unsigned lo = a < b ? a : b, hi = a < b ? b : a; a = lo; b = hi; }

// Entry/exit distances of the four child boxes of a node from its entry planes (pnx, pny, pnz) and exit planes (pfx, pfy, pfz);
// an empty slot holds [+inf, -inf] boxes: its entry distance is +inf and its exit distance -inf for every direction sign.
// On the device the 24 plane * (1/d) + (-o/d) products are 12 FFMA2: Blackwell's packed FP32 FMA (fma.rn.f32x2, sm_100+) takes the
// two plane values of a child pair as one 64-bit register pair - exactly how LDG.128 delivers them - and broadcasts the scalar
// reciprocal and offset (SASS: FFMA2 R, R.F32x2.HI_LO, R.F32, R.F32).  Each half is an IEEE fused multiply-add, so the
// distances are bit-identical to the scalar form; the walkers are issue-bound, and this removes 12 issue slots per node.
struct Slab4 { float n0, n1, n2, n3; bool h0, h1, h2, h3; };
#if defined(__CUDA_ARCH__)
__device__ __forceinline__ float2 rt_fma2(float a0, float a1, float b, float c) {
    unsigned long long va, vb, vc, r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(va) : "f"(a0), "f"(a1));
    asm("mov.b64 %0, {%1, %1};" : "=l"(vb) : "f"(b));
    asm("mov.b64 %0, {%1, %1};" : "=l"(vc) : "f"(c));
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(va), "l"(vb), "l"(vc));
    float2 o;
    asm("mov.b64 {%0, %1}, %2;" : "=f"(o.x), "=f"(o.y) : "l"(r));
    return o;
}
#else
inline float2 rt_fma2(float a0, float a1, float b, float c) { float2 o; o.x = fmaf(a0, b, c); o.y = fmaf(a1, b, c); return o; }
#endif
// USE_TMAX = false: the far distances are not clipped at all (walks whose tmax is +inf by construction: mesh shadows ignore t_max)
template <bool USE_TMAX>
RT_HD Slab4 slab4(const float4& pnx, const float4& pny, const float4& pnz, const float4& pfx, const float4& pfy, const float4& pfz,
                  const float3 inv, const float3 oinv, const float tmin, const float t0, const float t1, const float t2, const float t3) {
    const float2 nx01 = rt_fma2(pnx.x, pnx.y, inv.x, oinv.x), nx23 = rt_fma2(pnx.z, pnx.w, inv.x, oinv.x);
    const float2 ny01 = rt_fma2(pny.x, pny.y, inv.y, oinv.y), ny23 = rt_fma2(pny.z, pny.w, inv.y, oinv.y);
    const float2 nz01 = rt_fma2(pnz.x, pnz.y, inv.z, oinv.z), nz23 = rt_fma2(pnz.z, pnz.w, inv.z, oinv.z);
    const float2 fx01 = rt_fma2(pfx.x, pfx.y, inv.x, oinv.x), fx23 = rt_fma2(pfx.z, pfx.w, inv.x, oinv.x);
    const float2 fy01 = rt_fma2(pfy.x, pfy.y, inv.y, oinv.y), fy23 = rt_fma2(pfy.z, pfy.w, inv.y, oinv.y);
    const float2 fz01 = rt_fma2(pfz.x, pfz.y, inv.z, oinv.z), fz23 = rt_fma2(pfz.z, pfz.w, inv.z, oinv.z);
    Slab4 r;
    r.n0 = fmaxf(fmaxf(nx01.x, ny01.x), fmaxf(nz01.x, tmin)); r.n1 = fmaxf(fmaxf(nx01.y, ny01.y), fmaxf(nz01.y, tmin));
    r.n2 = fmaxf(fmaxf(nx23.x, ny23.x), fmaxf(nz23.x, tmin)); r.n3 = fmaxf(fmaxf(nx23.y, ny23.y), fmaxf(nz23.y, tmin));
    float f0, f1, f2, f3_;
    if (USE_TMAX) {
        f0 = fminf(fminf(fx01.x, fy01.x), fminf(fz01.x, t0)); f1 = fminf(fminf(fx01.y, fy01.y), fminf(fz01.y, t1));
        f2 = fminf(fminf(fx23.x, fy23.x), fminf(fz23.x, t2)); f3_ = fminf(fminf(fx23.y, fy23.y), fminf(fz23.y, t3));
    } else {
        f0 = fminf(fminf(fx01.x, fy01.x), fz01.x); f1 = fminf(fminf(fx01.y, fy01.y), fz01.y);
        f2 = fminf(fminf(fx23.x, fy23.x), fz23.x); f3_ = fminf(fminf(fx23.y, fy23.y), fz23.y);
    }
    r.h0 = r.n0 <= f0; r.h1 = r.n1 <= f1; r.h2 = r.n2 <= f2; r.h3 = r.n3 <= f3_;
    return r;
}
#define RT_PLANE_AT(NB, OFF) ld4((const float4*)((NB) + (OFF)))
#define RT_NODE_BASE(NP) const char* nb_ = (const char*)(NP);
#define RT_LOAD_NODE4(NP)                                                                                                       \
    RT_NODE_BASE(NP)                                                                                                            \
    const float4 pnx = RT_PLANE_AT(nb_, ri.nx), pny = RT_PLANE_AT(nb_, ri.ny), pnz = RT_PLANE_AT(nb_, ri.nz);                  \
    const float4 pfx = RT_PLANE_AT(nb_, ri.nx ^ 48), pfy = RT_PLANE_AT(nb_, 80 - ri.ny), pfz = RT_PLANE_AT(nb_, 112 - ri.nz);  \
    const int4 link = RT_LDG(&(NP)->link);

// The children a node step continues with.  Unordered (any-hit) walks: the first of l0..l3 that is a link to follow becomes
// `cur`, the others are pushed (v_k: child k is to be followed).  On the device the pushes are predicated stores at a running
// shared address when three entries still fit the shared part of the stack (no branches, no reconvergence points); the
// generic stk_push path handles the rare spill.
RT_HD void push_rest_unordered(TraceCtx& cx, int& cur, int& sp, int l0, int l1, int l2, int l3, bool v0, bool v1, bool v2, bool v3) {
    cur = v0 ? l0 : v1 ? l1 : v2 ? l2 : v3 ? l3 : RT_LINK_NONE;
    const bool t1 = v0 || v1, t2 = t1 || v2;
    const bool p1 = v1 && v0, p2 = v2 && t1, p3 = v3 && t2;
#if defined(__CUDA_ARCH__)
    if (sp + 3 <= RT_STACK_DEPTH) {
        unsigned a = cx.saddr + (unsigned)sp * RT_STACK_PITCH;
        if (p1) { asm volatile("st.shared.b32 [%0], %1;" :: "r"(a), "r"(l1)); a += RT_STACK_PITCH; }
        if (p2) { asm volatile("st.shared.b32 [%0], %1;" :: "r"(a), "r"(l2)); a += RT_STACK_PITCH; }
        if (p3) { asm volatile("st.shared.b32 [%0], %1;" :: "r"(a), "r"(l3)); }
        sp += (int)p1 + (int)p2 + (int)p3;
        return;
    }
#endif
    if (p1) stk_push(cx, sp, l1);
    if (p2) stk_push(cx, sp, l2);
    if (p3) stk_push(cx, sp, l3);
}
// Ordered (closest-hit) walks: a0 is entered next, a1..a3 (nearest first) are pushed farthest first; n = how many are valid.
// (The predicated form measured 3 % slower in k_hit than these three branches.)
RT_HD void push_rest_ordered(TraceCtx& cx, int& cur, int& sp, int a0, int a1, int a2, int a3, int n) {
    cur = n ? a0 : RT_LINK_NONE;
    if (n > 3) stk_push(cx, sp, a3);
    if (n > 2) stk_push(cx, sp, a2);
    if (n > 1) stk_push(cx, sp, a1);
}

// Advances the walk to the next leaf whose box the ray segment [tmin, tmax] touches; returns the leaf id
// (triangle slot / object index) or -1 when the walk is over.  `cur` = link being processed (RT_LINK_NONE: pop),
// stack entries [sbase, sp) are this walk's.  ORDERED: children entered nearest first (closest hit);
// FLAGS: children flagged RT_TLAS_NO_TMAX are tested without the tmax bound (shadow rays).
template <bool ORDERED, bool FLAGS>
RT_HD int bvh4_next(TraceCtx& cx, const BvhNode4* __restrict__ nodes, int& cur, int& sp, const int sbase, const RayInv& ri,
                    const float tmin, const float tmax, unsigned& visits) {
    for (;;) {
        if (cur == RT_LINK_NONE) {
            if (sp == sbase) return -1;
            cur = stk_pop(cx, sp);
        }
        if (cur < 0) { const int leaf = ~cur; cur = RT_LINK_NONE; return leaf; }
        const BvhNode4* np = nodes + cur;
        RT_LOAD_NODE4(np)
        visits++;
        float t0 = tmax, t1 = tmax, t2 = tmax, t3 = tmax;
        if (FLAGS && cx.S.tlas_flagged) {                      // no flagged child anywhere in the object BVH: the 16-byte flag row is not read
            const int4 fl = RT_LDG(&np->flags);
            if (fl.x & RT_TLAS_NO_TMAX) t0 = RT_INF;
            if (fl.y & RT_TLAS_NO_TMAX) t1 = RT_INF;
            if (fl.z & RT_TLAS_NO_TMAX) t2 = RT_INF;
            if (fl.w & RT_TLAS_NO_TMAX) t3 = RT_INF;
        }
        const Slab4 sl = slab4<true>(pnx, pny, pnz, pfx, pfy, pfz, ri.inv, ri.oinv, tmin, t0, t1, t2, t3);
        const float n0 = sl.n0, n1 = sl.n1, n2 = sl.n2, n3 = sl.n3; const bool h0 = sl.h0, h1 = sl.h1, h2 = sl.h2, h3 = sl.h3;
        if (ORDERED) {
            // sort (entry distance, slot) keys: entry distances are >= tmin >= 0, so their bit patterns order as integers
            unsigned k0 = h0 ? ((as_uint(n0) & 0x7ffffffcu) | 0u) : 0xffffffffu;
            unsigned k1 = h1 ? ((as_uint(n1) & 0x7ffffffcu) | 1u) : 0xffffffffu;
            unsigned k2 = h2 ? ((as_uint(n2) & 0x7ffffffcu) | 2u) : 0xffffffffu;
            unsigned k3 = h3 ? ((as_uint(n3) & 0x7ffffffcu) | 3u) : 0xffffffffu;
            key_swap(k0, k1); key_swap(k2, k3); key_swap(k0, k2); key_swap(k1, k3); key_swap(k1, k2);
            const int nhit = (int)h0 + (int)h1 + (int)h2 + (int)h3;
            push_rest_ordered(cx, cur, sp, link_of(link, k0 & 3u), link_of(link, k1 & 3u), link_of(link, k2 & 3u), link_of(link, k3 & 3u), nhit);
        } else {
            push_rest_unordered(cx, cur, sp, link.x, link.y, link.z, link.w, h0, h1, h2, h3);
        }
    }
}

RT_HD int first_bit(unsigned m) {
#if defined(__CUDA_ARCH__)
    return __ffs((int)m) - 1;
#else
    return __builtin_ffs((int)m) - 1;
#endif
}
RT_HD int sel4i(int a, int b, int c, int d, int k) { return (k & 2) ? ((k & 1) ? d : c) : ((k & 1) ? b : a); }
// bit 8k+7 set iff link k is negative (a leaf): the four sign bytes gathered with three byte permutes
RT_HD unsigned leaf_bits(int l0, int l1, int l2, int l3) {
#if defined(__CUDA_ARCH__)
    return __byte_perm(__byte_perm((unsigned)l0, (unsigned)l1, 0x0073), __byte_perm((unsigned)l2, (unsigned)l3, 0x0073), 0x5410) & 0x80808080u;
#else
    return ((unsigned)l0 >> 31) << 7 | ((unsigned)l1 >> 31) << 15 | ((unsigned)l2 >> 31) << 23 | ((unsigned)l3 >> 31) << 31;
#endif
}
RT_HD float sel4(float a, float b, float c, float d, int k) { return (k & 2) ? ((k & 1) ? d : c) : ((k & 1) ? b : a); }

// LBVH traversal of one mesh.  `onTri(slot, tmax)` tests triangle `slot`, may shrink tmax, returns true to stop.
// Replaces the face loops of mesh.py:77 and :125 (brute force kept as a mesh_mode bit for the equivalence test).
// ONE loop level: every iteration is a node step (four slab tests) followed by the triangle tests of that node's leaf
// children, so the lanes of a warp advance node by node in lock-step; an iterator that returns to the caller at every
// leaf makes each lane wait for the slowest lane's walk to ITS next leaf (measured: 12 of 32 lanes active).
template <bool ORDERED, class M, class OnTri>
RT_HD void mesh_visit(TraceCtx& cx, const M& m, const RayInv& ri, float tmin, float& tmax, OnTri&& onTri) {
    const int n = m.n_tris();
    if (n <= 0) return;
    if ((cx.S.mesh_mode & RT_MODE_BRUTE_MESH) || n == 1) {
        for (int s = 0; s < n; s++) { cx.tally.tris++; if (onTri(s, tmax)) return; }
        return;
    }
    const BvhNode4* __restrict__ nodes = m.nodes();
    const int sbase = cx.sbase;
    int cur = 0, sp = sbase;
    for (;;) {
        if (cur == RT_LINK_NONE) {
            if (sp == sbase) return;
            cur = stk_pop(cx, sp);
        }
        const BvhNode4* np = nodes + cur;
        RT_LOAD_NODE4(np)
        cx.tally.nodes++;
        // (the unordered walk is the mesh shadow test, which ignores t_max: its tmax is +inf and stays +inf, no far clip needed)
        const Slab4 sl = slab4<ORDERED>(pnx, pny, pnz, pfx, pfy, pfz, ri.inv, ri.oinv, tmin, tmax, tmax, tmax, tmax);
        const float n0 = sl.n0, n1 = sl.n1, n2 = sl.n2, n3 = sl.n3; bool h0 = sl.h0, h1 = sl.h1, h2 = sl.h2, h3 = sl.h3;
        // the links of the children whose boxes the ray touches (RT_LINK_NONE otherwise)
        const int l0 = h0 ? link.x : RT_LINK_NONE, l1 = h1 ? link.y : RT_LINK_NONE, l2 = h2 ? link.z : RT_LINK_NONE, l3 = h3 ? link.w : RT_LINK_NONE;
        // leaf children (negative links): their triangles are tested here and now
        unsigned lm = leaf_bits(l0, l1, l2, l3);
        h0 = (unsigned)l0 < (unsigned)RT_LINK_NONE; h1 = (unsigned)l1 < (unsigned)RT_LINK_NONE;
        h2 = (unsigned)l2 < (unsigned)RT_LINK_NONE; h3 = (unsigned)l3 < (unsigned)RT_LINK_NONE;
        while (lm) {
            const int k = first_bit(lm) >> 3;
            lm &= lm - 1u;
            if (ORDERED && sel4(n0, n1, n2, n3, k) > tmax) continue;        // an earlier leaf of this node already came closer
            const int code = ~sel4i(l0, l1, l2, l3, k);                      // first slot | (count - 1) << RT_LEAF_SLOT_BITS
            const int first = code & ((1 << RT_LEAF_SLOT_BITS) - 1), last = first + (code >> RT_LEAF_SLOT_BITS);
            for (int s = first; s <= last; s++) {
                cx.tally.tris++;
                if (onTri(s, tmax)) return;
            }
        }
        if (ORDERED) {
            h0 = h0 && n0 <= tmax; h1 = h1 && n1 <= tmax; h2 = h2 && n2 <= tmax; h3 = h3 && n3 <= tmax;
            unsigned k0 = h0 ? ((as_uint(n0) & 0x7ffffffcu) | 0u) : 0xffffffffu;
            unsigned k1 = h1 ? ((as_uint(n1) & 0x7ffffffcu) | 1u) : 0xffffffffu;
            unsigned k2 = h2 ? ((as_uint(n2) & 0x7ffffffcu) | 2u) : 0xffffffffu;
            unsigned k3 = h3 ? ((as_uint(n3) & 0x7ffffffcu) | 3u) : 0xffffffffu;
            key_swap(k0, k1); key_swap(k2, k3); key_swap(k0, k2); key_swap(k1, k3); key_swap(k1, k2);
            const int nhit = (int)h0 + (int)h1 + (int)h2 + (int)h3;
            push_rest_ordered(cx, cur, sp, sel4i(l0, l1, l2, l3, k0 & 3u), sel4i(l0, l1, l2, l3, k1 & 3u), sel4i(l0, l1, l2, l3, k2 & 3u), sel4i(l0, l1, l2, l3, k3 & 3u), nhit);
        } else {
            push_rest_unordered(cx, cur, sp, l0, l1, l2, l3, h0, h1, h2, h3);
        }
    }
}

// ------------------------------------------------------------------------------------------------
// Which top-level objects a ray has to be tested against, one at a time: replaces the plain
// `for obj in self.objects` of scene.py:87 and :160-164.  Objects without finite bounds come first (scene
// order), then the leaves of the object BVH whose (padded, motion-swept) boxes the ray segment [0, tmax]
// touches, nearer child first.  The order differs from the reference's list order, so the closest-hit
// consumer breaks equal-t ties by object index itself; the any-hit consumer is order-free.
struct TopIter { int lin_i; int cur; int sp; };
RT_HD TopIter top_begin() { TopIter it; it.lin_i = 0; it.cur = 0; it.sp = 0; return it; }
// the object BVH is not entered at all (the ray misses the scene's bounded part): the walk ends after the linear list
RT_HD void top_skip_bvh(TopIter& it) { it.cur = RT_LINK_NONE; }

// SHADOW: subtrees holding a mesh / CSG node are walked without the t_max bound because their shadow tests
// ignore it
template <bool SHADOW>
RT_HD int top_next(TraceCtx& cx, TopIter& it, const RayInv& ri, float tmax) {
    const SceneView& S = cx.S;
    if (S.mesh_mode & RT_MODE_LINEAR_TOP) return it.lin_i < S.n_top ? it.lin_i++ : -1;     // test aid: the reference's own loop (meshes skipped by the callers)
    if (it.lin_i < S.n_lin) return RT_LDG(S.lin_objs + it.lin_i++);
    if (S.tlas_nodes == nullptr) return -1;
    return bvh4_next<!SHADOW, SHADOW>(cx, S.tlas_nodes, it.cur, it.sp, 0, ri, 0.0f, tmax, cx.tally.prims);
}

template <class M>
RT_HD void tri_load(const M& m, int slot, float3& v0, float3& v1, float3& v2, int& face) {
    float4 a = ld4(m.tri() + 3 * (size_t)slot), b = ld4(m.tri() + 3 * (size_t)slot + 1), c = ld4(m.tri() + 3 * (size_t)slot + 2);
    v0 = f3(a); v1 = f3(b); v2 = f3(c); face = as_int(a.w);
}

// Shading normal of an accepted triangle hit (mesh.py:84,103-113); barycentrics as libigl computes them
template <class M>
RT_HD float3 tri_normal(TraceCtx& cx, const M& m, int slot, float3 p) {
    float3 v0, v1, v2; int face;
    tri_load(m, slot, v0, v1, v2, face);
    if (m.flat_shaded() || m.vnormals() == nullptr) return normalize(cross(v1 - v0, v2 - v0));
    cx.tally.smooth++;
    float3 e0 = v1 - v0, e1 = v2 - v0, e2 = p - v0;
    float d00 = dot(e0, e0), d01 = dot(e0, e1), d11 = dot(e1, e1), d20 = dot(e2, e0), d21 = dot(e2, e1);
    float denom = d00 * d11 - d01 * d01;
    float l1 = (d11 * d20 - d01 * d21) / denom;
    float l2 = (d00 * d21 - d01 * d20) / denom;
    float l0 = 1.0f - (l1 + l2);
    int4 vi = RT_LDG(m.face_vidx() + slot);
    float3 n0 = f3(ld4(m.vnormals() + vi.x)), n1 = f3(ld4(m.vnormals() + vi.y)), n2 = f3(ld4(m.vnormals() + vi.z));
    return normalize((n0 * l0 + n1 * l1) + n2 * l2);
}

// Mesh.shadow_intersect (mesh.py:121-153): gate, then any accepted face; t_max is ignored by the reference
template <class M>
RT_HD bool mesh_shadow(TraceCtx& cx, const M& m, const Ray& r, const RayInv& ri) {
    if (!mesh_gate_rcp(m, r, ri)) return false;
    bool occ = false;
    float tmax = RT_INF;
    mesh_visit<false>(cx, m, ri, RT_SHADOW_EPS * 0.5f, tmax, [&](int slot, float&) {
        float3 v0, v1, v2; int face;
        tri_load(m, slot, v0, v1, v2, face);
        if (tri_shadow(v0, v1, v2, r)) { occ = true; return true; }
        return false;
    });
    return occ;
}
// out-of-line copy for meshes under CSG nodes: keeps that traversal code out of the hot instruction stream
RT_HD_NOINLINE bool mesh_shadow_cold(TraceCtx& cx, const DevMesh* m, const Ray& r) { return mesh_shadow(cx, MeshG{m}, r, ray_inv(r)); }

// Closest accepted face of one top-level mesh against the running best (mesh.py:72-119 + scene.py:94).
// better(t): would a hit at t replace the current best of OTHER objects; among this mesh's own faces an equal t
// goes to the lower face index (list order).
template <class M, class Better>
RT_HD void mesh_closest(TraceCtx& cx, const M& m, const Ray& r, const RayInv& ri, float best, Better&& better, float& t_out, int& slot_out) {
    slot_out = -1; t_out = best;
    if (!mesh_gate_rcp(m, r, ri)) return;
    float tmax = best;
    int best_face = 0x7fffffff, best_slot = -1;
    mesh_visit<true>(cx, m, ri, 0.0f, tmax, [&](int slot, float& tm) {
        float3 v0, v1, v2; int face; float t;
        tri_load(m, slot, v0, v1, v2, face);
        if (tri_hit(v0, v1, v2, r, t)) {
            if ((best_slot < 0 && better(t)) || (best_slot >= 0 && (t < tm || (t == tm && face < best_face)))) {
                tm = t; best_face = face; best_slot = slot;
            }
        }
        return false;
    });
    slot_out = best_slot; t_out = tmax;
}
// ------------------------------------------------------------------------------------------------
// Texture / diffuse lookups (scene.py:143-146)
RT_HD float3 texel(const DevTexture* t, long long i, long long j) {
    i = i < 0 ? 0 : (i >= t->width ? t->width - 1 : i);     // the reference would raise IndexError; clamp
    j = j < 0 ? 0 : (j >= t->height ? t->height - 1 : j);
    const uint8_t* px = t->rgb + ((size_t)j * (size_t)t->width + (size_t)i) * 3;
    return f3((float)(RT_LDG(px) / 255.0), (float)(RT_LDG(px + 1) / 255.0), (float)(RT_LDG(px + 2) / 255.0));
}
RT_HD double py_mod(double a, double m) { double r = fmod(a, m); return (r != 0.0 && r < 0.0) ? r + m : r; }

// Plane.get_diffuse (simple_geometry.py:150-173): tiled nearest texel, or the checker material's diffuse
RT_HD float3 plane_diffuse(const SceneView& S, const DevObject* o, float3 p, float time, Tally& tally) {
    if (o->tex >= 0) {
        const DevTexture* tx = S.textures + o->tex;
        float3 n = f3(ld4(&o->b));
        float3 base = f3(ld4(&o->a));
        float3 position = moved(o, base, time);
        float3 pp = p - n * dot(p - base, n);                 // projection uses the UN-moved point (:163)
        double u = (double)dot(pp - position, f3(ld4(&o->ax_w))) * 1000.0 / o->tex_scale;
        double v = (double)dot(pp - position, f3(ld4(&o->ax_h))) * 1000.0 / o->tex_scale;
        return texel(tx, (long long)py_mod(u, (double)tx->width), (long long)py_mod(v, (double)tx->height));
    }
    int m = plane_material(o, p, time);
    if (m < 0) { tally.err |= 1u << RT_ERR_NO_MATERIAL; return f3(0, 0, 0); }
    return f3(ld4(&S.materials[m].diffuse));
}
// AABB.get_diffuse (simple_geometry.py:312-355): cube-face mapping (PIL truncates the float indices)
RT_HD float3 box_diffuse(const SceneView& S, const DevObject* o, float3 p, float time, Tally& tally) {
    if (o->tex >= 0) {
        const DevTexture* tx = S.textures + o->tex;
        float3 mn = moved(o, f3(ld4(&o->a)), time), mx = moved(o, f3(ld4(&o->b)), time);
        double W = tx->width, H = tx->height;
        double x = ((double)p.x - mn.x) / ((double)mx.x - mn.x);
        double y = ((double)p.y - mn.y) / ((double)mx.y - mn.y);
        double z = ((double)p.z - mn.z) / ((double)mx.z - mn.z);
        double i, j;
        if (fabs((double)p.x - mn.x) < 1e-4) { i = z * W; j = (1 - y) * H; }
        else if (fabs((double)p.x - mx.x) < 1e-4) { i = (1 - z) * W; j = (1 - y) * H; }
        else if (fabs((double)p.y - mn.y) < 1e-4) { i = x * W; j = (1 - z) * H; }
        else if (fabs((double)p.y - mx.y) < 1e-4) { i = x * W; j = z * H; }
        else if (fabs((double)p.z - mn.z) < 1e-4) { i = (1 - x) * W; j = (1 - y) * H; }
        else if (fabs((double)p.z - mx.z) < 1e-4) { i = x * W; j = (1 - y) * H; }
        else { i = 0; j = 0; }
        i = fmin(fmax(0.0, i), W - 1);
        j = fmin(fmax(0.0, j), H - 1);
        return texel(tx, (long long)i, (long long)j);
    }
    if (o->n_mats < 1) { tally.err |= 1u << RT_ERR_NO_MATERIAL; return f3(0, 0, 0); }
    return f3(ld4(&S.materials[o->mat0].diffuse));
}

// ------------------------------------------------------------------------------------------------
// CSG / transform nodes (hierarchy.py:42-138).

// shadow_intersect of a leaf primitive (simple_geometry.py:48-72,122-131,251-294; mesh.py:121-153)
template <bool WITH_MESH>
RT_HD bool leaf_shadow(TraceCtx& cx, int oi, const Ray& r, float tmax) {
    const DevObject* o = cx.S.objects + oi;
    cx.tally.prims++;
    switch (o->kind) {
    case RT_KIND_SPHERE: {                                     // simple_geometry.py:48-72
        float4 a = ld4(&o->a);
        float t1, t2;
        if (!sphere_roots(moved(cx, o, f3(a)), a.w, r, dot(r.d, r.d), t1, t2)) return false;
        return (RT_SHADOW_EPS_SPHERE < t1 && t1 < tmax) || (RT_SHADOW_EPS_SPHERE < t2 && t2 < tmax);
    }
    case RT_KIND_PLANE: {                                      // simple_geometry.py:122-131
        float t;
        if (!plane_t(moved(cx, o, f3(ld4(&o->a))), f3(ld4(&o->b)), r, t)) return false;
        return RT_SHADOW_EPS < t && t < tmax;
    }
    case RT_KIND_BOX: {                                        // simple_geometry.py:251-294 (entry t only, may be < 0)
        float en, ex; int label;
        if (!box_slabs(moved(cx, o, f3(ld4(&o->a))), moved(cx, o, f3(ld4(&o->b))), r, en, ex, label)) return false;
        if (en > ex) return false;
        return RT_SHADOW_EPS < en && en < tmax;
    }
    case RT_KIND_MESH: {                                       // top-level meshes: see occluded()
        if (!WITH_MESH) return false;
        Ray mr = r; mr.o = r.o - mesh_shift(cx, oi);
        return mesh_shadow_cold(cx, cx.S.meshes + o->mesh, mr);
    }
            default: return false;
    }
}

template <int V> RT_HD_NOINLINE bool difference_shadow(TraceCtx& cx, int c0, const Ray& mr);

// The recursions of hierarchy.py (is_inside :111-129, get_material :131-138, shadow_intersect :80-109) run as loops over
// an explicit frame stack of RT_MAX_NEST entries: ONE copy of each function in the instruction stream (the first version
// unrolled the recursion through templates, eight copies of everything, and the node kernels were instruction-fetch bound).
// Nesting deeper than RT_MAX_NEST is flagged (RT_ERR_LIMIT) and treated as empty.

// Pruning by the children's bounds (SceneView::child_lo/hi, parent space).  Both tests only ever say "cannot": the
// boxes are padded, an unbounded or top-level object carries [-inf, +inf] and always passes.
RT_HD float3 csg_ray_inv(const Ray& r) { return f3(safe_rcp(r.d.x), safe_rcp(r.d.y), safe_rcp(r.d.z)); }
RT_HD bool child_may_hit(const SceneView& S, int oi, const Ray& r, float3 inv) {
#ifdef RT_NO_CHILD_BOUNDS      // test aid (tests/hostsim variant): the pruning must not change any image
    return true;
#endif
    const float4 lo = ld4(S.child_lo + oi), hi = ld4(S.child_hi + oi);
    if (lo.x == -RT_INF) return true;
    const float ax = (lo.x - r.o.x) * inv.x, bx = (hi.x - r.o.x) * inv.x;
    const float ay = (lo.y - r.o.y) * inv.y, by = (hi.y - r.o.y) * inv.y;
    const float az = (lo.z - r.o.z) * inv.z, bz = (hi.z - r.o.z) * inv.z;
    const float entry = fmaxf(fmaxf(fminf(ax, bx), fminf(ay, by)), fmaxf(fminf(az, bz), 0.0f));
    const float exit_ = fminf(fminf(fmaxf(ax, bx), fmaxf(ay, by)), fmaxf(az, bz));
    return entry <= exit_;
}
RT_HD bool child_may_contain(const SceneView& S, int oi, float3 p) {
#ifdef RT_NO_CHILD_BOUNDS
    return true;
#endif
    const float4 lo = ld4(S.child_lo + oi), hi = ld4(S.child_hi + oi);
    return !(p.x < lo.x || p.x > hi.x || p.y < lo.y || p.y > hi.y || p.z < lo.z || p.z > hi.z);
}

// is_inside of a leaf (Sphere simple_geometry.py:74-80, AABB :296-307; Plane/Mesh: base class False)
RT_HD bool leaf_inside(const TraceCtx& cx, const DevObject* o, float3 p) {
    if (o->kind == RT_KIND_SPHERE) {
        float4 a = ld4(&o->a);
        float3 q = p - moved(cx, o, f3(a));
        return sqrtf(dot(q, q)) < a.w;
    }
    if (o->kind == RT_KIND_BOX) {
        float3 mn = moved(cx, o, f3(ld4(&o->a))), mx = moved(cx, o, f3(ld4(&o->b)));
        return mn.x < p.x && p.x < mx.x && mn.y < p.y && p.y < mx.y && mn.z < p.z && p.z < mx.z;
    }
    return false;
}

// How many children a node's boolean looks at, and what it is worth with none: union/intersection all of them
// (any([]) = False, all([]) = True), difference exactly two (fewer: nothing).
RT_HD int csg_arity(const DevObject* nd) {
    return (nd->csg == RT_CSG_UNION || nd->csg == RT_CSG_INTERSECTION) ? nd->n_children : ((nd->csg == RT_CSG_DIFFERENCE && nd->n_children >= 2) ? 2 : 0);
}
// After child number `k` of `nd` evaluated to `val`: true = the node's value is decided (in `val`), false = evaluate child k+1
RT_HD bool csg_step(const DevObject* nd, int k, int arity, bool& val) {
    if (nd->csg == RT_CSG_UNION) { if (val) return true; if (k + 1 < arity) return false; val = false; return true; }
    if (nd->csg == RT_CSG_INTERSECTION) { if (!val) return true; if (k + 1 < arity) return false; val = true; return true; }
    if (k == 0) { if (!val) return true; return false; }      // difference: c0 and not c1
    val = !val; return true;
}

struct InsideFrame { int obj, cursor; float3 q; };
RT_HD_NOINLINE bool csg_inside(TraceCtx& cx, int root, float3 p) {
    const DevObject* objs = cx.S.objects;
    InsideFrame st[RT_MAX_NEST];
    int d = -1, cur = root;
    float3 cp = p;
    for (;;) {
        const DevObject* o = objs + cur;
        bool val;
        if (!child_may_contain(cx.S, cur, cp)) val = false;      // outside the child's bound: not inside it
        else if (o->kind == RT_KIND_NODE) {
            const int ar = csg_arity(o);
            if (d + 1 >= RT_MAX_NEST) { cx.tally.err |= 1u << RT_ERR_LIMIT; val = false; }
            else if (ar == 0) val = o->csg == RT_CSG_INTERSECTION;
            else {
                d++;
                st[d].obj = cur; st[d].cursor = 0; st[d].q = xform_point(cx.S.xforms[o->xform].Minv, cp);
                cp = st[d].q; cur = o->first_child;
                continue;
            }
        } else val = leaf_inside(cx, o, cp);
        for (;;) {                                               // hand the value up until a node wants its next child
            if (d < 0) return val;
            const DevObject* nd = objs + st[d].obj;
            if (csg_step(nd, st[d].cursor, csg_arity(nd), val)) { d--; continue; }
            st[d].cursor++;
            cur = nd->first_child + st[d].cursor; cp = st[d].q;
            break;
        }
    }
}

// get_material (hierarchy.py:131-138: the first child that contains the point answers; Plane :133; others materials[0])
RT_HD_NOINLINE int csg_material(TraceCtx& cx, int oi, float3 p) {
    const DevObject* objs = cx.S.objects;
    for (int depth = 0;; depth++) {
        const DevObject* o = objs + oi;
        if (o->kind == RT_KIND_PLANE) return plane_material(o, p, cx.time);
        if (o->kind != RT_KIND_NODE) return o->n_mats > 0 ? o->mat0 : -2;
        if (depth >= RT_MAX_NEST) { cx.tally.err |= 1u << RT_ERR_LIMIT; return -1; }
        const float3 q = xform_point(cx.S.xforms[o->xform].Minv, p);
        int next = -1;
        for (int k = 0; k < o->n_children && next < 0; k++)
            if (csg_inside(cx, o->first_child + k, q)) next = o->first_child + k;
        if (next < 0) return -1;   // None
        oi = next; p = q;
    }
}

// shadow_intersect (hierarchy.py:80-109 and the leaf versions): union = any child, intersection = ALL children report
// shadow, difference goes through intersect() and ignores t_max
struct ShadowFrame { int obj, cursor; Ray ray; float3 inv; };
template <int V>
RT_HD_NOINLINE bool csg_shadow(TraceCtx& cx, int root, const Ray& r, float tmax) {
    const DevObject* objs = cx.S.objects;
    ShadowFrame st[RT_MAX_NEST];
    int d = -1, cur = root;
    Ray cr = r;
    float3 cinv = csg_ray_inv(r);
    for (;;) {
        const DevObject* o = objs + cur;
        bool val;
        if (!child_may_hit(cx.S, cur, cr, cinv)) val = false;    // the ray misses the child's bound: it casts no shadow on it
        else if (o->kind == RT_KIND_NODE) {
            cx.tally.prims++;
            if (d + 1 >= RT_MAX_NEST) { cx.tally.err |= 1u << RT_ERR_LIMIT; val = false; }
            else {
                const DevXform* xf = cx.S.xforms + o->xform;
                Ray mr; mr.o = xform_point(xf->Minv, cr.o); mr.d = xform_dir(xf->Minv, cr.d);
                const int ar = csg_arity(o);
                if (ar == 0) val = o->csg == RT_CSG_INTERSECTION;
                else if (o->csg == RT_CSG_DIFFERENCE) val = difference_shadow<V>(cx, o->first_child, mr);
                else {
                    d++;
                    cinv = csg_ray_inv(mr);
                    st[d].obj = cur; st[d].cursor = 0; st[d].ray = mr; st[d].inv = cinv;
                    cr = mr; cur = o->first_child;
                    continue;
                }
            }
        } else val = leaf_shadow<(V != RT_V_NODES)>(cx, cur, cr, tmax);
        for (;;) {
            if (d < 0) return val;
            const DevObject* nd = objs + st[d].obj;
            if (csg_step(nd, st[d].cursor, csg_arity(nd), val)) { d--; continue; }
            st[d].cursor++;
            cur = nd->first_child + st[d].cursor; cr = st[d].ray; cinv = st[d].inv;
            break;
        }
    }
}

// A mesh as a node's child (mesh.py:72-119: every accepted face).  Kept out of line so that the LBVH walk is not part of
// every node kernel's hot instruction stream.
template <class Want, class Cand>
RT_HD_NOINLINE void mesh_leaf_hits(TraceCtx& cx, int oi, const Ray& r, Want& want, Cand& cand) {
    const DevObject* o = cx.S.objects + oi;
    const MeshG m{cx.S.meshes + o->mesh};
    const float3 shift = mesh_shift(cx, oi);
    Ray mr = r; mr.o = r.o - shift;                                 // the mesh's rest frame (== r unless the mesh moves)
    if (!mesh_gate(m, mr)) return;
    float tmax = RT_INF;
    const int mat = o->n_mats > 0 ? o->mat0 : -2;
    const RayInv ri = ray_inv(mr);
    mesh_visit<true>(cx, m, ri, 0.0f, tmax, [&](int slot, float&) {
        float3 v0, v1, v2; int face; float t;
        tri_load(m, slot, v0, v1, v2, face);
        if (tri_hit(v0, v1, v2, mr, t) && want(t)) {
            Hit h; h.t = t; h.p = r.o + r.d * t; h.n = tri_normal(cx, m, slot, mr.o + mr.d * t); h.mat = mat; h.geom = oi;
            cand(h);
        }
        return false;
    });
}

// Candidate hits of ONE leaf primitive for `r`, in the reference's list order; `want(t)` prunes early.  The (at most two)
// hit times of an analytic leaf are collected first and handed to `cand` from ONE call site: `cand` is the whole climb
// through the ancestors' CSG filters, and a copy of it per primitive kind tripled the size of the node kernels.
template <int V, class Want, class Cand>
RT_HD void leaf_hits(TraceCtx& cx, int oi, const Ray& r, Want&& want, Cand&& cand) {
    const DevObject* o = cx.S.objects + oi;
    cx.tally.prims++;
    const int kind = o->kind;
    if (V != RT_V_NODES && kind == RT_KIND_MESH) { TraceCtx gx = cx.global(); mesh_leaf_hits(gx, oi, r, want, cand); return; }   // out of line: rare under nodes, and large
    float ts[2]; int nt = 0;
    float3 nrm = f3(0, 0, 0), center = f3(0, 0, 0);
    if (kind == RT_KIND_SPHERE) {                                  // simple_geometry.py:20-46: both roots with t > 0
        float4 a = ld4(&o->a);
        center = moved(cx, o, f3(a));
        float t1, t2;
        if (!sphere_roots(center, a.w, r, dot(r.d, r.d), t1, t2)) return;
        if (t1 > 0.0f) ts[nt++] = t1;
        if (t2 > 0.0f) ts[nt++] = t2;
    } else if (kind == RT_KIND_PLANE) {                            // simple_geometry.py:105-120
        nrm = f3(ld4(&o->b)); float t;
        if (!plane_t(moved(cx, o, f3(ld4(&o->a))), nrm, r, t)) return;
        if (t >= 0.0f) ts[nt++] = t;
    } else if (kind == RT_KIND_BOX) {                              // simple_geometry.py:188-249: entry AND exit, both with the entry normal
        float en, ex; int label;
        if (!box_slabs(moved(cx, o, f3(ld4(&o->a))), moved(cx, o, f3(ld4(&o->b))), r, en, ex, label)) return;
        if (en > ex || en < 0.0f) return;
        nrm = box_normal(r, label);
        ts[0] = en; ts[1] = ex; nt = 2;
    }
    for (int k = 0; k < nt; k++) {
        const float t = k == 0 ? ts[0] : ts[1];
        if (!want(t)) continue;
        Hit h; h.t = t; h.p = r.o + r.d * t; h.geom = oi;
        h.n = kind == RT_KIND_SPHERE ? normalize(h.p - center) : nrm;
        h.mat = kind == RT_KIND_PLANE ? plane_material(o, h.p, cx.time) : (o->n_mats > 0 ? o->mat0 : -2);
        cand(h);
    }
}

// intersect() of an arbitrary object (hierarchy.py:42-78 for nodes): every hit of `root` for `r`,
// positions/normals in the space `r` lives in.  Depth-first over the node tree with an explicit frame
// stack; a candidate from a leaf is carried UP through its ancestors' CSG filters and transforms
// (times never change on the way up, so `want(t)` prunes before any of that work).  The reference
// enumerates hits in list order; consumers here are order-free (strict-< running minimum in DFS
// order == list order; boolean any).
struct NodeFrame { int obj; int cursor; Ray ray; float3 inv; };

// ROOT_IS_NODE: the caller knows that `root` is a node (closest_hit): the leaf-root branch - a second copy of leaf_hits and of the
// whole climb - is not compiled (the node kernels are instruction-fetch bound)
template <int V, bool ROOT_IS_NODE = false, class Want, class Emit>
RT_HD void object_hits(TraceCtx& cx, int root, const Ray& r, Want&& want, Emit&& emit) {
    const DevObject* objs = cx.S.objects;
    NodeFrame fr[RT_MAX_NEST];
    int depth = -1;
    int leaf = root;
    TraceCtx gx = cx.global();                         // for the out-of-line is_inside / get_material calls of the climb
    auto cand = [&](Hit h) {
        int child = leaf;
        for (int lv = depth; lv >= 0; lv--) {
            const DevObject* nd = objs + fr[lv].obj;
            const int c0 = nd->first_child, nc = nd->n_children;
            if (nd->csg == RT_CSG_INTERSECTION) {                  // hierarchy.py:52-57: every OTHER child contains it
                for (int j = 0; j < nc; j++)
                    if (c0 + j != child && !csg_inside(gx, c0 + j, h.p)) return;
            } else if (nd->csg == RT_CSG_DIFFERENCE) {             // hierarchy.py:58-70
                if (child == c0) { if (csg_inside(gx, c0 + 1, h.p)) return; }
                else {
                    if (!csg_inside(gx, c0, h.p)) return;
                    h.mat = csg_material(gx, c0, h.p);
                    h.n = -h.n;
                }
            }
            if (h.mat == -1) h.mat = nd->n_mats > 0 ? nd->mat0 : -2;    // hierarchy.py:73-74
            const DevXform* xf = gx.S.xforms + nd->xform;      // (gx, not cx: this closure travels to the out-of-line mesh_leaf_hits)
            h.p = xform_point(xf->M, h.p);                          // hierarchy.py:75
            h.n = xform_normal(xf->Minv, h.n);                      // hierarchy.py:76
            child = fr[lv].obj;
        }
        emit(h);
    };
    if (!child_may_hit(cx.S, root, r, csg_ray_inv(r))) return;      // (roots that are top-level objects always pass)
    if (!ROOT_IS_NODE && objs[root].kind != RT_KIND_NODE) { leaf_hits<V>(cx, root, r, want, cand); return; }
    auto push = [&](int oi, const Ray& pr) {
        const DevXform* xf = cx.S.xforms + objs[oi].xform;
        depth++;
        fr[depth].obj = oi; fr[depth].cursor = 0;
        fr[depth].ray.o = xform_point(xf->Minv, pr.o);              // hierarchy.py:43-45 (direction NOT renormalised)
        fr[depth].ray.d = xform_dir(xf->Minv, pr.d);
        fr[depth].inv = csg_ray_inv(fr[depth].ray);
        cx.tally.prims++;
    };
    push(root, r);
    while (depth >= 0) {
        const DevObject* nd = objs + fr[depth].obj;
        const int csg = nd->csg, nc = nd->n_children;
        const int limit = (csg == RT_CSG_UNION || csg == RT_CSG_INTERSECTION) ? nc : ((csg == RT_CSG_DIFFERENCE && nc >= 2) ? 2 : 0);
        if (fr[depth].cursor >= limit) { depth--; continue; }
        const int child = nd->first_child + fr[depth].cursor++;
        if (!child_may_hit(cx.S, child, fr[depth].ray, fr[depth].inv)) continue;   // no hit of this child (or anything below it) is possible
        if (objs[child].kind == RT_KIND_NODE) {
            if (depth + 1 >= RT_MAX_NEST) { cx.tally.err |= 1u << RT_ERR_LIMIT; continue; }
            Ray pr = fr[depth].ray;
            push(child, pr);
        } else {
            leaf = child;
            leaf_hits<V>(cx, child, fr[depth].ray, want, cand);
        }
    }
}

// Hierarchy.shadow_intersect, difference case (hierarchy.py:94-107): goes through intersect() of both
// children, `time > shadow_epsilon`, and ignores t_max.
template <int V>
RT_HD_NOINLINE bool difference_shadow(TraceCtx& cx, int c0, const Ray& mr) {
    bool occ = false;
    for (int side = 0; side < 2 && !occ; side++)       // hits of c0 outside c1, then hits of c1 inside c0 (one instantiation of the walk)
        object_hits<V>(cx, c0 + side, mr, [&](float t) { return !occ && t > RT_SHADOW_EPS; }, [&](const Hit& h) {
            if (!occ && csg_inside(cx, c0 + 1 - side, h.p) == (side == 1)) occ = true;
        });
    return occ;
}

// Header of the k-th top-level mesh (the meshes traversed after the object walk).  The lean kernels (no CSG / transform nodes)
// read it from the parameter bank: the pointer chase sep_mesh[k] -> objects[oi].mesh -> meshes[] -> nodes / triangles (four dependent
// L1 round trips in front of every mesh walk, and the triangle array's address again at every leaf) becomes constant-bank operands.
// Scenes with more than RT_CONST_MESH top-level meshes run on the general kernels (convert_scene).
template <int HAS_NODES>
RT_HD int top_mesh_obj(const SceneView& S, int k) { return HAS_NODES ? RT_LDG(S.sep_mesh + k) : S.c.sep_obj[k]; }
template <int HAS_NODES> struct TopMesh { typedef MeshG type; static RT_HD MeshG get(const SceneView& S, int, int oi) { return MeshG{S.meshes + S.objects[oi].mesh}; } };
template <> struct TopMesh<0> { typedef MeshC type; static RT_HD MeshC get(const SceneView& S, int k, int) { return MeshC{S, k}; } };

// ------------------------------------------------------------------------------------------------
// Closest hit over the top-level objects (scene.py:85-94): min by time, ties -> first in list order.
// Leaves are resolved lazily (only t while scanning; position/normal/material for the winner).
template <int HAS_NODES>
RT_HD bool closest_hit(TraceCtx& cx, const Ray& r, Hit& out) {
    const SceneView& S = cx.S;
    float best = RT_INF;
    int best_obj = 0x7fffffff, best_aux = 0, best_k = 0;
    Hit node_hit; node_hit.t = RT_INF; node_hit.mat = -1; node_hit.geom = -1; node_hit.n = f3(0, 0, 0); node_hit.p = f3(0, 0, 0);
    const float A = dot(r.d, r.d);
    const RayInv ri = ray_inv(r);
    TopIter it = top_begin();
    // rays that miss the padded bound of everything bounded (object BVH + separate meshes) only see the linear list (planes)
    const bool in_world = !S.has_world || (S.mesh_mode & RT_MODE_LINEAR_TOP) || touches_world(S, ri, RT_INF);
    if (!in_world) top_skip_bvh(it);
    // a candidate at time t replaces the current best iff it is earlier, or equally early and earlier in scene.objects
    #define RT_BETTER(t) ((t) < best || ((t) == best && oi < best_obj))
    if (!HAS_NODES && S.c.n_lin_planes >= 0 && !(S.mesh_mode & RT_MODE_LINEAR_TOP)) {      // the linear list is all planes: point / normal / speed from the parameter bank
        for (int k = 0; k < S.c.n_lin_planes; k++) {
            const int oi = S.c.lin_obj[k];
            cx.tally.prims++;
            float3 pt = f3(S.c.lin_point[k]);
            if (S.any_motion) pt = pt + f3(S.c.lin_speed[k]) * cx.time;
            float t;
            if (plane_t(pt, f3(S.c.lin_normal[k]), r, t) && t >= 0.0f && RT_BETTER(t)) { best = t; best_obj = oi; }
        }
        it.lin_i = S.n_lin;
    }
    for (;;) {
        const int oi = top_next<false>(cx, it, ri, best);
        if (oi < 0) break;
        cx.sbase = it.sp;
        const DevObject* o = S.objects + oi;
        const int kind = o->kind;
        if (kind == RT_KIND_SPHERE) {
            cx.tally.prims++;
            float4 a = ld4(&o->a);
            float t1, t2;
            if (sphere_roots(moved(cx, o, f3(a)), a.w, r, A, t1, t2)) {
                float t = t1 > 0.0f ? t1 : t2;                       // t1 <= t2: first positive root is the list minimum
                if (t > 0.0f && RT_BETTER(t)) { best = t; best_obj = oi; }
            }
        } else if (kind == RT_KIND_PLANE) {
            cx.tally.prims++;
            float t;
            if (plane_t(moved(cx, o, f3(ld4(&o->a))), f3(ld4(&o->b)), r, t) && t >= 0.0f && RT_BETTER(t)) { best = t; best_obj = oi; }
        } else if (kind == RT_KIND_BOX) {
            cx.tally.prims++;
            float en, ex; int label;
            if (box_slabs(moved(cx, o, f3(ld4(&o->a))), moved(cx, o, f3(ld4(&o->b))), r, en, ex, label) &&
                !(en > ex || en < 0.0f) && RT_BETTER(en)) { best = en; best_obj = oi; best_aux = label; }
        } else if (HAS_NODES && kind == RT_KIND_NODE) {
            // inside one node the first hit in DFS (= list) order wins equal t: strict '<' once this node holds the best
            object_hits<HAS_NODES, true>(cx, oi, r, [&](float t) { return RT_BETTER(t); },
                        [&](const Hit& h) { if (RT_BETTER(h.t)) { best = h.t; best_obj = oi; node_hit = h; } });
        }
    }
    cx.sbase = 0;
    if (HAS_NODES != RT_V_NODES)
    for (int k = 0; k < S.n_sep_mesh && in_world; k++) {              // top-level meshes: all lanes of the warp enter the LBVH together
        {
            const int oi = top_mesh_obj<HAS_NODES>(S, k);
            const typename TopMesh<HAS_NODES>::type m = TopMesh<HAS_NODES>::get(S, k, oi);
            cx.tally.prims++;
            float t; int slot;
            // (moving meshes run on the GENERAL kernels only - HAS_NODES, see convert_scene - so the lean ones carry none of this)
            Ray mr = r; RayInv mri = ri;
            if (HAS_NODES && S.mesh_motion) { mr.o = r.o - mesh_shift(cx, oi); mri = ray_inv_moved(ri, mr.o); }
            mesh_closest(cx, m, mr, mri, best, [&](float tt) { return RT_BETTER(tt); }, t, slot);
            if (slot >= 0) { best = t; best_obj = oi; best_aux = slot; best_k = k; }
        }
    }
    #undef RT_BETTER
    if (best_obj == 0x7fffffff) return false;
    const DevObject* o = S.objects + best_obj;
    if (HAS_NODES && o->kind == RT_KIND_NODE) { out = node_hit; return true; }
    out.t = best; out.geom = best_obj;
    out.p = r.o + r.d * best;                                          // Ray.getPoint, helperclasses.py:21
    out.mat = o->n_mats > 0 ? o->mat0 : -2;
    switch (o->kind) {
        case RT_KIND_SPHERE: out.n = normalize(out.p - moved(cx, o, f3(ld4(&o->a)))); break;
        case RT_KIND_PLANE: out.n = f3(ld4(&o->b)); out.mat = plane_material(o, out.p, cx.time); break;
        case RT_KIND_BOX: out.n = box_normal(r, best_aux); break;
        default: if (HAS_NODES != RT_V_NODES) out.n = tri_normal(cx, TopMesh<HAS_NODES>::get(S, best_k, best_obj), best_aux, (HAS_NODES && S.mesh_motion) ? (r.o - mesh_shift(cx, best_obj)) + r.d * best : out.p); break;
    }
    return true;
}

// shadow_intersect of ONE top-level object (the body of the loop of scene.py:160-164); A = d.d
template <int HAS_NODES>
RT_HD bool object_shadow(TraceCtx& cx, int oi, const Ray& r, float A, float tmax) {
    const DevObject* o = cx.S.objects + oi;
    const int kind = o->kind;
    if (kind == RT_KIND_SPHERE) {
        cx.tally.prims++;
        float4 a = ld4(&o->a);
        float t1, t2;
        return sphere_roots(moved(cx, o, f3(a)), a.w, r, A, t1, t2) &&
               ((RT_SHADOW_EPS_SPHERE < t1 && t1 < tmax) || (RT_SHADOW_EPS_SPHERE < t2 && t2 < tmax));
    }
    if (kind == RT_KIND_MESH) return false;                           // (linear test mode only) traversed after the walk
    if (kind != RT_KIND_NODE) return leaf_shadow<false>(cx, oi, r, tmax);
    if constexpr (HAS_NODES != 0) { TraceCtx gx = cx.global(); return csg_shadow<HAS_NODES>(gx, oi, r, tmax); }
    else return false;
}

// Any-hit over the top-level objects (scene.py:160-164)
template <int HAS_NODES>
RT_HD bool occluded(TraceCtx& cx, const Ray& r, float tmax) {
    const SceneView& S = cx.S;
    const float A = dot(r.d, r.d);
    const RayInv ri = ray_inv(r);
    TopIter it = top_begin();
    bool occ = false;
    // mesh and CSG-difference shadows ignore t_max, so the world test is made with the unbounded ray
    const bool in_world = !S.has_world || (S.mesh_mode & RT_MODE_LINEAR_TOP) || touches_world(S, ri, RT_INF);
    if (!in_world) top_skip_bvh(it);
    if (!HAS_NODES && S.c.n_lin_planes >= 0 && !(S.mesh_mode & RT_MODE_LINEAR_TOP)) {      // the linear list is all planes (simple_geometry.py:122-131 from the parameter bank)
        for (int k = 0; k < S.c.n_lin_planes; k++) {
            cx.tally.prims++;
            float3 pt = f3(S.c.lin_point[k]);
            if (S.any_motion) pt = pt + f3(S.c.lin_speed[k]) * cx.time;
            float t;
            if (plane_t(pt, f3(S.c.lin_normal[k]), r, t) && RT_SHADOW_EPS < t && t < tmax) return true;
        }
        it.lin_i = S.n_lin;
    }
    for (;;) {
        const int oi = top_next<true>(cx, it, ri, tmax);
        if (oi < 0) break;
        cx.sbase = it.sp;
        if (object_shadow<HAS_NODES>(cx, oi, r, A, tmax)) { occ = true; break; }
    }
    cx.sbase = 0;
    if (HAS_NODES != RT_V_NODES)
    for (int k = 0; k < S.n_sep_mesh && !occ && in_world; k++) {
        cx.tally.prims++;
        const int oi = top_mesh_obj<HAS_NODES>(S, k);
        Ray mr = r; RayInv mri = ri;
        if (HAS_NODES && S.mesh_motion) { mr.o = r.o - mesh_shift(cx, oi); mri = ray_inv_moved(ri, mr.o); }
        occ = mesh_shadow(cx, TopMesh<HAS_NODES>::get(S, k, oi), mr, mri);
    }
    return occ;
}

// ------------------------------------------------------------------------------------------------
// Shadow ray of light `li` from hit point p (scene.py:148-167): un-offset origin, un-normalised direction (Q12)
// DevLight fields of light `li`: the first RT_CONST_LIGHTS from the parameter bank (SceneConst)
// (CONST = false: always the global record - the general kernels are instruction-fetch bound and take no second code path)
template <bool CONST = true>
RT_HD float4 light_vector(const SceneView& S, int li) { return (CONST && li < RT_CONST_LIGHTS) ? S.c.light_vec[li] : ld4(&S.lights[li].vector); }
RT_HD float4 light_colour_power(const SceneView& S, int li) { return li < RT_CONST_LIGHTS ? S.c.light_colour[li] : ld4(&S.lights[li].colour_power); }
RT_HD float4 light_dir_norm(const SceneView& S, int li) { return li < RT_CONST_LIGHTS ? S.c.light_dirn[li] : ld4(&S.lights[li].dir_norm); }

template <int HAS_NODES>
RT_HD bool light_occluded(TraceCtx& cx, float3 p, int li) {
    const float4 lv = light_vector<!HAS_NODES>(cx.S, li);
    const bool point = as_int(lv.w) == RT_LIGHT_POINT;
    Ray sr; sr.o = p;
    sr.d = point ? f3(lv) - p : -f3(lv);
    cx.tally.shadow++;
    return occluded<HAS_NODES>(cx, sr, point ? 1.0f : RT_INF);
}
// Blinn-Phong's two normalisations and its power per light (scene.py:170-181) on the device: MUFU.RSQ instead of IEEE sqrt + divide
// (2 ulp), and x ** y as exp2(y * log2(x)) with the MUFU approximations (relative error about y * 2e-7: 1e-5 at hardness 64, two
// orders below 1/255) with Python's corner cases kept exact (x ** 0 == 1, also for x == 0; 0 ** y == 0).  They feed only the
// colour, never a ray: k_shade executed 622 instructions per ray, a third of them in the correctly rounded powf / sqrt / divide
// sequences of three lights.  The host twin (tests/hostsim) keeps the exact forms.
#if defined(__CUDA_ARCH__)
__device__ __forceinline__ float3 shade_normalize(float3 a) { return a * rsqrtf(dot(a, a)); }
__device__ __forceinline__ float shade_pow(float x, float y) { return y == 0.0f ? 1.0f : (x > 0.0f ? __powf(x, y) : 0.0f); }
#else
inline float3 shade_normalize(float3 a) { return normalize(a); }
inline float shade_pow(float x, float y) { return powf(x, y); }
#endif

// _compute_regular_lighting (scene.py:140-187): Blinn-Phong + ambient; lit(li) says whether light li reaches the
// hit point (the shadow-ray results: traced in place by trace_level, read back from k_shadow's flags by k_shade)
template <class Lit>
RT_HD float3 shade_local(TraceCtx& cx, const Ray& ray, const Hit& hit, Lit&& lit) {
    const SceneView& S = cx.S;
    const DevMaterial* mat = S.materials + hit.mat;
    float4 md = ld4(&mat->diffuse), ms = ld4(&mat->specular);
    const DevObject* g = S.objects + hit.geom;
    float3 diffuse;
    if (g->kind == RT_KIND_BOX) diffuse = box_diffuse(S, g, hit.p, cx.time, cx.tally);
    else if (g->kind == RT_KIND_PLANE) diffuse = plane_diffuse(S, g, hit.p, cx.time, cx.tally);
    else diffuse = f3(md);
    cx.tally.shaded++;
    float3 colour = f3(0, 0, 0);
    for (int li = 0; li < S.n_lights; li++) {
        if (!lit(li)) continue;
        float4 lv = light_vector(S, li);
        const bool point = as_int(lv.w) == RT_LIGHT_POINT;
        float3 light_dir = point ? shade_normalize(f3(lv) - hit.p) : f3(light_dir_norm(S, li));
        float3 lambert = diffuse * fmaxf(0.0f, dot(hit.n, light_dir));
        float3 half_vect = shade_normalize(light_dir - ray.d);
        float3 specular = f3(ms) * shade_pow(fmaxf(0.0f, dot(hit.n, half_vect)), md.w);   // 0**0 == 1 both sides
        colour = colour + f3(light_colour_power(S, li)) * (lambert + specular);
    }
    return colour + S.ambient * diffuse;
}

// The clamped-affine fold of the per-level clamp (scene.py:104,108,113-116).  A path carries
// F(c) = clamp(a + w*c, lo, hi): the map from "colour returned by the next cast_ray" to the sample
// value.  Composing with one more level g(c) = clamp(A + W*c, 0, 1) stays in the same family.
struct PathFold { float3 a; float w; float3 lo, hi; };
RT_HD PathFold fold_identity() { PathFold f; f.a = f3(0, 0, 0); f.w = 1.0f; f.lo = f3(0, 0, 0); f.hi = f3(1, 1, 1); return f; }
RT_HD float3 fold_eval(const PathFold& f, float3 c) { return clamp3(f.a + c * f.w, f.lo, f.hi); }
RT_HD PathFold fold_compose(const PathFold& f, float3 A, float W) {
    PathFold g;
    float3 e0 = f.a, e1 = f.a + f3(f.w, f.w, f.w);                     // F's pre-clamp value at g = 0 and g = 1
    g.lo = clamp3(f3(fminf(e0.x, e1.x), fminf(e0.y, e1.y), fminf(e0.z, e1.z)), f.lo, f.hi);
    g.hi = clamp3(f3(fmaxf(e0.x, e1.x), fmaxf(e0.y, e1.y), fmaxf(e0.z, e1.z)), f.lo, f.hi);
    g.a = f.a + A * f.w;
    g.w = f.w * W;
    return g;
}

// One level of cast_ray (scene.py:81-116) for one path.  Returns true when the path continues with
// `next` (reflection / refraction ray); otherwise `sample` holds the finished sample value.
struct PathState { Ray ray; PathFold fold; int depth_left; int in_shape; };

// Everything of cast_ray after the closest hit is known (scene.py:96-116): material branch, local lighting, fold,
// next ray.  `hit` may get its normal flipped (Q19).
template <class Lit>
RT_HD bool shade_level(TraceCtx& cx, PathState& ps, Hit& hit, Lit&& lit, float3& sample) {
    if (hit.mat < 0) { cx.tally.err |= 1u << RT_ERR_NO_MATERIAL; sample = fold_eval(ps.fold, f3(0, 0, 0)); return false; }
    const DevMaterial* mat = cx.S.materials + hit.mat;
    const int type = mat->type;
    // secondary ray first (scene.py:99-101,191-206), then ONE local-lighting evaluation shared by the three material branches
    float3 rd = f3(0, 0, 0), ro = f3(0, 0, 0);
    bool dead = false;                                                 // total internal reflection
    if (type == RT_MAT_MIRROR) {                                       // scene.py:97-104
        rd = reflect3(ps.ray.d, hit.n);
        ro = hit.p + rd * 0.01f;
    } else if (type == RT_MAT_REFRACTIVE) {                            // scene.py:105-108,189-209
        float eta;
        if (ps.in_shape) { eta = mat->refr_index; hit.n = -hit.n; }    // normal flipped BEFORE local lighting (Q19)
        else eta = mat->inv_refr;
        rd = refract3(ps.ray.d, hit.n, eta, dead);
        if (!dead && rd.x == 0.0f && rd.y == 0.0f && rd.z == 0.0f) dead = true;   // `refract_dir == vec3(0)`
        ro = hit.p + rd * 0.0001f;
    }
    const float3 local = shade_local(cx, ps.ray, hit, lit);
    if (type != RT_MAT_MIRROR && type != RT_MAT_REFRACTIVE) {          // scene.py:109-116
        sample = fold_eval(ps.fold, clamp01(local));
        return false;
    }
    ps.fold = fold_compose(ps.fold, local * ld4(&mat->specular).w, mat->one_minus_tint);
    if (dead || ps.depth_left - 1 <= 0) { sample = fold_eval(ps.fold, f3(0, 0, 0)); return false; }
    if (type == RT_MAT_MIRROR) { cx.tally.reflect++; ps.in_shape = 0; }
    else { cx.tally.refract++; ps.in_shape = !ps.in_shape; }
    ps.ray.o = ro; ps.ray.d = rd; ps.depth_left -= 1;
    return true;
}

// One whole level in one thread (unit-test kernels and tests/hostsim; the render kernels run the three stages as
// separate launches: k_hit, k_shadow, k_shade)
template <int HAS_NODES>
RT_HD bool trace_level(TraceCtx& cx, PathState& ps, float3& sample) {
    Hit hit;
    if (ps.depth_left <= 0 || !closest_hit<HAS_NODES>(cx, ps.ray, hit)) { sample = fold_eval(ps.fold, f3(0, 0, 0)); return false; }
    const float3 p = hit.p;
    return shade_level(cx, ps, hit, [&](int li) { return !light_occluded<HAS_NODES>(cx, p, li); }, sample);
}

// ------------------------------------------------------------------------------------------------
// Sample generation (scene.py:47-69): primary ray of (col, row, dof d, aa s)
// direction of the primary rays of (col, row, dof sample d): the AA point and the jitter move only the origin (scene.py:58-65)
RT_HD float3 primary_direction(const SceneView& S, int col, int row, int d) {
    float x = RT_LDG(S.col_x + col), y = RT_LDG(S.row_y + row);
    float3 base_dir = normalize((S.cam_u * x + S.cam_v * y) - S.cam_w);          // d = 1: d*w == w
    float3 focal_point = S.cam_pos + base_dir * S.focal_length;
    return normalize(focal_point - f3(ld4(S.dof_origin + d)));
}
RT_HD Ray generate_ray(const SceneView& S, int col, int row, int d, int s) {
    Ray r;
    r.d = primary_direction(S, col, row, d);
    r.o = f3(ld4(S.ray_origin + d * S.S + s));
    if (S.jitter) {
        uint64_t idx = (((uint64_t)col * (uint64_t)S.height + (uint64_t)row) * (uint64_t)S.D + (uint64_t)d) * (uint64_t)S.S + (uint64_t)s;
        uint32_t c0 = (uint32_t)idx, c1 = (uint32_t)(idx >> 32), c2 = 0, c3 = 0;
        philox4x32_10(c0, c1, c2, c3, S.seed_lo, S.seed_hi);
        float3 rnd = f3((float)c0 * 2.3283064365386963e-10f, (float)c1 * 2.3283064365386963e-10f, (float)c2 * 2.3283064365386963e-10f);
        r.o = r.o + normalize(rnd) * S.noise_scale;
    }
    return r;
}

// ------------------------------------------------------------------------------------------------
// Which pixels a launch covers (see RtRenderJob in include/b200rt.h)
// Division of a 32-bit unsigned by a launch-invariant divisor without the divide sequence (Granlund-Montgomery):
// q = (t + ((n - t) >> s1)) >> s2 with t = mulhi(n, mul).  The index arithmetic of one ray (pixel, tile, sample slots) is six
// divisions; as plain 64-bit '/' and '%' they were a fifth of k_hit's instructions on a scene with nothing to intersect.
struct FastDiv { unsigned mul, s1, s2, d; };
RT_HD unsigned umulhi32(unsigned a, unsigned b) {
#if defined(__CUDA_ARCH__)
    return __umulhi(a, b);
#else
    return (unsigned)(((unsigned long long)a * b) >> 32);
#endif
}
RT_HD unsigned fdiv(unsigned n, const FastDiv& f) { const unsigned t = umulhi32(n, f.mul); return (t + ((n - t) >> f.s1)) >> f.s2; }
inline FastDiv make_fastdiv(unsigned d) {
    FastDiv f; f.d = d < 1 ? 1 : d;
    unsigned l = 0; while ((1ull << l) < f.d) l++;                          // ceil(log2 d)
    f.mul = (unsigned)(((1ull << 32) * ((1ull << l) - f.d)) / f.d + 1);
    f.s1 = l < 1 ? l : 1; f.s2 = l < 1 ? 0 : l - 1;
    return f;
}

struct PixelMap {
    int col0, ncols, row0, nrows;
    int tile_w, tile_h, rank, world;
    int n_tile_rows;            // tiles per tile-column
    int tile_px;                // tile_w * tile_h
    long long n_local_pixels;   // n_local_tiles * tile_px (partial tiles padded); < 2^31
    FastDiv fd_tile_px, fd_tile_rows, fd_tile_w;      // filled by make_pixel_map
    FastDiv fd_N, fd_T, fd_S;                         // samples per pixel, motion times, AA samples: filled by the renderer
};

// local pixel slot lp (compact tile order of this rank) -> (col, row); false for the padding of partial tiles
RT_HD bool pixel_of(const PixelMap& pm, unsigned lp, int& col, int& row) {
    const unsigned lt = fdiv(lp, pm.fd_tile_px);
    const unsigned q = lp - lt * (unsigned)pm.tile_px;
    const unsigned gt = lt * (unsigned)pm.world + (unsigned)pm.rank;
    const unsigned tcol = fdiv(gt, pm.fd_tile_rows), trow = gt - tcol * (unsigned)pm.n_tile_rows;
    const unsigned qr = fdiv(q, pm.fd_tile_w), qc = q - qr * (unsigned)pm.tile_w;
    col = pm.col0 + (int)(tcol * (unsigned)pm.tile_w + qc);
    row = pm.row0 + (int)(trow * (unsigned)pm.tile_h + qr);
    return col < pm.col0 + pm.ncols && row < pm.row0 + pm.nrows;
}

