// rt_kernels.cuh - the wavefront kernels of the render path (one thread per (pixel, sample) ray).
//
// One cast_ray level (scene.py:81-116) = three launches over the same rays, each a SMALL kernel whose code stays
// in the SM's instruction caches (a single fused kernel measured 94 % of the GPC instruction-fetch bandwidth and
// a third of its issue slots stalled on "no instruction"):
//   k_hit     : raygen (level 0) or queue load, closest hit over the scene -> hit record planes
//   k_shadow  : any-hit shadow rays -> one occlusion bit (level 0) / byte (queued levels) per (light, ray)
//   k_shade   : Blinn-Phong with those flags, per-level clamp fold; finished samples go to the sample buffer,
//               reflection/refraction rays are compacted (warp ballot + prefix popcount + one atomic per warp)
//               into the next level's ray queue (replaces the recursion of scene.py:102,203,208)
//   k_resolve : per-pixel sum in the reference's loop order and division (scene.py:69,73)
//   k_assemble: scatter gathered tile buffers into the reference layout (glue.py:19-22)
//   k_finalize: rot90 + *255 truncation (main.py:31-33)
#pragma once
#include "rt_core.cuh"

// resident blocks per SM each kernel's register allocation is held to (tuned on B200)
#ifndef RT_MIN_BLOCKS_HIT
#define RT_MIN_BLOCKS_HIT 7
#endif
#ifndef RT_MIN_BLOCKS_SHADOW
#define RT_MIN_BLOCKS_SHADOW 8
#endif
// scenes with CSG / transform nodes (HAS_NODES): the node code needs far more registers than the plain walkers
#ifndef RT_MIN_BLOCKS_HIT_CSG
#define RT_MIN_BLOCKS_HIT_CSG 4
#endif
#ifndef RT_MIN_BLOCKS_SHADOW_CSG
#define RT_MIN_BLOCKS_SHADOW_CSG 4
#endif
#ifndef RT_MIN_BLOCKS_SHADE
#define RT_MIN_BLOCKS_SHADE 8
#endif

// The per-ray wavefront records (hit planes, occlusion flags, samples, queue entries) are written once and read once, tens of
// GB per frame, next to a ~100 MB BVH that every ray re-reads: they are accessed with the streaming (evict-first) cache hints so
// they do not push BVH lines out of L1/L2 (measured 0.5 ms of 124 per frame at 4K x 64; -DRT_STREAM_RECORDS=0 for plain accesses).
#ifndef RT_STREAM_RECORDS
#define RT_STREAM_RECORDS 1
#endif
#if RT_STREAM_RECORDS
#define RT_ST_REC(ptr, val) __stcs((ptr), (val))
#define RT_LD_REC(ptr) __ldcs(ptr)
#else
#define RT_ST_REC(ptr, val) (*(ptr) = (val))
#define RT_LD_REC(ptr) (*(ptr))
#endif

// Compacted ray queue, SoA of float4 planes (coalesced 128-bit accesses):
//   plane 0: origin.xyz, time           plane 1: dir.xyz, bits(sample slot)
//   plane 2: fold.a.xyz, fold.w         plane 3: fold.lo.xyz, bits(depth_left | in_shape << 8)
//   plane 4: fold.hi.xyz, -
struct RayQueue { float4* planes; long long cap; };
#define RT_QUEUE_PLANES 5

__device__ __forceinline__ void queue_store(const RayQueue& q, long long i, const PathState& ps, float time, unsigned slot) {
    q.planes[0 * q.cap + i] = f4(ps.ray.o.x, ps.ray.o.y, ps.ray.o.z, time);
    q.planes[1 * q.cap + i] = f4(ps.ray.d.x, ps.ray.d.y, ps.ray.d.z, as_float((int)slot));
    q.planes[2 * q.cap + i] = f4(ps.fold.a.x, ps.fold.a.y, ps.fold.a.z, ps.fold.w);
    q.planes[3 * q.cap + i] = f4(ps.fold.lo.x, ps.fold.lo.y, ps.fold.lo.z, as_float(ps.depth_left | (ps.in_shape << 8)));
    q.planes[4 * q.cap + i] = f4(ps.fold.hi.x, ps.fold.hi.y, ps.fold.hi.z, 0.0f);
}
__device__ __forceinline__ void queue_load(const RayQueue& q, long long i, PathState& ps, float& time, unsigned& slot) {
    float4 p0 = q.planes[0 * q.cap + i], p1 = q.planes[1 * q.cap + i], p2 = q.planes[2 * q.cap + i];
    float4 p3 = q.planes[3 * q.cap + i], p4 = q.planes[4 * q.cap + i];
    ps.ray.o = f3(p0); time = p0.w;
    ps.ray.d = f3(p1); slot = (unsigned)as_int(p1.w);
    ps.fold.a = f3(p2); ps.fold.w = p2.w;
    ps.fold.lo = f3(p3); int fl = as_int(p3.w); ps.depth_left = fl & 0xff; ps.in_shape = (fl >> 8) & 1;
    ps.fold.hi = f3(p4);
}

// Warp-aggregated append: ballot -> exclusive prefix by popcount -> one atomicAdd per warp.
// Must be reached by all 32 lanes of the warp.
__device__ __forceinline__ long long queue_reserve(unsigned long long* count, bool want) {
    unsigned mask = __ballot_sync(0xffffffffu, want);
    if (mask == 0) return -1;
    unsigned lane = threadIdx.x & 31;
    int leader = __ffs(mask) - 1;
    unsigned long long base = 0;
    if ((int)lane == leader) base = atomicAdd(count, (unsigned long long)__popc(mask));
    base = __shfl_sync(0xffffffffu, base, leader);
    return want ? (long long)(base + __popc(mask & ((1u << lane) - 1))) : -1;
}

__device__ __forceinline__ void tally_flush(const Tally& t, unsigned long long primary, unsigned long long queued, RtCounters* c) {
    unsigned v[9] = {t.shadow, t.reflect, t.refract, t.nodes, t.tris, t.prims, t.shaded, t.smooth, t.err};
    unsigned long long w[9];
#pragma unroll
    for (int k = 0; k < 9; k++) {
        unsigned long long x = v[k];
        if (k == 8) {   // error bits: OR
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) x |= __shfl_xor_sync(0xffffffffu, x, o);
        } else {
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) x += __shfl_xor_sync(0xffffffffu, x, o);
        }
        w[k] = x;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) { primary += __shfl_xor_sync(0xffffffffu, primary, o); queued += __shfl_xor_sync(0xffffffffu, queued, o); }
    if ((threadIdx.x & 31) == 0) {
        if (primary) atomicAdd((unsigned long long*)&c->rays_primary, primary);
        if (w[0]) atomicAdd((unsigned long long*)&c->rays_shadow, w[0]);
        if (w[1]) atomicAdd((unsigned long long*)&c->rays_reflect, w[1]);
        if (w[2]) atomicAdd((unsigned long long*)&c->rays_refract, w[2]);
        if (w[3]) atomicAdd((unsigned long long*)&c->bvh_nodes, w[3]);
        if (w[4]) atomicAdd((unsigned long long*)&c->tri_tests, w[4]);
        if (w[5]) atomicAdd((unsigned long long*)&c->prim_tests, w[5]);
        if (w[6]) atomicAdd((unsigned long long*)&c->shaded_hits, w[6]);
        if (w[7]) atomicAdd((unsigned long long*)&c->mesh_hits_smooth, w[7]);
        if (queued) atomicAdd((unsigned long long*)&c->queued_rays, queued);
        if (w[8]) atomicOr((unsigned long long*)&c->error_flags, w[8]);
    }
}

__device__ __forceinline__ Tally tally_zero() { Tally t; t.shadow = t.reflect = t.refract = t.nodes = t.tris = t.prims = t.shaded = t.smooth = t.err = 0; return t; }

// How a kernel receives the scene: SceneView BY VALUE in the parameter bank (every field a constant-bank operand for the code inlined
// into the kernel) and, for the general kernels, also a pointer to ONE copy in global memory for their out-of-line node functions
// (csg_inside, csg_shadow, ...), which take the scene by reference (TraceCtx::global()): a reference to the by-value parameter that
// reaches a real call made every THREAD copy the whole SceneView into its local-memory stack frame at kernel entry and read it back
// from there (ptxas: 2.0 - 3.2 KB frames, 512 threads per SM).
template <bool GLOBAL> struct ViewArg;
template <> struct ViewArg<false> { SceneView v; const SceneView* p; __device__ __forceinline__ const SceneView& get() const { return v; } };
template <> struct ViewArg<true> { SceneView v; const SceneView* __restrict__ p; __device__ __forceinline__ const SceneView& get() const { return v; } };

extern __shared__ int rt_smem_stack[];   // RT_STACK_ROWS * blockDim.x ints when the scene has a mesh or an object BVH
// the launches allocate the shared stack area only for scenes with something to walk (b200rt.cu: has_mesh || has_tlas)
__device__ __forceinline__ bool rt_smem_stack_present(const SceneView& S) { return S.tlas_nodes != nullptr || S.n_sep_mesh > 0; }

// Hit records of one level, SoA float4 planes of `cap` entries, indexed like the level's rays: 32 bytes per ray
//   plane 0: hit position xyz, w = ray time (NaN: the ray hit nothing / ended)
//   plane 1: shading normal xyz, w = bits((material index + 2) | leaf object index << RT_HIT_MAT_BITS)
// The ray direction k_shade needs is NOT stored (it was a third plane: 17 GB of DRAM traffic per 4K x 64 frame): a queued ray's
// direction is in its queue entry, a primary ray's is recomputed from its pixel (primary_direction: two table look-ups and two
// normalisations; the jitter only moves the origin).
struct HitPlanes { float4* planes; long long cap; };
#define RT_HIT_PLANES 2
#define RT_HIT_MAT_BITS 12            // materials < 4094 (the two specials -1 / -2 included), objects < 2^20: checked by rt_scene_create
__device__ __forceinline__ float hit_pack(int mat, int geom) { return as_float((mat + 2) | (geom << RT_HIT_MAT_BITS)); }
__device__ __forceinline__ void hit_unpack(float w, int& mat, int& geom) {
    const unsigned u = (unsigned)as_int(w);
    mat = (int)(u & ((1u << RT_HIT_MAT_BITS) - 1u)) - 2; geom = (int)(u >> RT_HIT_MAT_BITS);
}
// ray i of a level-0 batch -> pixel and sample slot (loop order of scene.py:57-67: slot = (d*S + s)*T + t)
struct PrimaryIndex { int col, row; unsigned d, s, t; bool live; };
__device__ __forceinline__ PrimaryIndex primary_index(const SceneView& S, const PixelMap& pm, long long pixel_begin, unsigned iu, unsigned N) {
    PrimaryIndex pi;
    const unsigned px = fdiv(iu, pm.fd_N), slot = iu - px * N;
    pi.live = pixel_of(pm, (unsigned)pixel_begin + px, pi.col, pi.row) && S.max_depth > 0;
    const unsigned ds = fdiv(slot, pm.fd_T);
    pi.t = slot - ds * (unsigned)S.T;
    pi.d = fdiv(ds, pm.fd_S); pi.s = ds - pi.d * (unsigned)S.S;
    return pi;
}

// ---- stage 1: closest hit ------------------------------------------------------------------------------
// PRIMARY: ray i of the batch is generated here: pixel = pixel_begin + i / N, slot = i % N = (d*S + s)*T + t (loop
// order of scene.py:57-67).  Otherwise ray i is entry i of the compacted queue `qin` (count on the device).
template <int HAS_NODES, bool PRIMARY>
__global__ void __launch_bounds__(RT_BLOCK, HAS_NODES ? RT_MIN_BLOCKS_HIT_CSG : RT_MIN_BLOCKS_HIT) k_hit(const ViewArg<(HAS_NODES != 0)> SA, PixelMap pm, long long pixel_begin, long long n_rays,
                                                                     RayQueue qin, const unsigned long long* qin_count,
                                                                     HitPlanes H, RtCounters* counters) {
    const SceneView& S = SA.get();
    Tally tally = tally_zero();
    int spill[RT_STACK_SPILL];       // overflow of the shared-memory traversal stack (thread-local memory)
    unsigned long long n_primary = 0;
    const int N = S.D * S.S * S.T;
    (void)N;
    const long long n = PRIMARY ? n_rays : (long long)*qin_count;
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        Ray ray; float time; bool live = true;
        if (PRIMARY) {
            const PrimaryIndex pi = primary_index(S, pm, pixel_begin, (unsigned)i, (unsigned)N);      // a batch holds < 2^30 rays
            live = pi.live;
            if (live) { ray = generate_ray(S, pi.col, pi.row, (int)pi.d, (int)pi.s); n_primary++; }
            time = RT_LDG(S.times + pi.t);
        } else {
            const float4 p0 = qin.planes[0 * qin.cap + i], p1 = qin.planes[1 * qin.cap + i];
            ray.o = f3(p0); ray.d = f3(p1); time = p0.w;
        }
        Hit hit;
        if (live) {
            TraceCtx cx{S, time, tally, rt_smem_stack + threadIdx.x, (int)blockDim.x, spill, SA.p};
            live = closest_hit<HAS_NODES>(cx, ray, hit);
        }
        if (live) {
            RT_ST_REC(&H.planes[0 * H.cap + i], f4(hit.p.x, hit.p.y, hit.p.z, time));
            RT_ST_REC(&H.planes[1 * H.cap + i], f4(hit.n.x, hit.n.y, hit.n.z, hit_pack(hit.mat, hit.geom)));
        } else {
            RT_ST_REC(&H.planes[0 * H.cap + i], f4(0.0f, 0.0f, 0.0f, as_float(0x7fc00000)));
        }
    }
    tally_flush(tally, n_primary, 0, counters);
}

// ---- stage 2: shadow rays (scene.py:148-167) -------------------------------------------------------------
// a warp works on ONE light and 32 neighbouring hit points at a time
// LIGHT_ROWS: gridDim.y = lights (small queues of the deeper levels: every (light, ray) pair gets its own thread);
// otherwise a thread walks the lights of its ray one after the other (big level-0 batches: the 16-B hit position is read
// once instead of once per light; measured 3 ms/frame faster than sweeping the rays once per light).
// MODE: how the (light, ray) pairs are laid over the threads
//   SHADOW_PER_RAY : a thread walks the lights of its ray one after the other (big level-0 batches: the 16-B hit position is read once
//                    instead of once per light; measured 3 ms/frame faster than sweeping the rays once per light)
//   SHADOW_ROWS    : gridDim.y = lights (small queues of the deeper levels: every (light, ray) pair gets its own thread)
// (A third mode walked the BVHs warp by warp - the 32 samples of a pixel as one packet with a warp-uniform stack - and was verified
// lane for lane against the per-ray walk: the same warp instructions per launch, 98 % of the GPC's instruction-fetch bandwidth, 63.3
// instead of 62.7 ms.  Removed; DESIGN.md keeps the measurement.)
enum { SHADOW_PER_RAY = 0, SHADOW_ROWS = 1 };
template <int HAS_NODES, int MODE>
__global__ void __launch_bounds__(RT_BLOCK, HAS_NODES ? RT_MIN_BLOCKS_SHADOW_CSG : RT_MIN_BLOCKS_SHADOW) k_shadow(const ViewArg<(HAS_NODES != 0)> SA, long long n_rays, const unsigned long long* n_dev,
                                                                           HitPlanes H, uint8_t* __restrict__ occ, RtCounters* counters) {
    const SceneView& S = SA.get();
    Tally tally = tally_zero();
    int spill[RT_STACK_SPILL];       // overflow of the shared-memory traversal stack (thread-local memory)
    const long long n = n_dev ? (long long)*n_dev : n_rays;
    const long long stride = (long long)gridDim.x * blockDim.x;
    {
        const int li0 = MODE == SHADOW_ROWS ? (int)blockIdx.y : 0, li1 = MODE == SHADOW_ROWS ? li0 + 1 : S.n_lights;
        for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
            const float4 h0 = RT_LD_REC(&H.planes[0 * H.cap + i]);
            if (h0.w != h0.w) continue;                                       // no hit: no shadow rays
            TraceCtx cx{S, h0.w, tally, rt_smem_stack + threadIdx.x, (int)blockDim.x, spill, SA.p};
            if (MODE == SHADOW_ROWS) {                                        // one byte per (light, ray)
                RT_ST_REC(&occ[(long long)li0 * H.cap + i], (uint8_t)(light_occluded<HAS_NODES>(cx, f3(h0), li0) ? 1 : 0));
            } else {                                                          // one BIT per (light, ray): byte plane li / 8; the lanes of a warp still work on ONE light at a time
                unsigned bits = 0;
                for (int li = li0; li < li1; li++) {
                    bits |= (light_occluded<HAS_NODES>(cx, f3(h0), li) ? 1u : 0u) << (li & 7);
                    if ((li & 7) == 7 || li == li1 - 1) { RT_ST_REC(&occ[(long long)(li >> 3) * H.cap + i], (uint8_t)bits); bits = 0; }
                }
            }
        }
    }
    tally_flush(tally, 0, 0, counters);
}

// ---- stage 3: shading, fold, next ray ----------------------------------------------------------------------
// PRIMARY (level 0): the occlusion flags are one BIT per (light, ray) (k_shadow SHADOW_PER_RAY), the ray direction is
// recomputed from the pixel; queued levels: one byte per (light, ray) (SHADOW_ROWS), direction from the queue entry.
template <bool PRIMARY>
__global__ void __launch_bounds__(RT_BLOCK, RT_MIN_BLOCKS_SHADE) k_shade(SceneView S, PixelMap pm, long long pixel_begin, long long n_rays, RayQueue qin, const unsigned long long* qin_count,
                                                                         HitPlanes H, const uint8_t* __restrict__ occ, float4* __restrict__ samples,
                                                                         RayQueue qout, unsigned long long* qout_count, RtCounters* counters) {
    Tally tally = tally_zero();
    unsigned long long n_queued = 0;
    const long long n = PRIMARY ? n_rays : (long long)*qin_count;
    const long long stride = (long long)gridDim.x * blockDim.x;
    const unsigned N = (unsigned)(S.D * S.S * S.T);
    for (long long base = (long long)blockIdx.x * blockDim.x; base < n; base += stride) {
        const long long i = base + threadIdx.x;
        bool cont = false;
        PathState ps; float time = 0.0f; unsigned slot = (unsigned)i;
        if (i < n) {
            // every load of this ray issued up front (the kernel is latency-bound: hit planes + the flag bytes)
            const float4 h0 = RT_LD_REC(&H.planes[0 * H.cap + i]), h1 = RT_LD_REC(&H.planes[1 * H.cap + i]);
            unsigned lit_mask = 0;
            const int n_pre = S.n_lights < 32 ? S.n_lights : 32;
            if (PRIMARY) {
                for (int b = 0; b * 8 < n_pre; b++) lit_mask |= (unsigned)RT_LD_REC(&occ[(long long)b * H.cap + i]) << (8 * b);
                lit_mask = ~lit_mask;
                ps.fold = fold_identity(); ps.depth_left = S.max_depth; ps.in_shape = 0; ps.ray.o = f3(0, 0, 0);
            } else {
                for (int li = 0; li < n_pre; li++) lit_mask |= (RT_LD_REC(&occ[(long long)li * H.cap + i]) == 0 ? 1u : 0u) << li;
                queue_load(qin, i, ps, time, slot);
            }
            float3 sample;
            if (h0.w != h0.w) sample = fold_eval(ps.fold, f3(0, 0, 0));          // miss: cast_ray returns black (scene.py:91-92)
            else {
                Hit hit; hit.p = f3(h0); hit.n = f3(h1); hit_unpack(h1.w, hit.mat, hit.geom); hit.t = 0.0f;
                time = h0.w;
                if (PRIMARY) {
                    const PrimaryIndex pi = primary_index(S, pm, pixel_begin, (unsigned)i, N);
                    ps.ray.d = primary_direction(S, pi.col, pi.row, (int)pi.d);
                }
                TraceCtx cx{S, time, tally, nullptr, 0, nullptr};
                cont = shade_level(cx, ps, hit, [&](int li) {
                    if (li < 32) return ((lit_mask >> li) & 1u) != 0;
                    return PRIMARY ? ((occ[(long long)(li >> 3) * H.cap + i] >> (li & 7)) & 1) == 0 : occ[(long long)li * H.cap + i] == 0; }, sample);
            }
            if (!cont) RT_ST_REC(&samples[slot], f4(sample.x, sample.y, sample.z, 0.0f));
        }
        long long qi = queue_reserve(qout_count, cont);
        if (cont) { queue_store(qout, qi, ps, time, slot); n_queued++; }
    }
    tally_flush(tally, 0, n_queued, counters);
}

// ---- the deep levels in one launch ----------------------------------------------------------------------------
// After a few levels the queues hold a small fraction of a batch, and every further level costs three launches whose
// time is the slowest warp's, not the ray count's.  From level RT_TAIL_LEVEL on, each queued path is followed to its end
// by ONE thread (closest hit, shadow rays and shading in place, like the reference's recursion): same functions, same
// results, one launch instead of 3 x (max_depth - RT_TAIL_LEVEL).
#ifndef RT_MIN_BLOCKS_TAIL
#define RT_MIN_BLOCKS_TAIL 4
#endif
// Paths end at different depths, so a thread that walked "its" path to the end in an inner loop left its lane idle until the
// longest of the warp's 32 paths was done (ncu: 9 of 32 lanes active).  Here every lane runs ONE level of whatever path it holds
// per iteration, and a lane whose path has ended fetches the next queued path right away (warp-aggregated atomic on `next`, a
// zeroed counter): the lanes of a warp stay together at the granularity of a level.
template <int HAS_NODES>
__global__ void __launch_bounds__(RT_BLOCK, HAS_NODES ? RT_MIN_BLOCKS_HIT_CSG : RT_MIN_BLOCKS_TAIL) k_tail(const ViewArg<(HAS_NODES != 0)> SA, RayQueue qin, const unsigned long long* qin_count,
                                                                                          unsigned long long* next, float4* __restrict__ samples, RtCounters* counters) {
    const SceneView& S = SA.get();
    Tally tally = tally_zero();
    int spill[RT_STACK_SPILL];
    unsigned long long n_queued = 0;
    const long long n = (long long)*qin_count;
    const unsigned lane = threadIdx.x & 31u;
    PathState ps; float time = 0.0f; unsigned slot = 0;
    bool have = false, exhausted = false;
    for (;;) {
        const unsigned want = __ballot_sync(0xffffffffu, !have && !exhausted);
        if (want) {
            const int leader = __ffs(want) - 1;
            unsigned long long base = 0;
            if ((int)lane == leader) base = atomicAdd(next, (unsigned long long)__popc(want));
            base = __shfl_sync(0xffffffffu, base, leader);
            if (!have && !exhausted) {
                const long long i = (long long)(base + __popc(want & ((1u << lane) - 1u)));
                if (i < n) { queue_load(qin, i, ps, time, slot); have = true; }
                else exhausted = true;
            }
        }
        if (!__any_sync(0xffffffffu, have)) break;
        if (have) {
            TraceCtx cx{S, time, tally, rt_smem_stack + threadIdx.x, (int)blockDim.x, spill, SA.p};
            float3 sample;
            if (trace_level<HAS_NODES>(cx, ps, sample)) n_queued++;      // counted like the rays the wavefront levels would have queued
            else { samples[slot] = f4(sample.x, sample.y, sample.z, 0.0f); have = false; }
        }
    }
    tally_flush(tally, 0, n_queued, counters);
}

// ---- per-pixel average -------------------------------------------------------------------------------
// colour accumulates in float32 in loop order (scene.py:69), then one float32 division (scene.py:73)
__global__ void k_resolve(PixelMap pm, long long pixel_begin, int n_pixels, int N, const float4* __restrict__ samples,
                          float* __restrict__ out, int direct) {
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n_pixels) return;
    const unsigned lp = (unsigned)(pixel_begin + p);
    int col, row;
    const bool valid = pixel_of(pm, lp, col, row);
    float3 c = f3(0, 0, 0);
    if (valid) {
        const float4* sp = samples + (long long)p * N;
        for (int k = 0; k < N; k++) { float4 v = RT_LD_REC(&sp[k]); c = c + f3(v); }
        const float n = (float)N;
        c = f3(c.x / n, c.y / n, c.z / n);
    }
    long long o;
    if (direct) { if (!valid) return; o = ((long long)(col - pm.col0) * pm.nrows + (row - pm.row0)) * 3; }
    else o = lp * 3;
    out[o] = c.x; out[o + 1] = c.y; out[o + 2] = c.z;
}

// ---- fan-in (glue.py:19-22) ----------------------------------------------------------------------------
#define RT_MAX_WORLD 64
struct RankOffsets { long long v[RT_MAX_WORLD]; };      // first float of every rank's compact buffer inside `gathered`
__global__ void k_assemble(PixelMap pm, const float* __restrict__ gathered, const RankOffsets rank_off,
                           float* __restrict__ out) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (long long)pm.ncols * pm.nrows) return;
    const int c = (int)(i / pm.nrows), r = (int)(i - (long long)c * pm.nrows);
    const int tcol = c / pm.tile_w, trow = r / pm.tile_h;
    const long long gt = (long long)tcol * pm.n_tile_rows + trow;
    const int rank = (int)(gt % pm.world);
    const long long lt = gt / pm.world;
    const int q = (r - trow * pm.tile_h) * pm.tile_w + (c - tcol * pm.tile_w);
    const float* src = gathered + rank_off.v[rank] + (lt * pm.tile_px + q) * 3;
    out[i * 3] = src[0]; out[i * 3 + 1] = src[1]; out[i * 3 + 2] = src[2];
}

// ---- main.py:31-33: rot90 + (image * 255).astype(uint8) (double product, truncation) --------------------
__global__ void k_finalize_u8(const float* __restrict__ img, int ncols, int nrows, uint8_t* __restrict__ out) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (long long)ncols * nrows) return;
    const int line = (int)(i / ncols), j = (int)(i - (long long)line * ncols);
    const float* src = img + ((long long)j * nrows + (nrows - 1 - line)) * 3;
#pragma unroll
    for (int k = 0; k < 3; k++) out[i * 3 + k] = (uint8_t)(int)((double)src[k] * 255.0);
}

// ---- unit-test aids: explicit rays through the same device functions ---------------------------------------
template <int HAS_NODES>
__global__ void __launch_bounds__(RT_BLOCK) k_trace(const ViewArg<(HAS_NODES != 0)> SA, long long n, const float* __restrict__ org, const float* __restrict__ dir,
                                                     const float* __restrict__ times, float* rgb, float* hit_t, float* hit_n, float* hit_p,
                                                     int* hit_geom, int* hit_mat, RtCounters* counters) {
    const SceneView& S = SA.get();
    Tally tally = tally_zero();
    int spill[RT_STACK_SPILL];       // overflow of the shared-memory traversal stack (thread-local memory)
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) {
        PathState ps;
        ps.ray.o = f3(org[3 * i], org[3 * i + 1], org[3 * i + 2]);
        ps.ray.d = f3(dir[3 * i], dir[3 * i + 1], dir[3 * i + 2]);
        TraceCtx cx{S, times[i], tally, rt_smem_stack + threadIdx.x, (int)blockDim.x, spill, SA.p};
        Hit h;
        if (closest_hit<HAS_NODES>(cx, ps.ray, h)) {
            hit_t[i] = h.t; hit_geom[i] = h.geom; hit_mat[i] = h.mat;
            hit_n[3 * i] = h.n.x; hit_n[3 * i + 1] = h.n.y; hit_n[3 * i + 2] = h.n.z;
            hit_p[3 * i] = h.p.x; hit_p[3 * i + 1] = h.p.y; hit_p[3 * i + 2] = h.p.z;
        } else {
            hit_t[i] = RT_INF; hit_geom[i] = -1; hit_mat[i] = -1;
            for (int k = 0; k < 3; k++) { hit_n[3 * i + k] = 0.0f; hit_p[3 * i + k] = 0.0f; }
        }
        ps.fold = fold_identity(); ps.depth_left = S.max_depth; ps.in_shape = 0;
        float3 sample;
        while (trace_level<HAS_NODES>(cx, ps, sample)) {}
        rgb[3 * i] = sample.x; rgb[3 * i + 1] = sample.y; rgb[3 * i + 2] = sample.z;
    }
    tally_flush(tally, 0, 0, counters);
}

template <int HAS_NODES>
__global__ void __launch_bounds__(RT_BLOCK) k_occluded(const ViewArg<(HAS_NODES != 0)> SA, long long n, const float* __restrict__ org, const float* __restrict__ dir,
                                                        const float* __restrict__ times, const float* __restrict__ tmax, int* occ,
                                                        RtCounters* counters) {
    const SceneView& S = SA.get();
    Tally tally = tally_zero();
    int spill[RT_STACK_SPILL];       // overflow of the shared-memory traversal stack (thread-local memory)
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const bool valid = i < n;
    const long long j = valid ? i : 0;
    Ray r; r.o = f3(org[3 * j], org[3 * j + 1], org[3 * j + 2]); r.d = f3(dir[3 * j], dir[3 * j + 1], dir[3 * j + 2]);
    TraceCtx cx{S, times[j], tally, rt_smem_stack + threadIdx.x, (int)blockDim.x, spill, SA.p};
    bool o;
    o = valid && occluded<HAS_NODES>(cx, r, tmax[j]);
    if (valid) occ[i] = o ? 1 : 0;
    tally_flush(tally, 0, 0, counters);
}
