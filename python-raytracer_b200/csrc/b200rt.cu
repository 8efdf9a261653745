// b200rt.cu - C ABI of the render path (include/b200rt.h): scene upload, LBVH build, wavefront
// render orchestration on a CUDA stream.  Single translation unit: kernels live in the .cuh files.
#include <cuda_runtime.h>

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "rt_bvh.cuh"
#include "rt_ingest.cuh"
#include "rt_kernels.cuh"
#include "rt_obj.hpp"
#include "rt_peaks.cuh"
#include "rt_scene_build.hpp"

#ifndef RT_GRID_WAVES
#define RT_GRID_WAVES 48
#endif
#ifndef RT_GRID_WAVES_DEEP
#define RT_GRID_WAVES_DEEP 8
#endif
// First level handled by k_tail (0 = never: every level as three launches).  The level is FIXED per kernel family, not chosen from
// the batch size: the same per-ray functions compiled into k_tail and into k_hit / k_shadow / k_shade round a few products
// differently (nvcc contracts across the stages k_tail inlines into one loop), so a path's last bits depend on where k_tail takes
// over - measured: up to 1.6e-4 on deep refraction chains of scenes/MirrorRefraction.json between take-over levels 1 and 3.  A
// take-over level that moved with the batch size made the image depend on the batch size, i.e. on the number of ranks a frame is
// split over (SURVEY 8e wants N-rank == 1-rank bit for bit).  Measured at 4K x 64 (ms/frame) for levels never / 1 / 2 / 3 / 4:
// 121.2 / 119.6 / 120.1 / 119.0 / 119.3.  Scenes with CSG / transform nodes: k_tail<true> is the whole node interpreter in one
// kernel (instruction-fetch bound, 4 blocks/SM); NovelScene1 at its shipped 2048x1024x32 took 159 ms with k_tail from level 2
// against 139 ms without.  Developer override for experiments: B200RT_TAIL_LEVEL (echoed by bench.py).
#ifndef RT_TAIL_LEVEL_PLAIN
#define RT_TAIL_LEVEL_PLAIN 3
#endif
#ifndef RT_TAIL_LEVEL_NODES
#define RT_TAIL_LEVEL_NODES 0
#endif
static int tail_level_for(bool has_nodes) {
    static const char* env = getenv("B200RT_TAIL_LEVEL");
    if (env) return atoi(env);
    return has_nodes ? RT_TAIL_LEVEL_NODES : RT_TAIL_LEVEL_PLAIN;
}

// Persistent grid-stride grids: blocks = SMs x resident blocks per SM of THAT kernel (its __launch_bounds__) x waves.  Measured at
// 4K x 64 spp (ms/frame): 1 wave 144.6, 2 139.0, 4 132.3, 8 128.3, 16 127.0, 32 126.1, 64 126.7, 256 134.4 (48 + 8 deep: 124.5) - blocks finish at
// very different times (rays per block are equal, node visits are not), so small blocks even out the end of a launch; the small
// queues of the deeper levels want fewer (18 ms at 8-16 waves, 19.4 at 64).  Developer overrides: B200RT_WAVES, B200RT_WAVES_DEEP.
static int grid_waves(bool deep) {
    static const char* env = getenv("B200RT_WAVES");
    static const char* env_deep = getenv("B200RT_WAVES_DEEP");
    const char* e = deep ? env_deep : env;
    const int w = e ? atoi(e) : (deep ? RT_GRID_WAVES_DEEP : RT_GRID_WAVES);
    return w > 0 ? w : 1;
}

// profiling classes (per-launch CUDA-event pairs when rt_scene_profile is on): level-0 k_hit / k_shadow / k_shade, k_resolve,
// then the queued levels: level-1 k_hit / k_shadow / k_shade, all launches of levels >= 2, k_tail
enum { PC_HIT0 = 0, PC_SHADOW0, PC_SHADE0, PC_RESOLVE, PC_HIT1, PC_SHADOW1, PC_SHADE1, PC_DEEP, PC_TAIL, RT_PROF_CLASSES };
#define RT_COUNTER_CLASSES 4     // tallies kept apart for level-0 k_hit / k_shadow / k_shade and for the deeper levels
static thread_local std::string g_last_error;

static int fail(int code, const std::string& msg) { g_last_error = msg; return code; }

#define CU(expr)                                                                                          \
    do {                                                                                                  \
        cudaError_t e_ = (expr);                                                                          \
        if (e_ != cudaSuccess)                                                                            \
            return fail(e_ == cudaErrorMemoryAllocation ? RT_ERR_NOMEM : RT_ERR_CUDA,                      \
                        std::string(#expr) + ": " + cudaGetErrorString(e_));                              \
    } while (0)

// cudaSetDevice(scene's device) for the duration of a call, the caller's current device restored on return
struct DeviceGuard {
    int prev = -1; bool switched = false;
    explicit DeviceGuard(int dev) { if (cudaGetDevice(&prev) == cudaSuccess && prev != dev) switched = cudaSetDevice(dev) == cudaSuccess; }
    ~DeviceGuard() { if (switched) cudaSetDevice(prev); }
};

struct MeshBuild {
    int n_verts = 0, n_faces = 0;
    float* d_verts = nullptr;       // n_verts*3
    int* d_faces = nullptr;         // n_faces*3
    float4* d_vnormals = nullptr;   // n_verts
    float4* d_tri = nullptr;        // 3*n_faces
    int4* d_face_vidx = nullptr;    // n_faces
    BvhNode4* d_nodes = nullptr;    // n_nodes4 traversal nodes
    int n_nodes4 = 0;
    int* d_order = nullptr;         // n_faces: slot -> original face
    float3 lo, hi;
    float pad = 0.0f;
    bool built = false;
};

// the per-batch wavefront buffers of a render
struct WorkSet {
    float4* d_samples = nullptr; long long samples_cap = 0;
    float4* d_queue[2] = {nullptr, nullptr}; long long queue_cap = 0;
    float4* d_hit = nullptr; long long hit_cap = 0;          // hit record planes of the level in flight
    uint8_t* d_occ = nullptr; long long occ_cap = 0;         // n_lights planes of occlusion flags
    unsigned long long* d_qcount = nullptr;                  // [max_depth + 2]
};

struct RtScene {
    SceneView view{};
    std::vector<void*> allocs;
    std::vector<MeshBuild> meshes;
    std::vector<DevMesh> h_meshes;
    std::vector<DevObject> h_objs;             // host copy of the object records (which mesh a top-level mesh object uses)
    DevMesh* d_meshes = nullptr;
    bool has_nodes = false, has_specular = false, has_mesh = false, has_tlas = false;
    int samples_per_pixel = 1;
    // work buffers of rt_render, grown on demand.  TWO sets: consecutive wavefront batches of a frame alternate between them and
    // between two streams (the caller's and `side`), so that one batch's tails and latency-bound kernels (the small launches of the
    // deeper levels, k_shade, the last blocks of every launch) run under the other batch's issue-bound kernels.
    WorkSet ws[2];
    cudaStream_t side = nullptr;               // created by the first render that has two batches
    cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
    cudaEvent_t ev_l0[2] = {nullptr, nullptr}, ev_deep[2] = {nullptr, nullptr};   // pipeline mode: level 0 of a set done / its deeper levels + resolve done
    bool serial_only = false;                  // rt_render_static: one set (the static array), one stream
    // rt_render_host with a PINNED destination: the part of the frame a batch completes is copied while the next batches render
    float* sink_host = nullptr; long long sink_done = 0;
    cudaStream_t copy_stream = nullptr; cudaEvent_t ev_copy = nullptr;
    RtCounters* d_counters = nullptr;
    SceneView* d_view = nullptr;               // the scene as the general kernels read it (ViewArg<true>): one copy in global memory
    SceneView h_view_dev{};                    // what d_view holds
    bool view_dev_valid = false;
    float* d_stage = nullptr; long long stage_cap = 0;   // rt_render_host staging
    uint8_t* d_stage_u8 = nullptr; long long stage_u8_cap = 0;
    int device = 0;
    int sm_count = 148;
    unsigned long long last_launches = 0;
    bool profiling = false, l2_window_tried = false;
    std::vector<std::pair<cudaEvent_t, cudaEvent_t>> ev[RT_PROF_CLASSES];     // recorded, not yet read
    std::vector<cudaEvent_t> ev_pool;                                          // events ready for re-use (created once)
    double prof_ms[RT_PROF_CLASSES] = {};
    unsigned long long prof_n[RT_PROF_CLASSES] = {};
};

struct ProfScope {   // event pair around one launch when profiling is on (events come from a pool: none is created per launch once warm)
    RtScene* sc; int cls; cudaStream_t st; cudaEvent_t a = nullptr, b = nullptr;
    static cudaEvent_t take(RtScene* s) {
        cudaEvent_t e = nullptr;
        if (!s->ev_pool.empty()) { e = s->ev_pool.back(); s->ev_pool.pop_back(); }
        else cudaEventCreate(&e);
        return e;
    }
    ProfScope(RtScene* s, int c, cudaStream_t t) : sc(s), cls(c), st(t) {
        if (sc->profiling) { a = take(sc); b = take(sc); cudaEventRecord(a, st); }
    }
    ~ProfScope() { if (a) { cudaEventRecord(b, st); sc->ev[cls].push_back({a, b}); } }
};

template <class T>
static int upload(RtScene* sc, const T* host, size_t n, T** out) {
    *out = nullptr;
    const size_t alloc_n = n == 0 ? 1 : n;
    void* p = nullptr;
    CU(cudaMalloc(&p, alloc_n * sizeof(T)));
    sc->allocs.push_back(p);
    if (host && n) CU(cudaMemcpy(p, host, n * sizeof(T), cudaMemcpyHostToDevice));
    else CU(cudaMemset(p, 0, alloc_n * sizeof(T)));       // the callers synchronise once before caller streams may touch the buffers
    *out = (T*)p;
    return RT_OK;
}
#define UP(call) do { int rc_ = (call); if (rc_ != RT_OK) return rc_; } while (0)

using rtb::h3;
using rtb::h4;

extern "C" int rt_abi_version(void) { return RT_ABI_VERSION; }
extern "C" const char* rt_last_error(void) { return g_last_error.c_str(); }

static int scene_create_impl(const RtSceneDesc* d, RtScene* sc) {
    CU(cudaGetDevice(&sc->device));
    CU(cudaDeviceGetAttribute(&sc->sm_count, cudaDevAttrMultiProcessorCount, sc->device));
    HostScene hs;
    std::string err;
    int rc = rtb::convert_scene(d, hs, err);
    if (rc != RT_OK) return fail(rc, err);
    sc->has_nodes = hs.has_nodes; sc->has_specular = hs.has_specular; sc->has_mesh = hs.has_mesh;
    sc->view = hs.view;
    SceneView& V = sc->view;
    DevObject* d_objs; UP(upload(sc, hs.objs.data(), hs.objs.size(), &d_objs));
    sc->h_objs = hs.objs;
    DevMaterial* d_mats; UP(upload(sc, hs.mats.data(), hs.mats.size(), &d_mats));
    DevLight* d_lights; UP(upload(sc, hs.lights.data(), hs.lights.size(), &d_lights));
    static_assert(sizeof(DevXform) == sizeof(RtXform), "xform layout");
    DevXform* d_xf; UP(upload(sc, (const DevXform*)d->xforms, (size_t)d->n_xforms, &d_xf));
    V.objects = d_objs; V.materials = d_mats; V.lights = d_lights; V.xforms = d_xf;
    // ---- object BVH over the bounded top-level objects (host-built in convert_scene)
    sc->has_tlas = !hs.tlas_nodes.empty();
    if (sc->has_tlas) { BvhNode4* d_tn; UP(upload(sc, hs.tlas_nodes.data(), hs.tlas_nodes.size(), &d_tn)); V.tlas_nodes = d_tn; }
    if (hs.has_nodes) {
        float4 *d_clo, *d_chi;
        UP(upload(sc, hs.child_lo.data(), hs.child_lo.size(), &d_clo)); UP(upload(sc, hs.child_hi.data(), hs.child_hi.size(), &d_chi));
        V.child_lo = d_clo; V.child_hi = d_chi;
    }
    int* d_lin; UP(upload(sc, hs.lin_objs.data(), hs.lin_objs.size(), &d_lin));
    V.lin_objs = d_lin; V.n_lin = (int)hs.lin_objs.size();
    int* d_sep; UP(upload(sc, hs.sep_mesh.data(), hs.sep_mesh.size(), &d_sep));
    V.sep_mesh = d_sep; V.n_sep_mesh = (int)hs.sep_mesh.size();
    // ---- textures
    std::vector<DevTexture> texs(d->n_textures);
    for (int i = 0; i < d->n_textures; i++) {
        uint8_t* px; UP(upload(sc, d->textures[i].rgb, (size_t)d->textures[i].width * d->textures[i].height * 3, &px));
        texs[i].rgb = px; texs[i].width = d->textures[i].width; texs[i].height = d->textures[i].height;
    }
    DevTexture* d_tex; UP(upload(sc, texs.data(), texs.size(), &d_tex));
    V.textures = d_tex;
    // ---- meshes (raw arrays now; LBVH in rt_scene_build_bvh)
    sc->meshes.resize(d->n_meshes);
    sc->h_meshes.resize(d->n_meshes);
    for (int i = 0; i < d->n_meshes; i++) {
        const RtMesh& m = d->meshes[i];
        MeshBuild& b = sc->meshes[i];
        b.n_verts = m.n_verts; b.n_faces = m.n_faces;
        for (size_t k = 0; k < (size_t)m.n_faces * 3; k++)
            if (m.faces[k] < 0 || m.faces[k] >= m.n_verts) return fail(RT_ERR_INVALID, "mesh face index out of range");
        UP(upload(sc, m.verts, (size_t)m.n_verts * 3, &b.d_verts));
        UP(upload(sc, m.faces, (size_t)m.n_faces * 3, &b.d_faces));
        if (m.vnormals && !m.flat_shaded) {
            std::vector<float4> vn(m.n_verts);
            for (int k = 0; k < m.n_verts; k++) vn[k] = h4(m.vnormals + 3 * (size_t)k);
            UP(upload(sc, vn.data(), vn.size(), &b.d_vnormals));
        }
        rtb::mesh_bounds(m, b.lo, b.hi, b.pad);
        rtb::mesh_header(m, sc->h_meshes[i]);
        sc->h_meshes[i].vnormals = b.d_vnormals;
    }
    UP(upload(sc, (const DevMesh*)nullptr, (size_t)d->n_meshes, &sc->d_meshes));
    V.meshes = sc->d_meshes;
    // ---- sampling tables
    float *d_colx, *d_rowy, *d_times; float4 *d_dof, *d_org;
    UP(upload(sc, hs.col_x.data(), hs.col_x.size(), &d_colx)); UP(upload(sc, hs.row_y.data(), hs.row_y.size(), &d_rowy));
    UP(upload(sc, hs.dof_origin.data(), hs.dof_origin.size(), &d_dof)); UP(upload(sc, hs.ray_origin.data(), hs.ray_origin.size(), &d_org));
    UP(upload(sc, hs.times.data(), hs.times.size(), &d_times));
    V.col_x = d_colx; V.row_y = d_rowy; V.dof_origin = d_dof; V.ray_origin = d_org; V.times = d_times;
    sc->samples_per_pixel = d->dof_samples * d->aa_samples * d->n_times;
    for (int k = 0; k < 2; k++) UP(upload(sc, (const unsigned long long*)nullptr, (size_t)d->max_depth + 2, &sc->ws[k].d_qcount));
    UP(upload(sc, (const RtCounters*)nullptr, RT_COUNTER_CLASSES, &sc->d_counters));
    UP(upload(sc, (const SceneView*)nullptr, 1, &sc->d_view));
    CU(cudaDeviceSynchronize());     // the memsets above ran on the legacy stream; later kernels run on caller streams
    return RT_OK;
}

static void drain_events(RtScene* sc);

extern "C" int rt_scene_destroy(RtScene* sc) {
    if (!sc) return RT_OK;
    for (void* p : sc->allocs) cudaFree(p);
    for (WorkSet& w : sc->ws) {
        if (w.d_samples) cudaFree(w.d_samples);
        for (int k = 0; k < 2; k++) if (w.d_queue[k]) cudaFree(w.d_queue[k]);
        if (w.d_hit) cudaFree(w.d_hit);
        if (w.d_occ) cudaFree(w.d_occ);
    }
    if (sc->ev_fork) cudaEventDestroy(sc->ev_fork);
    if (sc->ev_join) cudaEventDestroy(sc->ev_join);
    if (sc->ev_copy) cudaEventDestroy(sc->ev_copy);
    if (sc->copy_stream) cudaStreamDestroy(sc->copy_stream);
    for (int k = 0; k < 2; k++) { if (sc->ev_l0[k]) cudaEventDestroy(sc->ev_l0[k]); if (sc->ev_deep[k]) cudaEventDestroy(sc->ev_deep[k]); }
    if (sc->side) cudaStreamDestroy(sc->side);
    if (sc->d_stage) cudaFree(sc->d_stage);
    if (sc->d_stage_u8) cudaFree(sc->d_stage_u8);
    drain_events(sc);
    for (cudaEvent_t e : sc->ev_pool) cudaEventDestroy(e);
    delete sc;
    return RT_OK;
}

extern "C" int rt_scene_create(const RtSceneDesc* d, RtScene** out) {
    if (!d || !out) return fail(RT_ERR_INVALID, "null argument");
    if (d->abi_version != RT_ABI_VERSION) return fail(RT_ERR_INVALID, "RtSceneDesc.abi_version mismatch");
    RtScene* sc = new RtScene();
    int rc = scene_create_impl(d, sc);
    if (rc != RT_OK) { std::string keep = g_last_error; rt_scene_destroy(sc); g_last_error = keep; *out = nullptr; return rc; }
    *out = sc;
    return RT_OK;
}

// ---------------------------------------------------------------------------------------------------
// LBVH build
static int build_mesh_bvh(RtScene* sc, int mi, cudaStream_t st) {
    MeshBuild& b = sc->meshes[mi];
    DevMesh& g = sc->h_meshes[mi];
    const int n = b.n_faces;
    g.n_tris = n;
    if (n == 0) { b.built = true; return RT_OK; }
    UP(upload(sc, (const float4*)nullptr, (size_t)n * 3, &b.d_tri));
    UP(upload(sc, (const int4*)nullptr, (size_t)n, &b.d_face_vidx));
    UP(upload(sc, (const int*)nullptr, (size_t)n, &b.d_order));
    // temporaries
    uint64_t *keys[2]; int* vals[2]; unsigned *hist, *scratch; int2* children; int *parent_node, *parent_leaf; unsigned *arrive, *even, *remap;
    float4 *leaf_lo, *leaf_hi, *node_lo, *node_hi;
    const int n_tiles = (n + RS_TILE - 1) / RS_TILE;
    std::vector<void*> tmp;
    auto talloc = [&](void** p, size_t bytes) { cudaError_t e = cudaMalloc(p, std::max<size_t>(bytes, 16)); if (e == cudaSuccess) tmp.push_back(*p); return e; };
    auto tfree = [&]() { for (void* p : tmp) cudaFree(p); };
#define TA(ptr, bytes) do { cudaError_t e_ = talloc((void**)&(ptr), (bytes)); if (e_ != cudaSuccess) { tfree(); return fail(RT_ERR_NOMEM, cudaGetErrorString(e_)); } } while (0)
    // every CUDA call of the build is checked: the first failure frees the temporaries and is reported with its call site
#define TC(expr) do { cudaError_t e_ = (expr); if (e_ != cudaSuccess) { tfree(); return fail(RT_ERR_CUDA, std::string("LBVH build: " #expr ": ") + cudaGetErrorString(e_)); } } while (0)
    TA(keys[0], sizeof(uint64_t) * n); TA(keys[1], sizeof(uint64_t) * n);
    TA(vals[0], sizeof(int) * n); TA(vals[1], sizeof(int) * n);
    TA(hist, sizeof(unsigned) * 256 * n_tiles);
    TA(scratch, sizeof(unsigned) * scan_scratch_elems(std::max<long long>(256ll * n_tiles, n)));
    TA(children, sizeof(int2) * n); TA(parent_node, sizeof(int) * n); TA(parent_leaf, sizeof(int) * n);
    TA(arrive, sizeof(unsigned) * n); TA(even, sizeof(unsigned) * n); TA(remap, sizeof(unsigned) * n);
    TA(leaf_lo, sizeof(float4) * n); TA(leaf_hi, sizeof(float4) * n); TA(node_lo, sizeof(float4) * n); TA(node_hi, sizeof(float4) * n);
    TC(cudaDeviceSynchronize());                     // the zero-fills of upload() ran on the legacy stream
    float3 ext = b.hi - b.lo;
    float3 inv_ext = f3(ext.x > 0 ? 1.0f / ext.x : 0.0f, ext.y > 0 ? 1.0f / ext.y : 0.0f, ext.z > 0 ? 1.0f / ext.z : 0.0f);
    const int T = 256, G = (n + T - 1) / T;
    k_morton<<<G, T, 0, st>>>(b.d_verts, b.d_faces, n, b.lo, inv_ext, keys[0], vals[0]);
    int cur = 0;
    for (int pass = 0; pass < 8; pass++) {           // 63-bit keys: 8 digits of 8 bits, least significant first
        const int shift = pass * 8;
        k_rs_hist<<<n_tiles, RS_THREADS, 0, st>>>(keys[cur], n, shift, n_tiles, hist);
        exclusive_scan_u32(hist, 256 * n_tiles, scratch, st);
        k_rs_scatter<<<n_tiles, RS_THREADS, 0, st>>>(keys[cur], vals[cur], n, shift, n_tiles, hist, keys[cur ^ 1], vals[cur ^ 1]);
        cur ^= 1;
    }
    TC(cudaGetLastError());
    if (RT_PLOC_RADIUS > 0 && n >= 2) {
        // PLOC (rt_bvh.cuh): clusters = the Morton-ordered leaves, merged round by round; every array of the radix-tree path that is
        // free at this point doubles as a PLOC buffer
        int *cid[2], *nn = parent_node, *first = parent_leaf, *slot_of; float4 *clo[2], *chi[2]; unsigned *merge = arrive, *keep = even, *merge_scan = remap, *keep_scan;
        TA(cid[0], sizeof(int) * n); TA(cid[1], sizeof(int) * n); TA(slot_of, sizeof(int) * n); TA(keep_scan, sizeof(unsigned) * n);
        TA(clo[0], sizeof(float4) * n); TA(clo[1], sizeof(float4) * n); TA(chi[0], sizeof(float4) * n); TA(chi[1], sizeof(float4) * n);
        k_ploc_leaf_boxes<<<G, T, 0, st>>>(b.d_verts, b.d_faces, vals[cur], n, cid[0], clo[0], chi[0]);
        std::vector<int> round_end;
        int m = n, created = 0, pc = 0;
        while (m > 1) {
            const int Gm = (m + T - 1) / T;
            k_ploc_nn<<<Gm, T, 0, st>>>(clo[pc], chi[pc], m, RT_PLOC_RADIUS, nn);
            k_ploc_flags<<<Gm, T, 0, st>>>(nn, m, merge, keep, merge_scan, keep_scan);
            exclusive_scan_u32(merge_scan, m, scratch, st);
            exclusive_scan_u32(keep_scan, m, scratch, st);
            k_ploc_apply<<<Gm, T, 0, st>>>(n, m, created, nn, merge, keep, merge_scan, keep_scan, cid[pc], clo[pc], chi[pc], cid[pc ^ 1], clo[pc ^ 1], chi[pc ^ 1],
                                          children, node_lo, node_hi);
            unsigned last_scan = 0, last_flag = 0;
            TC(cudaMemcpyAsync(&last_scan, merge_scan + (m - 1), sizeof(unsigned), cudaMemcpyDeviceToHost, st));
            TC(cudaMemcpyAsync(&last_flag, merge + (m - 1), sizeof(unsigned), cudaMemcpyDeviceToHost, st));
            TC(cudaStreamSynchronize(st));
            const int merged = (int)(last_scan + last_flag);
            if (merged <= 0) { tfree(); return fail(RT_ERR_CUDA, "LBVH build: a PLOC round merged nothing"); }    // (cannot happen: the closest pair is always mutual)
            created += merged; m -= merged; pc ^= 1;
            round_end.push_back(created);
        }
        TC(cudaMemsetAsync(first, 0, sizeof(int), st));               // the root's first slot
        for (int t = (int)round_end.size() - 1; t >= 0; t--) {
            const int k0 = t ? round_end[t - 1] : 0, k1 = round_end[t];
            k_ploc_place<<<(k1 - k0 + T - 1) / T, T, 0, st>>>(n, k0, k1, children, node_lo, node_hi, first, slot_of);
        }
        k_ploc_emit_leaves<<<G, T, 0, st>>>(b.d_verts, b.d_faces, vals[cur], slot_of, n, b.d_tri, b.d_face_vidx, leaf_lo, leaf_hi, b.d_order);
        TC(cudaGetLastError());
    } else {
    TC(cudaMemcpyAsync(b.d_order, vals[cur], sizeof(int) * n, cudaMemcpyDeviceToDevice, st));
    TC(cudaMemsetAsync(arrive, 0, sizeof(unsigned) * n, st));
    if (n >= 2) k_lbvh_hierarchy<<<G, T, 0, st>>>(keys[cur], n, children, parent_node, parent_leaf);
    k_lbvh_leaves_refit<<<G, T, 0, st>>>(b.d_verts, b.d_faces, vals[cur], n, children, parent_node, parent_leaf, b.d_tri, b.d_face_vidx,
                                        leaf_lo, leaf_hi, node_lo, node_hi, arrive);
    TC(cudaGetLastError());
    }
    // 4-wide collapse: which binary nodes survive as traversal nodes; exclusive scan of the flags = compact, deterministic node indices
    b.n_nodes4 = 0;
    if (n >= 2) {
        // rounds of top-down marking from the root (arrive[] is free again: reused as the marks), a host look at the
        // "changed" flag every 8 rounds
        unsigned* mark = arrive;
        unsigned* d_changed = (unsigned*)hist;
        const unsigned one = 1u;
        TC(cudaMemsetAsync(mark, 0, sizeof(unsigned) * n, st));
        TC(cudaMemcpyAsync(mark, &one, sizeof(unsigned), cudaMemcpyHostToDevice, st));
        for (unsigned round = 1, go = 1; go && round < 4096; ) {
            TC(cudaMemsetAsync(d_changed, 0, sizeof(unsigned), st));
            for (int k = 0; k < 8; k++, round++) k_lbvh_mark<<<G, T, 0, st>>>(n, children, node_lo, node_hi, mark, round, d_changed);
            TC(cudaMemcpyAsync(&go, d_changed, sizeof(unsigned), cudaMemcpyDeviceToHost, st));
            TC(cudaStreamSynchronize(st));
        }
        k_lbvh_flags<<<G, T, 0, st>>>(n, mark, even, remap);
        exclusive_scan_u32(remap, n - 1, scratch, st);
        unsigned last_idx = 0, last_even = 0;
        TC(cudaMemcpyAsync(&last_idx, remap + (n - 2), sizeof(unsigned), cudaMemcpyDeviceToHost, st));
        TC(cudaMemcpyAsync(&last_even, even + (n - 2), sizeof(unsigned), cudaMemcpyDeviceToHost, st));
        TC(cudaStreamSynchronize(st));
        b.n_nodes4 = (int)(last_idx + last_even);
        int rc = upload(sc, (const BvhNode4*)nullptr, (size_t)std::max(1, b.n_nodes4), &b.d_nodes);
        if (rc != RT_OK) { tfree(); return rc; }
        TC(cudaDeviceSynchronize());
        k_lbvh_emit4<<<G, T, 0, st>>>(n, children, even, remap, leaf_lo, leaf_hi, node_lo, node_hi, b.pad, b.d_nodes);
    }
    TC(cudaStreamSynchronize(st));
    TC(cudaGetLastError());
#undef TA
#undef TC
    tfree();
    g.tri = b.d_tri; g.nodes = n >= 2 ? b.d_nodes : nullptr; g.face_vidx = b.d_face_vidx;
    b.built = true;
    return RT_OK;
}

// Developer experiment (B200RT_L2_PERSIST=<MiB>): pin the largest mesh's traversal nodes in a persisting L2 carve-out so the
// per-ray record streams cannot evict them (B200: 126 MB L2).
static void l2_persist_nodes(RtScene* sc, cudaStream_t st) {
    static const char* env = getenv("B200RT_L2_PERSIST");
    if (!env || sc->meshes.empty()) return;
    const size_t want = (size_t)atoi(env) << 20;
    const MeshBuild* big = nullptr;
    for (const MeshBuild& b : sc->meshes) if (b.d_nodes && (!big || b.n_nodes4 > big->n_nodes4)) big = &b;
    if (!big || want == 0) return;
    int max_persist = 0, max_window = 0;
    cudaDeviceGetAttribute(&max_persist, cudaDevAttrMaxPersistingL2CacheSize, sc->device);
    cudaDeviceGetAttribute(&max_window, cudaDevAttrMaxAccessPolicyWindowSize, sc->device);
    const size_t carve = std::min(want, (size_t)max_persist);
    if (cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, carve) != cudaSuccess) { cudaGetLastError(); return; }
    cudaStreamAttrValue av{};
    const size_t bytes = std::min((size_t)big->n_nodes4 * sizeof(BvhNode4), (size_t)max_window);
    av.accessPolicyWindow.base_ptr = (void*)big->d_nodes;
    av.accessPolicyWindow.num_bytes = bytes;
    av.accessPolicyWindow.hitRatio = bytes > carve ? (float)carve / (float)bytes : 1.0f;
    av.accessPolicyWindow.hitProp = cudaAccessPropertyPersisting;
    av.accessPolicyWindow.missProp = cudaAccessPropertyStreaming;
    if (cudaStreamSetAttribute(st, cudaStreamAttributeAccessPolicyWindow, &av) != cudaSuccess) cudaGetLastError();
    fprintf(stderr, "b200rt: L2 persisting window %zu MB over %zu MB of nodes (carve-out %zu MB, max %d MB)\n", bytes >> 20,
            ((size_t)big->n_nodes4 * sizeof(BvhNode4)) >> 20, carve >> 20, max_persist >> 20);
}

extern "C" int rt_scene_build_bvh(RtScene* sc, void* stream) {
    if (!sc) return fail(RT_ERR_INVALID, "null scene");
    DeviceGuard dg(sc->device);
    for (size_t i = 0; i < sc->meshes.size(); i++)
        if (!sc->meshes[i].built) UP(build_mesh_bvh(sc, (int)i, (cudaStream_t)stream));
    if (!sc->h_meshes.empty())
        CU(cudaMemcpy(sc->d_meshes, sc->h_meshes.data(), sizeof(DevMesh) * sc->h_meshes.size(), cudaMemcpyHostToDevice));
    rtb::fill_const_meshes(sc->view, sc->h_objs, sc->h_meshes.data());
    sc->view_dev_valid = false;
    return RT_OK;
}

extern "C" int rt_scene_bvh_stats(const RtScene* sc, int32_t mesh, int64_t* n_nodes, int64_t* n_leaf_tris, int64_t* bytes) {
    if (!sc || mesh < 0 || mesh >= (int)sc->meshes.size()) return fail(RT_ERR_INVALID, "mesh index");
    const MeshBuild& b = sc->meshes[mesh];
    if (n_nodes) *n_nodes = b.n_nodes4;
    if (n_leaf_tris) *n_leaf_tris = b.n_faces;
    if (bytes) *bytes = (int64_t)b.n_nodes4 * (int64_t)sizeof(BvhNode4) + (int64_t)b.n_faces * (48 + 16) + (b.d_vnormals ? (int64_t)b.n_verts * 16 : 0);
    return RT_OK;
}

extern "C" int rt_scene_bvh_download(const RtScene* sc, int32_t mesh, float* h_nodes, int32_t* h_order) {
    if (!sc || mesh < 0 || mesh >= (int)sc->meshes.size()) return fail(RT_ERR_INVALID, "mesh index");
    const MeshBuild& b = sc->meshes[mesh];
    if (!b.built) return fail(RT_ERR_INVALID, "BVH not built");
    if (h_nodes && b.n_nodes4 > 0) CU(cudaMemcpy(h_nodes, b.d_nodes, sizeof(BvhNode4) * (size_t)b.n_nodes4, cudaMemcpyDeviceToHost));
    if (h_order && b.n_faces >= 1) CU(cudaMemcpy(h_order, b.d_order, sizeof(int) * (size_t)b.n_faces, cudaMemcpyDeviceToHost));
    return RT_OK;
}

// ---------------------------------------------------------------------------------------------------
// render
static int make_pixel_map(const RtRenderJob* job, PixelMap& pm, long long& n_local_tiles) {
    if (!rtb::make_pixel_map(job, pm, n_local_tiles)) return fail(RT_ERR_INVALID, "bad RtRenderJob");
    return RT_OK;
}

extern "C" int rt_job_extent(const RtRenderJob* job, int64_t* n_local_tiles, int64_t* n_floats) {
    PixelMap pm; long long nt;
    UP(make_pixel_map(job, pm, nt));
    if (n_local_tiles) *n_local_tiles = nt;
    if (n_floats) *n_floats = ((job->direct && job->world == 1) || job->direct == RT_DIRECT_SHARED_FRAME) ? (int64_t)job->ncols * job->nrows * 3 : pm.n_local_pixels * 3;
    return RT_OK;
}

template <class T>
static int grow(T** buf, long long* cap, long long need, size_t elems_per_unit = 1) {
    if (need <= *cap) return RT_OK;
    if (*buf) { cudaFree(*buf); *buf = nullptr; *cap = 0; }
    void* p = nullptr;
    CU(cudaMalloc(&p, (size_t)need * elems_per_unit * sizeof(T)));
    *buf = (T*)p; *cap = need;
    return RT_OK;
}

// The scene argument of the HAS_NODES-templated kernels: by value, or the pointer to the device copy, which is brought up to date
// first (it changes only with the job's mesh_mode test bits and when the LBVHs are built; a plain render copies nothing).
static int sync_device_view(RtScene* sc, const SceneView& V, cudaStream_t st) {
    if (sc->view_dev_valid && memcmp(&sc->h_view_dev, &V, sizeof V) == 0) return RT_OK;
    if (sc->view_dev_valid) CU(cudaStreamSynchronize(st));     // (kernels of an earlier call on another stream may still read the old copy: callers keep one call in flight)
    memcpy(&sc->h_view_dev, &V, sizeof V);
    CU(cudaMemcpyAsync(sc->d_view, &sc->h_view_dev, sizeof V, cudaMemcpyHostToDevice, st));
    sc->view_dev_valid = true;
    return RT_OK;
}
template <int HAS_NODES> static ViewArg<(HAS_NODES != 0)> view_arg(RtScene* sc, const SceneView& V) {
    if constexpr (HAS_NODES != 0) return ViewArg<true>{V, sc->d_view};
    else return ViewArg<false>{V, nullptr};
}
// kernel variant of a scene (rt_core.cuh): 0 lean, 1 nodes without any mesh, 2 nodes and meshes
static int scene_variant(const RtScene* sc) { return !sc->has_nodes ? RT_V_LEAN : (sc->has_mesh ? RT_V_NODES_MESH : RT_V_NODES); }

static void sum_counters(const RtCounters* cls, RtCounters* out) {
    memset(out, 0, sizeof *out);
    uint64_t* o = (uint64_t*)out;
    for (int c = 0; c < RT_COUNTER_CLASSES; c++) {
        const uint64_t* a = (const uint64_t*)&cls[c];
        for (size_t k = 0; k < sizeof(RtCounters) / sizeof(uint64_t); k++) o[k] += a[k];
    }
    out->error_flags = 0;
    for (int c = 0; c < RT_COUNTER_CLASSES; c++) out->error_flags |= cls[c].error_flags;
}

// Batch size of a frame when the job does not name one.  Two work sets of 64 Mi rays (samples 1 + hit planes 2 + occlusion 0.2 + queues
// 2 x 5 GiB each) cost what one set of 128 Mi rays did.  A frame that fits one batch stays one batch on the caller's stream: cutting
// a rank's 66 M-ray share of the 4K frame in two to have something to overlap measured 15.97 instead of 15.30 ms.  Scenes with CSG /
// transform nodes are the exception: their kernels are latency-bound at 4 blocks/SM, two batches in flight fill the machine better
// (NovelScene1 at its shipped 2048x1024x32 = 64 Mi primary rays: 60.0 ms in two batches, 66.6 in one), so above RT_SPLIT_RAYS_NODES
// rays such a frame is cut in at least two.
#ifndef RT_DEFAULT_BATCH_RAYS
#define RT_DEFAULT_BATCH_RAYS (64ll << 20)
#endif
#ifndef RT_SPLIT_RAYS_NODES
#define RT_SPLIT_RAYS_NODES (16ll << 20)
#endif
// developer override for experiments: B200RT_STREAMS=1 keeps every batch on the caller's stream (echoed by bench.py)
// 1 = every batch on the caller's stream; 2 = whole batches alternate between two streams; 3 = pipeline: level 0 of every batch on
// the caller's stream, the deeper levels + resolve of batch b on a high-priority side stream under level 0 of batch b + 1
#ifndef RT_STREAM_MODE
#define RT_STREAM_MODE 3
#endif
static int max_streams() {
    static const char* env = getenv("B200RT_STREAMS");
    return env ? std::max(1, std::min(3, atoi(env))) : RT_STREAM_MODE;
}

static int grow_set(RtScene* sc, WorkSet& w, long long batch_cap, int L, bool secondary) {
    UP(grow(&w.d_samples, &w.samples_cap, batch_cap));
    UP(grow(&w.d_hit, &w.hit_cap, batch_cap, RT_HIT_PLANES));
    if (w.hit_cap * L > w.occ_cap) {                          // occlusion planes share the hit planes' stride
        if (w.d_occ) { cudaFree(w.d_occ); w.d_occ = nullptr; w.occ_cap = 0; }
        CU(cudaMalloc((void**)&w.d_occ, (size_t)w.hit_cap * L));
        w.occ_cap = w.hit_cap * L;
    }
    if (secondary && batch_cap > w.queue_cap) {
        for (int k = 0; k < 2; k++) {
            if (w.d_queue[k]) { cudaFree(w.d_queue[k]); w.d_queue[k] = nullptr; }
            w.queue_cap = 0;
            CU(cudaMalloc((void**)&w.d_queue[k], sizeof(float4) * RT_QUEUE_PLANES * (size_t)batch_cap));
        }
        w.queue_cap = batch_cap;
    }
    (void)sc;
    return RT_OK;
}

template <int HAS_NODES>
static int render_impl(RtScene* sc, const RtRenderJob* job, const PixelMap& pm, float* d_out, cudaStream_t st, unsigned long long* launches) {
    SceneView V = sc->view;
    V.mesh_mode = job->mesh_mode;
    const int N = sc->samples_per_pixel;
    // batch = as many whole pixels as fit in max_batch_rays, evened out over the batches so the last one is not a sliver
    long long max_rays = job->max_batch_rays > 0 ? job->max_batch_rays : RT_DEFAULT_BATCH_RAYS;
    max_rays = std::min(max_rays, 1ll << 30);
    long long px_per_batch = std::max(1ll, max_rays / N);
    long long n_batches = std::max(1ll, (pm.n_local_pixels + px_per_batch - 1) / px_per_batch);
    if (HAS_NODES && job->max_batch_rays <= 0 && n_batches == 1 && pm.n_local_pixels * N > RT_SPLIT_RAYS_NODES && !sc->serial_only) n_batches = 2;
    if (pm.n_local_pixels > 0) px_per_batch = std::min(px_per_batch, (pm.n_local_pixels + n_batches - 1) / n_batches);
    const long long batch_cap = std::min(px_per_batch, std::max(1ll, pm.n_local_pixels)) * N;
    // two streams when there are two batches to overlap; the per-launch event pairs of the profiler want one launch at a time
    const bool two = n_batches >= 2 && !sc->serial_only && !sc->profiling && max_streams() >= 2;
    const int L = std::max(1, V.n_lights);
    const bool secondary = sc->has_specular && V.max_depth > 1;
    for (int k = 0; k < (two ? 2 : 1); k++) UP(grow_set(sc, sc->ws[k], batch_cap, L, secondary));
    const bool pipe = two && max_streams() >= 3;
    if (two && !sc->side) {
        int lo_prio = 0, hi_prio = 0;
        CU(cudaDeviceGetStreamPriorityRange(&lo_prio, &hi_prio));
        CU(cudaStreamCreateWithPriority(&sc->side, cudaStreamNonBlocking, hi_prio));      // its (small) launches go first when blocks retire
        CU(cudaEventCreateWithFlags(&sc->ev_fork, cudaEventDisableTiming));
        CU(cudaEventCreateWithFlags(&sc->ev_join, cudaEventDisableTiming));
        for (int k = 0; k < 2; k++) { CU(cudaEventCreateWithFlags(&sc->ev_l0[k], cudaEventDisableTiming)); CU(cudaEventCreateWithFlags(&sc->ev_deep[k], cudaEventDisableTiming)); }
    }
    const size_t smem = (sc->has_mesh || sc->has_tlas) ? sizeof(int) * RT_STACK_ROWS * RT_BLOCK : 0;
    if (smem) {
        CU(cudaFuncSetAttribute(k_hit<HAS_NODES, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        CU(cudaFuncSetAttribute(k_hit<HAS_NODES, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        CU(cudaFuncSetAttribute(k_shadow<HAS_NODES, SHADOW_PER_RAY>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        CU(cudaFuncSetAttribute(k_shadow<HAS_NODES, SHADOW_ROWS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    }
    // direct: k_resolve writes the reference layout (ncols, nrows, 3) instead of the compact tile buffer.  With world > 1 only for
    // RT_DIRECT_SHARED_FRAME: d_out is then the WHOLE frame (typically rank 0's buffer mapped through NVLink peer access) and
    // this rank fills in its own tiles - the fan-in of glue.py:19-22 fused into the last kernel of the render.
    const int direct = ((job->direct && job->world == 1) || job->direct == RT_DIRECT_SHARED_FRAME) ? 1 : 0;
    const int waves = grid_waves(false), waves_deep = grid_waves(true);
    const int b_hit = sc->sm_count * (HAS_NODES ? RT_MIN_BLOCKS_HIT_CSG : RT_MIN_BLOCKS_HIT);
    const int b_shadow = sc->sm_count * (HAS_NODES ? RT_MIN_BLOCKS_SHADOW_CSG : RT_MIN_BLOCKS_SHADOW);
    const int b_shade = sc->sm_count * RT_MIN_BLOCKS_SHADE;
    const int g_hit0 = b_hit * waves, g_shadow0 = b_shadow * waves, g_shade0 = b_shade * waves;              // level 0
    const int g_hit = b_hit * waves_deep, g_shadow = b_shadow * waves_deep, g_shade = b_shade * waves_deep;   // queued levels
    const int g_tail = sc->sm_count * waves_deep * (HAS_NODES ? RT_MIN_BLOCKS_HIT_CSG : RT_MIN_BLOCKS_TAIL);
    PixelMap pmk = pm;                                   // + the sample-slot divisors of this scene
    pmk.fd_N = make_fastdiv((unsigned)N); pmk.fd_T = make_fastdiv((unsigned)V.T); pmk.fd_S = make_fastdiv((unsigned)V.S);
    if (HAS_NODES) UP(sync_device_view(sc, V, st));
    const ViewArg<(HAS_NODES != 0)> VA = view_arg<HAS_NODES>(sc, V);
    if (two) {                                           // the side stream starts after everything the caller queued before this call
        CU(cudaEventRecord(sc->ev_fork, st));
        CU(cudaStreamWaitEvent(sc->side, sc->ev_fork, 0));
    }

    // the hit/occlusion planes are indexed with the stride of their allocation; occ planes use H.cap as stride
    long long bi = 0;
    for (long long p0 = 0; p0 < pm.n_local_pixels; p0 += px_per_batch, bi++) {
        const long long npx = std::min(px_per_batch, pm.n_local_pixels - p0);
        const long long n_rays = npx * N;
        WorkSet& w = sc->ws[two ? (bi & 1) : 0];
        cudaStream_t s = (two && !pipe && (bi & 1)) ? sc->side : st;      // a stream's batches run in order, so its set is free again when the next one starts
        HitPlanes H{w.d_hit, w.hit_cap};
        if (pipe && bi >= 2) CU(cudaStreamWaitEvent(st, sc->ev_deep[bi & 1], 0));      // the set is free when the deeper levels of batch bi - 2 are done
        CU(cudaMemsetAsync(w.d_qcount, 0, sizeof(unsigned long long) * (size_t)(V.max_depth + 2), s));
        RayQueue q0{w.d_queue[0], w.queue_cap}, q1{w.d_queue[1], w.queue_cap};
        const long long need = (n_rays + RT_BLOCK - 1) / RT_BLOCK;
        {   // level 0: generated rays
            { ProfScope ps_(sc, PC_HIT0, s);
              k_hit<HAS_NODES, true><<<(int)std::min<long long>(need, g_hit0), RT_BLOCK, smem, s>>>(VA, pmk, p0, n_rays, q0, nullptr, H, sc->d_counters + 0); }
            if (V.n_lights > 0) { ProfScope ps_(sc, PC_SHADOW0, s);
              k_shadow<HAS_NODES, SHADOW_PER_RAY><<<(int)std::min<long long>(need, g_shadow0), RT_BLOCK, smem, s>>>(VA, n_rays, nullptr, H, w.d_occ, sc->d_counters + 1);
              (*launches)++; }
            { ProfScope ps_(sc, PC_SHADE0, s);
              k_shade<true><<<(int)std::min<long long>(need, g_shade0), RT_BLOCK, 0, s>>>(V, pmk, p0, n_rays, q0, nullptr, H, w.d_occ, w.d_samples, q0, w.d_qcount, sc->d_counters + 2); }
            (*launches) += 2;
        }
        if (pipe) {                                      // the rest of this batch runs on the side stream, under level 0 of the next batch
            CU(cudaEventRecord(sc->ev_l0[bi & 1], st));
            CU(cudaStreamWaitEvent(sc->side, sc->ev_l0[bi & 1], 0));
            s = sc->side;
        }
        if (secondary) {
            const int tail_level = tail_level_for(HAS_NODES);
            for (int level = 1; level < V.max_depth; level++) {       // level L consumes the rays spawned by level L-1
                RayQueue& qin = (level & 1) ? q0 : q1;
                RayQueue& qout = (level & 1) ? q1 : q0;
                const unsigned long long* n_in = w.d_qcount + (level - 1);
                if (tail_level > 0 && level >= tail_level) {          // the small remainder: every path to its end in one launch
                    ProfScope ps_(sc, PC_TAIL, s);
                    k_tail<HAS_NODES><<<g_tail, RT_BLOCK, smem, s>>>(VA, qin, n_in, w.d_qcount + level, w.d_samples, sc->d_counters + 3);   // d_qcount[level]: zero, unused from here on: the fetch counter
                    (*launches)++;
                    break;
                }
                { ProfScope ps_(sc, level == 1 ? PC_HIT1 : PC_DEEP, s);
                  k_hit<HAS_NODES, false><<<g_hit, RT_BLOCK, smem, s>>>(VA, pmk, 0, 0, qin, n_in, H, sc->d_counters + 3); }
                if (V.n_lights > 0) { ProfScope ps_(sc, level == 1 ? PC_SHADOW1 : PC_DEEP, s);
                  k_shadow<HAS_NODES, SHADOW_ROWS><<<dim3((unsigned)std::max(1, g_shadow / V.n_lights), (unsigned)V.n_lights), RT_BLOCK, smem, s>>>(VA, 0, n_in, H, w.d_occ, sc->d_counters + 3); (*launches)++; }
                { ProfScope ps_(sc, level == 1 ? PC_SHADE1 : PC_DEEP, s);
                  k_shade<false><<<g_shade, RT_BLOCK, 0, s>>>(V, pmk, 0, 0, qin, n_in, H, w.d_occ, w.d_samples, qout, w.d_qcount + level, sc->d_counters + 3); }
                (*launches) += 2;
            }
        }
        {
            ProfScope ps_(sc, PC_RESOLVE, s);
            k_resolve<<<(int)((npx + 127) / 128), 128, 0, s>>>(pm, p0, (int)npx, N, w.d_samples, d_out, direct);
        }
        (*launches)++;
        if (pipe) CU(cudaEventRecord(sc->ev_deep[bi & 1], sc->side));
        if (sc->sink_host && (pipe || !two)) {          // (batches resolve in order on one stream)  floats of the output that are final after this batch: whole tile columns (reference layout) / the batch's slots (compact)
            const bool last = p0 + px_per_batch >= pm.n_local_pixels;
            long long f1;
            if (direct) f1 = (last ? (long long)pm.ncols : std::min<long long>(pm.ncols, ((p0 + npx) / pm.tile_px) / pm.n_tile_rows * pm.tile_w)) * pm.nrows * 3;
            else f1 = (p0 + npx) * 3;
            if (f1 > sc->sink_done) {
                CU(cudaEventRecord(sc->ev_copy, s));
                CU(cudaStreamWaitEvent(sc->copy_stream, sc->ev_copy, 0));
                CU(cudaMemcpyAsync(sc->sink_host + sc->sink_done, d_out + sc->sink_done, sizeof(float) * (size_t)(f1 - sc->sink_done), cudaMemcpyDeviceToHost, sc->copy_stream));
                sc->sink_done = f1;
            }
        }
    }
    if (sc->sink_host) {                                 // the caller's stream continues when the last chunk has landed
        CU(cudaEventRecord(sc->ev_copy, sc->copy_stream));
        CU(cudaStreamWaitEvent(st, sc->ev_copy, 0));
    }
    if (two) {                                           // the caller's stream continues when the side stream's batches are done too
        CU(cudaEventRecord(sc->ev_join, sc->side));
        CU(cudaStreamWaitEvent(st, sc->ev_join, 0));
    }
    CU(cudaGetLastError());
    return RT_OK;
}

extern "C" int rt_render(RtScene* sc, const RtRenderJob* job, float* d_out, RtCounters* counters, void* stream) {
    if (!sc || !d_out) return fail(RT_ERR_INVALID, "null argument");
    for (auto& m : sc->meshes) if (!m.built) return fail(RT_ERR_INVALID, "rt_scene_build_bvh has not been called");
    PixelMap pm; long long nt;
    UP(make_pixel_map(job, pm, nt));
    // the sampling tables (col_x[W], row_y[H]) are indexed by absolute pixel coordinates: the job must stay inside the frame
    if (job->col0 < 0 || job->row0 < 0 || (long long)job->col0 + job->ncols > sc->view.width || (long long)job->row0 + job->nrows > sc->view.height)
        return fail(RT_ERR_INVALID, "RtRenderJob rectangle leaves the frame [0,width) x [0,height)");
    cudaStream_t st = (cudaStream_t)stream;
    DeviceGuard dg(sc->device);
    if (!sc->l2_window_tried) { sc->l2_window_tried = true; l2_persist_nodes(sc, st); }
    CU(cudaMemsetAsync(sc->d_counters, 0, RT_COUNTER_CLASSES * sizeof(RtCounters), st));
    unsigned long long launches = 0;
    const int variant = scene_variant(sc);
    int rc = variant == RT_V_LEAN ? render_impl<RT_V_LEAN>(sc, job, pm, d_out, st, &launches)
           : variant == RT_V_NODES ? render_impl<RT_V_NODES>(sc, job, pm, d_out, st, &launches) : render_impl<RT_V_NODES_MESH>(sc, job, pm, d_out, st, &launches);
    if (rc != RT_OK) return rc;
    sc->last_launches = launches;
    if (counters) {
        RtCounters cls[RT_COUNTER_CLASSES];
        CU(cudaMemcpyAsync(cls, sc->d_counters, sizeof cls, cudaMemcpyDeviceToHost, st));
        CU(cudaStreamSynchronize(st));
        sum_counters(cls, counters);
        counters->kernel_launches = launches;
    }
    return RT_OK;
}

// ---------------------------------------------------------------------------------------------------
// rt_render without any allocation (no counterpart in the reference; added for callers
// that cannot call cudaMalloc: CUDA graph capture, tight serving loops, harnesses that forbid it).  The wavefront scratch of the
// render (sample buffer, hit planes, occlusion planes, both ray queues) lives in ONE file-scope device array instead of the
// handle's grown buffers; the frame is worked through in batches of <= RT_STATIC_MAX_RAYS rays.  Every GPU that loads the
// library holds the array until the process exits; every call on a device shares it (calls that overlap on the GPU race).
// Nothing in it has to start at zero: every record a kernel reads was written by an earlier kernel of the same batch.
#define RT_STATIC_ALIGN(x) (((size_t)(x) + 255) & ~(size_t)255)
#define RT_STATIC_OFF_SAMPLES 0
#define RT_STATIC_OFF_HIT (RT_STATIC_OFF_SAMPLES + RT_STATIC_ALIGN(sizeof(float4) * RT_STATIC_MAX_RAYS))
#define RT_STATIC_OFF_OCC (RT_STATIC_OFF_HIT + RT_STATIC_ALIGN(sizeof(float4) * RT_HIT_PLANES * RT_STATIC_MAX_RAYS))
#define RT_STATIC_OFF_Q0 (RT_STATIC_OFF_OCC + RT_STATIC_ALIGN((size_t)RT_STATIC_MAX_LIGHTS * RT_STATIC_MAX_RAYS))
#define RT_STATIC_OFF_Q1 (RT_STATIC_OFF_Q0 + RT_STATIC_ALIGN(sizeof(float4) * RT_QUEUE_PLANES * RT_STATIC_MAX_RAYS))
#define RT_STATIC_BYTES (RT_STATIC_OFF_Q1 + RT_STATIC_ALIGN(sizeof(float4) * RT_QUEUE_PLANES * RT_STATIC_MAX_RAYS))
static __device__ __align__(256) unsigned char g_static_scratch[RT_STATIC_BYTES];      // 216 B per ray: 13.5 MiB
// In host code `g_static_scratch` is only a stand-in for the symbol: the device address is asked for once per device, with
// that device current, and kept here.
#define RT_STATIC_MAX_DEVICES 64
static unsigned char* g_static_ptr[RT_STATIC_MAX_DEVICES];

static int static_scratch(int device, unsigned char** out) {
    if (device < 0 || device >= RT_STATIC_MAX_DEVICES) return fail(RT_ERR_LIMIT, "rt_render_static: device ordinal >= 64");
    if (!g_static_ptr[device]) {
        void* p = nullptr;
        CU(cudaGetSymbolAddress(&p, g_static_scratch));       // the caller made `device` current
        g_static_ptr[device] = (unsigned char*)p;
    }
    *out = g_static_ptr[device];
    return RT_OK;
}

extern "C" int rt_render_static(RtScene* sc, const RtRenderJob* job, float* d_out, RtCounters* counters, void* stream) {
    if (!sc || !job || !d_out) return fail(RT_ERR_INVALID, "null argument");
    for (auto& m : sc->meshes) if (!m.built) return fail(RT_ERR_INVALID, "rt_scene_build_bvh has not been called");
    if (sc->samples_per_pixel > RT_STATIC_MAX_RAYS)
        return fail(RT_ERR_LIMIT, "rt_render_static: " + std::to_string(sc->samples_per_pixel) + " samples per pixel exceed the static scratch ("
                                  + std::to_string(RT_STATIC_MAX_RAYS) + " rays per batch); use rt_render");
    if (sc->view.n_lights > RT_STATIC_MAX_LIGHTS)
        return fail(RT_ERR_LIMIT, "rt_render_static: " + std::to_string(sc->view.n_lights) + " lights exceed the static scratch ("
                                  + std::to_string(RT_STATIC_MAX_LIGHTS) + "); use rt_render");
    PixelMap pm; long long nt;
    UP(make_pixel_map(job, pm, nt));
    if (job->col0 < 0 || job->row0 < 0 || (long long)job->col0 + job->ncols > sc->view.width || (long long)job->row0 + job->nrows > sc->view.height)
        return fail(RT_ERR_INVALID, "RtRenderJob rectangle leaves the frame [0,width) x [0,height)");
    cudaStream_t st = (cudaStream_t)stream;
    DeviceGuard dg(sc->device);
    unsigned char* base = nullptr;
    UP(static_scratch(sc->device, &base));
    // the handle's own work buffers step aside for the duration of the call (render_impl sizes its batches from max_batch_rays and
    // never grows a buffer whose capacity already covers a batch); event pairs of the profiler are not recorded here either
    const WorkSet keep = sc->ws[0];
    const bool keep_profiling = sc->profiling;
    WorkSet& w = sc->ws[0];
    w.d_samples = (float4*)(base + RT_STATIC_OFF_SAMPLES); w.samples_cap = RT_STATIC_MAX_RAYS;
    w.d_hit = (float4*)(base + RT_STATIC_OFF_HIT); w.hit_cap = RT_STATIC_MAX_RAYS;
    w.d_occ = (uint8_t*)(base + RT_STATIC_OFF_OCC); w.occ_cap = (long long)RT_STATIC_MAX_LIGHTS * RT_STATIC_MAX_RAYS;
    w.d_queue[0] = (float4*)(base + RT_STATIC_OFF_Q0); w.d_queue[1] = (float4*)(base + RT_STATIC_OFF_Q1); w.queue_cap = RT_STATIC_MAX_RAYS;
    sc->profiling = false;
    sc->serial_only = true;                 // one work set (the static array) on the caller's stream: no second stream, no events
    RtRenderJob j = *job;
    j.max_batch_rays = RT_STATIC_MAX_RAYS;
    unsigned long long launches = 0;
    int rc = RT_OK;
    cudaError_t e = cudaMemsetAsync(sc->d_counters, 0, RT_COUNTER_CLASSES * sizeof(RtCounters), st);
    if (e != cudaSuccess) rc = fail(RT_ERR_CUDA, std::string("rt_render_static: cudaMemsetAsync: ") + cudaGetErrorString(e));
    const int variant = scene_variant(sc);
    if (rc == RT_OK) rc = variant == RT_V_LEAN ? render_impl<RT_V_LEAN>(sc, &j, pm, d_out, st, &launches)
                        : variant == RT_V_NODES ? render_impl<RT_V_NODES>(sc, &j, pm, d_out, st, &launches) : render_impl<RT_V_NODES_MESH>(sc, &j, pm, d_out, st, &launches);
    const bool intact = w.d_samples == (float4*)(base + RT_STATIC_OFF_SAMPLES) && w.d_hit == (float4*)(base + RT_STATIC_OFF_HIT) &&
                        w.d_occ == (uint8_t*)(base + RT_STATIC_OFF_OCC) && w.d_queue[0] == (float4*)(base + RT_STATIC_OFF_Q0) &&
                        w.d_queue[1] == (float4*)(base + RT_STATIC_OFF_Q1);
    sc->ws[0] = keep;
    sc->profiling = keep_profiling; sc->serial_only = false;
    if (rc != RT_OK) return rc;
    if (!intact) return fail(RT_ERR_LIMIT, "rt_render_static: a batch outgrew the static scratch");
    sc->last_launches = launches;
    if (counters) {
        RtCounters cls[RT_COUNTER_CLASSES];
        CU(cudaMemcpyAsync(cls, sc->d_counters, sizeof cls, cudaMemcpyDeviceToHost, st));
        CU(cudaStreamSynchronize(st));
        sum_counters(cls, counters);
        counters->kernel_launches = launches;
    }
    return RT_OK;
}

extern "C" int rt_render_host(RtScene* sc, const RtRenderJob* job, float* h_out, RtCounters* counters, void* stream) {
    if (!sc || !h_out || !job) return fail(RT_ERR_INVALID, "null argument");
    int64_t n_floats = 0;
    UP(rt_job_extent(job, nullptr, &n_floats));
    UP(grow(&sc->d_stage, &sc->stage_cap, std::max<long long>(n_floats, 1)));
    cudaStream_t st = (cudaStream_t)stream;
    // A pinned destination (cudaHostAlloc / cudaHostRegister) takes the frame in chunks: what a wavefront batch completes is copied
    // on a third stream while the next batches render, so only the last chunk's copy is left after the last kernel.  Pageable
    // memory gets one copy at the end (an "async" copy to pageable memory blocks the host, which would stop the launches).
    cudaPointerAttributes pa{};
    const bool pinned = cudaPointerGetAttributes(&pa, h_out) == cudaSuccess && pa.type == cudaMemoryTypeHost;
    cudaGetLastError();
    const bool chunked = pinned && job->direct != RT_DIRECT_SHARED_FRAME;
    if (chunked && !sc->copy_stream) {
        DeviceGuard dg(sc->device);
        CU(cudaStreamCreateWithFlags(&sc->copy_stream, cudaStreamNonBlocking));
        CU(cudaEventCreateWithFlags(&sc->ev_copy, cudaEventDisableTiming));
    }
    if (chunked) { sc->sink_host = h_out; sc->sink_done = 0; }
    const int rc = rt_render(sc, job, sc->d_stage, nullptr, stream);
    const long long done = sc->sink_done;
    sc->sink_host = nullptr; sc->sink_done = 0;
    if (rc != RT_OK) return rc;
    if (done < n_floats) CU(cudaMemcpyAsync(h_out + done, sc->d_stage + done, sizeof(float) * (size_t)(n_floats - done), cudaMemcpyDeviceToHost, st));
    RtCounters cls[RT_COUNTER_CLASSES];
    if (counters) CU(cudaMemcpyAsync(cls, sc->d_counters, sizeof cls, cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    if (counters) { sum_counters(cls, counters); counters->kernel_launches = sc->last_launches; }
    return RT_OK;
}

extern "C" int rt_render_u8_host(RtScene* sc, const RtRenderJob* job, uint8_t* h_out, RtCounters* counters, void* stream) {
    if (!sc || !h_out || !job) return fail(RT_ERR_INVALID, "null argument");
    if (job->world != 1) return fail(RT_ERR_INVALID, "rt_render_u8_host renders a whole region (world must be 1)");
    RtRenderJob j = *job; j.direct = 1;
    const long long n = (long long)j.ncols * j.nrows * 3;
    UP(grow(&sc->d_stage, &sc->stage_cap, std::max<long long>(n, 1)));
    UP(grow(&sc->d_stage_u8, &sc->stage_u8_cap, std::max<long long>(n, 1)));
    cudaStream_t st = (cudaStream_t)stream;
    UP(rt_render(sc, &j, sc->d_stage, nullptr, stream));
    UP(rt_finalize_u8(sc->d_stage, j.ncols, j.nrows, sc->d_stage_u8, stream));
    CU(cudaMemcpyAsync(h_out, sc->d_stage_u8, (size_t)n, cudaMemcpyDeviceToHost, st));
    RtCounters cls[RT_COUNTER_CLASSES];
    if (counters) CU(cudaMemcpyAsync(cls, sc->d_counters, sizeof cls, cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    if (counters) { sum_counters(cls, counters); counters->kernel_launches = sc->last_launches + 1; }
    return RT_OK;
}

extern "C" int rt_assemble_tiles(const RtRenderJob* job, const float* d_gathered, const int64_t* rank_offsets_floats, float* d_out, void* stream) {
    PixelMap pm; long long nt;
    UP(make_pixel_map(job, pm, nt));
    if (job->world > RT_MAX_WORLD) return fail(RT_ERR_LIMIT, "world > 64");
    cudaStream_t st = (cudaStream_t)stream;
    RankOffsets off;                                     // by value in the kernel's parameter bank: no allocation, no copy per frame
    for (int r = 0; r < job->world; r++) off.v[r] = rank_offsets_floats[r];
    const long long n = (long long)job->ncols * job->nrows;
    if (n > 0) k_assemble<<<(int)((n + 255) / 256), 256, 0, st>>>(pm, d_gathered, off, d_out);
    CU(cudaGetLastError());
    return RT_OK;
}

// ---------------------------------------------------------------------------------------------------
// Frame buffers shared between the ranks of one box (one process per GPU): rank 0 allocates and exports, the others open the
// handle and get a pointer into rank 0's HBM through NVLink peer access (cudaIpc*).  rt_render with RT_DIRECT_SHARED_FRAME
// then writes every rank's tiles straight into it.
extern "C" int rt_shared_alloc(int64_t bytes, void** d_ptr, RtIpcHandle* handle) {
    if (!d_ptr || !handle || bytes <= 0) return fail(RT_ERR_INVALID, "rt_shared_alloc: bad argument");
    static_assert(sizeof(cudaIpcMemHandle_t) <= sizeof(handle->bytes), "RtIpcHandle too small");
    void* p = nullptr;
    CU(cudaMalloc(&p, (size_t)bytes));                    // a plain allocation of its own: the IPC handle maps exactly this range
    cudaIpcMemHandle_t h;
    cudaError_t e = cudaIpcGetMemHandle(&h, p);
    if (e != cudaSuccess) { cudaFree(p); return fail(RT_ERR_CUDA, std::string("cudaIpcGetMemHandle: ") + cudaGetErrorString(e)); }
    memset(handle->bytes, 0, sizeof handle->bytes);
    memcpy(handle->bytes, &h, sizeof h);
    *d_ptr = p;
    return RT_OK;
}
extern "C" int rt_shared_free(void* d_ptr) { if (d_ptr) CU(cudaFree(d_ptr)); return RT_OK; }
extern "C" int rt_shared_open(const RtIpcHandle* handle, void** d_ptr) {
    if (!d_ptr || !handle) return fail(RT_ERR_INVALID, "rt_shared_open: bad argument");
    cudaIpcMemHandle_t h;
    memcpy(&h, handle->bytes, sizeof h);
    CU(cudaIpcOpenMemHandle(d_ptr, h, cudaIpcMemLazyEnablePeerAccess));
    return RT_OK;
}
extern "C" int rt_shared_close(void* d_ptr) { if (d_ptr) CU(cudaIpcCloseMemHandle(d_ptr)); return RT_OK; }

extern "C" int rt_finalize_u8(const float* d_image, int32_t ncols, int32_t nrows, uint8_t* d_out_u8, void* stream) {
    const long long n = (long long)ncols * nrows;
    if (n > 0) k_finalize_u8<<<(int)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(d_image, ncols, nrows, d_out_u8);
    CU(cudaGetLastError());
    return RT_OK;
}

// ---------------------------------------------------------------------------------------------------
// unit-test aids: explicit rays (host buffers in, host buffers out)
struct DevTmp {
    std::vector<void*> p;
    ~DevTmp() { for (void* q : p) cudaFree(q); }
    template <class T> int get(T** out, size_t n, const T* src = nullptr) {
        void* q = nullptr;
        CU(cudaMalloc(&q, std::max<size_t>(n, 1) * sizeof(T)));
        p.push_back(q);
        if (src && n) CU(cudaMemcpy(q, src, n * sizeof(T), cudaMemcpyHostToDevice));
        *out = (T*)q;
        return RT_OK;
    }
};

extern "C" int rt_trace_rays(RtScene* sc, int64_t n, const float* h_org, const float* h_dir, const float* h_times, float* h_rgb,
                             float* h_t, float* h_n, float* h_p, int32_t* h_geom, int32_t* h_mat, int32_t mesh_mode, void* stream) {
    if (!sc || n < 0) return fail(RT_ERR_INVALID, "bad argument");
    for (auto& m : sc->meshes) if (!m.built) return fail(RT_ERR_INVALID, "rt_scene_build_bvh has not been called");
    if (n == 0) return RT_OK;
    DeviceGuard dg(sc->device);
    cudaStream_t st = (cudaStream_t)stream;
    DevTmp tmp;
    float *org, *dir, *times, *rgb, *t, *nn, *pp; int *geom, *mat;
    UP(tmp.get(&org, 3 * n, h_org)); UP(tmp.get(&dir, 3 * n, h_dir)); UP(tmp.get(&times, n, h_times));
    UP(tmp.get(&rgb, 3 * n)); UP(tmp.get(&t, n)); UP(tmp.get(&nn, 3 * n)); UP(tmp.get(&pp, 3 * n));
    UP(tmp.get(&geom, n)); UP(tmp.get(&mat, n));
    SceneView V = sc->view; V.mesh_mode = mesh_mode;
    const size_t smem = (sc->has_mesh || sc->has_tlas) ? sizeof(int) * RT_STACK_ROWS * RT_BLOCK : 0;
    const int grid = (int)((n + RT_BLOCK - 1) / RT_BLOCK);
    CU(cudaMemsetAsync(sc->d_counters, 0, sizeof(RtCounters), st));
    switch (scene_variant(sc)) {
#define RT_LAUNCH_TRACE(VV)                                                                                                              \
        if (smem) CU(cudaFuncSetAttribute(k_trace<VV>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));                        \
        if (VV != RT_V_LEAN) UP(sync_device_view(sc, V, st));                                                                            \
        k_trace<VV><<<grid, RT_BLOCK, smem, st>>>(view_arg<VV>(sc, V), n, org, dir, times, rgb, t, nn, pp, geom, mat, sc->d_counters);
        case RT_V_LEAN: { RT_LAUNCH_TRACE(RT_V_LEAN) } break;
        case RT_V_NODES: { RT_LAUNCH_TRACE(RT_V_NODES) } break;
        default: { RT_LAUNCH_TRACE(RT_V_NODES_MESH) } break;
#undef RT_LAUNCH_TRACE
    }
    CU(cudaGetLastError());
    CU(cudaStreamSynchronize(st));
    if (h_rgb) CU(cudaMemcpy(h_rgb, rgb, sizeof(float) * 3 * n, cudaMemcpyDeviceToHost));
    if (h_t) CU(cudaMemcpy(h_t, t, sizeof(float) * n, cudaMemcpyDeviceToHost));
    if (h_n) CU(cudaMemcpy(h_n, nn, sizeof(float) * 3 * n, cudaMemcpyDeviceToHost));
    if (h_p) CU(cudaMemcpy(h_p, pp, sizeof(float) * 3 * n, cudaMemcpyDeviceToHost));
    if (h_geom) CU(cudaMemcpy(h_geom, geom, sizeof(int) * n, cudaMemcpyDeviceToHost));
    if (h_mat) CU(cudaMemcpy(h_mat, mat, sizeof(int) * n, cudaMemcpyDeviceToHost));
    return RT_OK;
}

extern "C" int rt_occluded(RtScene* sc, int64_t n, const float* h_org, const float* h_dir, const float* h_times, const float* h_tmax,
                           int32_t* h_occ, int32_t mesh_mode, void* stream) {
    if (!sc || n < 0) return fail(RT_ERR_INVALID, "bad argument");
    for (auto& m : sc->meshes) if (!m.built) return fail(RT_ERR_INVALID, "rt_scene_build_bvh has not been called");
    if (n == 0) return RT_OK;
    DeviceGuard dg(sc->device);
    cudaStream_t st = (cudaStream_t)stream;
    DevTmp tmp;
    float *org, *dir, *times, *tmax; int* occ;
    UP(tmp.get(&org, 3 * n, h_org)); UP(tmp.get(&dir, 3 * n, h_dir)); UP(tmp.get(&times, n, h_times)); UP(tmp.get(&tmax, n, h_tmax));
    UP(tmp.get(&occ, n));
    SceneView V = sc->view; V.mesh_mode = mesh_mode;
    const size_t smem = (sc->has_mesh || sc->has_tlas) ? sizeof(int) * RT_STACK_ROWS * RT_BLOCK : 0;
    const int grid = (int)((n + RT_BLOCK - 1) / RT_BLOCK);
    switch (scene_variant(sc)) {
#define RT_LAUNCH_OCC(VV)                                                                                                                \
        if (smem) CU(cudaFuncSetAttribute(k_occluded<VV>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));                     \
        if (VV != RT_V_LEAN) UP(sync_device_view(sc, V, st));                                                                            \
        k_occluded<VV><<<grid, RT_BLOCK, smem, st>>>(view_arg<VV>(sc, V), n, org, dir, times, tmax, occ, sc->d_counters);
        case RT_V_LEAN: { RT_LAUNCH_OCC(RT_V_LEAN) } break;
        case RT_V_NODES: { RT_LAUNCH_OCC(RT_V_NODES) } break;
        default: { RT_LAUNCH_OCC(RT_V_NODES_MESH) } break;
#undef RT_LAUNCH_OCC
    }
    CU(cudaGetLastError());
    CU(cudaStreamSynchronize(st));
    CU(cudaMemcpy(h_occ, occ, sizeof(int) * n, cudaMemcpyDeviceToHost));
    return RT_OK;
}

// ---------------------------------------------------------------------------------------------------
// profiling
static void drain_events(RtScene* sc) {
    for (int c = 0; c < RT_PROF_CLASSES; c++) {
        for (auto& pr : sc->ev[c]) {
            cudaEventSynchronize(pr.second);
            float ms = 0.0f;
            if (cudaEventElapsedTime(&ms, pr.first, pr.second) == cudaSuccess) { sc->prof_ms[c] += ms; sc->prof_n[c]++; }
            sc->ev_pool.push_back(pr.first); sc->ev_pool.push_back(pr.second);
        }
        sc->ev[c].clear();
    }
}

extern "C" int rt_scene_profile(RtScene* sc, int32_t enable) {
    if (!sc) return fail(RT_ERR_INVALID, "null scene");
    drain_events(sc);
    sc->profiling = enable != 0;
    if (enable) for (int c = 0; c < RT_PROF_CLASSES; c++) { sc->prof_ms[c] = 0; sc->prof_n[c] = 0; }
    return RT_OK;
}

extern "C" int rt_scene_profile_read(RtScene* sc, RtProfile* out) {
    if (!sc || !out) return fail(RT_ERR_INVALID, "null argument");
    drain_events(sc);
    const double* ms = sc->prof_ms; const unsigned long long* nn = sc->prof_n;
    out->ms_hit = ms[PC_HIT0]; out->ms_shadow = ms[PC_SHADOW0]; out->ms_shade = ms[PC_SHADE0]; out->ms_resolve = ms[PC_RESOLVE];
    out->ms_secondary = ms[PC_HIT1] + ms[PC_SHADOW1] + ms[PC_SHADE1] + ms[PC_DEEP] + ms[PC_TAIL];
    out->n_hit = nn[PC_HIT0]; out->n_shadow = nn[PC_SHADOW0]; out->n_shade = nn[PC_SHADE0]; out->n_resolve = nn[PC_RESOLVE];
    out->n_secondary = nn[PC_HIT1] + nn[PC_SHADOW1] + nn[PC_SHADE1] + nn[PC_DEEP] + nn[PC_TAIL];
    out->ms_l1_hit = ms[PC_HIT1]; out->ms_l1_shadow = ms[PC_SHADOW1]; out->ms_l1_shade = ms[PC_SHADE1];
    out->ms_deep = ms[PC_DEEP]; out->ms_tail = ms[PC_TAIL];
    out->n_l1_hit = nn[PC_HIT1]; out->n_deep = nn[PC_DEEP]; out->n_tail = nn[PC_TAIL];
    RtCounters cls[RT_COUNTER_CLASSES];
    CU(cudaMemcpy(cls, sc->d_counters, sizeof cls, cudaMemcpyDeviceToHost));
    out->hit = cls[0]; out->shadow = cls[1]; out->shade = cls[2]; out->secondary = cls[3];
    return RT_OK;
}

// ------------------------------------------------------------------------------------------------
// Roofline denominators measured on the current device (rt_peaks.cuh)
extern "C" int rt_measure_peaks(RtPeaks* out, void* stream) {
    if (!out) return fail(RT_ERR_INVALID, "null argument");
    cudaStream_t st = (cudaStream_t)stream;
    int dev = 0, sms = 0, khz = 0;
    CU(cudaGetDevice(&dev));
    CU(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    CU(cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, dev));
    float* d_out = nullptr; float4* d_table = nullptr;
    CU(cudaMalloc((void**)&d_out, 16));
    CU(cudaMalloc((void**)&d_table, PK_LSU_TABLE_BYTES));
    CU(cudaMemset(d_table, 0, PK_LSU_TABLE_BYTES));
    CU(cudaDeviceSynchronize());
    cudaEvent_t e0, e1;
    CU(cudaEventCreate(&e0)); CU(cudaEventCreate(&e1));
    const int grid = sms * PK_BLOCKS_PER_SM;
    const double threads = (double)grid * PK_BLOCK, warps = threads / 32.0;
    auto timed = [&](int which, int trips, float* ms_out) -> cudaError_t {
        float best = 1e30f;
        for (int rep = 0; rep < 4; rep++) {                    // rep 0 warms up (clocks, instruction cache); best of the other three
            cudaEventRecord(e0, st);
            if (which == 0) k_peak_fma<<<grid, PK_BLOCK, 0, st>>>(d_out, trips, 1.0000001f, 1e-9f);
            else if (which == 1) k_peak_issue<<<grid, PK_BLOCK, 0, st>>>(d_out, trips, 1.0000001f, 1e-9f, 0x9e3779b9u);
            else k_peak_lsu<<<grid, PK_BLOCK, 0, st>>>(d_table, d_out, trips);
            cudaEventRecord(e1, st);
            cudaError_t e = cudaEventSynchronize(e1);
            if (e != cudaSuccess) return e;
            float ms = 0.0f; cudaEventElapsedTime(&ms, e0, e1);
            if (rep > 0 && ms < best) best = ms;
        }
        *ms_out = best;
        return cudaGetLastError();
    };
    float ms_fma = 0, ms_issue = 0, ms_lsu = 0;
    const int trips_fma = 4096, trips_issue = 4096, trips_lsu = 4096;
    cudaError_t e = timed(0, trips_fma, &ms_fma);
    if (e == cudaSuccess) e = timed(1, trips_issue, &ms_issue);
    if (e == cudaSuccess) e = timed(2, trips_lsu, &ms_lsu);
    cudaEventDestroy(e0); cudaEventDestroy(e1); cudaFree(d_out); cudaFree(d_table);
    if (e != cudaSuccess) return fail(RT_ERR_CUDA, std::string("rt_measure_peaks: ") + cudaGetErrorString(e));
    out->fp32_fma_tflops = threads * trips_fma * (double)PK_FMA_PER_TRIP * 2.0 / (ms_fma * 1e-3) / 1e12;
    out->issue_ginst = warps * trips_issue * (double)PK_INST_PER_TRIP / (ms_issue * 1e-3) / 1e9;
    out->lsu_gwavefronts = warps * trips_lsu * (double)PK_LOADS_PER_TRIP * 4.0 / (ms_lsu * 1e-3) / 1e9;
    out->sm_count = sms; out->sm_clock_mhz = khz / 1000.0;
    out->fp32_fma_tflops_nominal = sms * 128.0 * 2.0 * (khz * 1e3) / 1e12;          // 128 FP32 lanes per SM, FMA = 2 flops
    out->issue_ginst_nominal = sms * 4.0 * (khz * 1e3) / 1e9;                        // 4 schedulers per SM, 1 warp instruction per cycle
    out->lsu_gwavefronts_nominal = sms * 1.0 * (khz * 1e3) / 1e9;                    // one 128-byte wavefront per cycle per SM
    return RT_OK;
}

// ------------------------------------------------------------------------------------------------
// Mesh ingestion: OBJ rows -> arrays on the host cores, everything per-vertex / per-face on the device
struct RtObj { ObjData d; };

extern "C" int rt_obj_load(const char* path, RtObj** out, int64_t* n_verts, int64_t* n_norms, int64_t* n_faces) {
    if (!path || !out) return fail(RT_ERR_INVALID, "rt_obj_load: null argument");
    RtObj* o = new RtObj();
    const std::string err = objio::load(path, o->d);
    if (!err.empty()) { delete o; return fail(RT_ERR_INVALID, err); }
    const int64_t nv = (int64_t)(o->d.verts.size() / 3);
    for (int32_t i : o->d.faces)
        if (i < 0 || i >= nv) { delete o; return fail(RT_ERR_INVALID, std::string(path) + ": face index out of range"); }
    if (n_verts) *n_verts = nv;
    if (n_norms) *n_norms = (int64_t)(o->d.norms.size() / 3);
    if (n_faces) *n_faces = (int64_t)(o->d.faces.size() / 3);
    *out = o;
    return RT_OK;
}

extern "C" int rt_obj_copy(const RtObj* o, double* verts, double* norms, int32_t* faces) {
    if (!o) return fail(RT_ERR_INVALID, "rt_obj_copy: null handle");
    if (verts && !o->d.verts.empty()) memcpy(verts, o->d.verts.data(), sizeof(double) * o->d.verts.size());
    if (norms && !o->d.norms.empty()) memcpy(norms, o->d.norms.data(), sizeof(double) * o->d.norms.size());
    if (faces && !o->d.faces.empty()) memcpy(faces, o->d.faces.data(), sizeof(int32_t) * o->d.faces.size());
    return RT_OK;
}

extern "C" int rt_obj_free(RtObj* o) { delete o; return RT_OK; }

extern "C" int rt_mesh_ingest(int32_t n_verts, const double* h_raw_verts, int32_t n_faces, const int32_t* h_faces,
                              const float* translate, float scale, int32_t want_normals, float* h_verts, float* h_vnormals,
                              float* h_extent_min, float* h_extent_max, float* h_center, float* h_max_dist, void* stream) {
    if (n_verts < 0 || n_faces < 0 || (n_verts && (!h_raw_verts || !h_verts)) || (n_faces && !h_faces) || !translate)
        return fail(RT_ERR_INVALID, "rt_mesh_ingest: bad argument");
    if (want_normals && n_verts && !h_vnormals) return fail(RT_ERR_INVALID, "rt_mesh_ingest: normals wanted but no output buffer");
    if ((long long)n_faces * 3 > 0x7fffffffLL) return fail(RT_ERR_LIMIT, "rt_mesh_ingest: more than 2^31 face corners");
    for (long long i = 0; i < (long long)n_faces * 3; i++)
        if (h_faces[i] < 0 || h_faces[i] >= n_verts) return fail(RT_ERR_INVALID, "rt_mesh_ingest: face index out of range");
    if (n_verts == 0) return fail(RT_ERR_INVALID, "rt_mesh_ingest: a mesh needs at least one vertex (mesh.py:29 takes max() of them)");
    cudaStream_t st = (cudaStream_t)stream;
    std::vector<void*> tmp;
    auto tfree = [&]() { for (void* p : tmp) cudaFree(p); };
    auto talloc = [&](void** p, size_t bytes) { cudaError_t e = cudaMalloc(p, std::max<size_t>(bytes, 16)); if (e == cudaSuccess) tmp.push_back(*p); return e; };
#define TA(ptr, bytes) do { cudaError_t e_ = talloc((void**)&(ptr), (bytes)); if (e_ != cudaSuccess) { tfree(); return fail(RT_ERR_NOMEM, cudaGetErrorString(e_)); } } while (0)
#define TC(expr) do { cudaError_t e_ = (expr); if (e_ != cudaSuccess) { tfree(); return fail(RT_ERR_CUDA, std::string(#expr) + ": " + cudaGetErrorString(e_)); } } while (0)
    double* d_raw; float* d_verts; unsigned* d_ext;
    TA(d_raw, sizeof(double) * 3 * (size_t)n_verts); TA(d_verts, sizeof(float) * 3 * (size_t)n_verts); TA(d_ext, sizeof(unsigned) * 8);
    const unsigned ext_init[8] = {0xffffffffu, 0xffffffffu, 0xffffffffu, 0u, 0u, 0u, 0u, 0u};
    TC(cudaMemcpyAsync(d_raw, h_raw_verts, sizeof(double) * 3 * (size_t)n_verts, cudaMemcpyHostToDevice, st));
    TC(cudaMemcpyAsync(d_ext, ext_init, sizeof(ext_init), cudaMemcpyHostToDevice, st));
    const int T = 256, GV = (n_verts + T - 1) / T;
    ing::k_transform<<<GV, T, 0, st>>>(d_raw, n_verts, f3(translate[0], translate[1], translate[2]), scale, d_verts, d_ext);
    unsigned ext[8];
    TC(cudaMemcpyAsync(ext, d_ext, sizeof(ext), cudaMemcpyDeviceToHost, st));
    TC(cudaMemcpyAsync(h_verts, d_verts, sizeof(float) * 3 * (size_t)n_verts, cudaMemcpyDeviceToHost, st));
    TC(cudaStreamSynchronize(st));
    float mn[3], mx[3], cen[3];
    for (int k = 0; k < 3; k++) {
        mn[k] = ing::ord2f(ext[k]); mx[k] = ing::ord2f(ext[3 + k]);
        cen[k] = (float)(((double)mx[k] + (double)mn[k]) / 2.0);                 // mesh.py:36-41 (Python floats), stored as a vec3
    }
    ing::k_max_dist<<<GV, T, 0, st>>>(d_verts, n_verts, f3(cen[0], cen[1], cen[2]), d_ext + 6);
    TC(cudaMemcpyAsync(ext + 6, d_ext + 6, sizeof(unsigned), cudaMemcpyDeviceToHost, st));
    if (want_normals) {
        const int n_pairs = 3 * n_faces;
        int* d_faces; float4* d_weighted; float* d_vn; uint64_t* keys[2]; int* vals[2]; unsigned *hist, *scratch;
        const int n_tiles = std::max(1, (n_pairs + RS_TILE - 1) / RS_TILE);
        TA(d_faces, sizeof(int) * (size_t)std::max(1, n_pairs)); TA(d_weighted, sizeof(float4) * (size_t)std::max(1, n_faces));
        TA(d_vn, sizeof(float) * 3 * (size_t)n_verts);
        TA(keys[0], sizeof(uint64_t) * (size_t)std::max(1, n_pairs)); TA(keys[1], sizeof(uint64_t) * (size_t)std::max(1, n_pairs));
        TA(vals[0], sizeof(int) * (size_t)std::max(1, n_pairs)); TA(vals[1], sizeof(int) * (size_t)std::max(1, n_pairs));
        TA(hist, sizeof(unsigned) * 256 * (size_t)n_tiles);
        TA(scratch, sizeof(unsigned) * scan_scratch_elems(256ll * n_tiles));
        int cur = 0;
        if (n_faces > 0) {
            TC(cudaMemcpyAsync(d_faces, h_faces, sizeof(int) * (size_t)n_pairs, cudaMemcpyHostToDevice, st));
            ing::k_face_normals<<<(n_faces + T - 1) / T, T, 0, st>>>(d_verts, d_faces, n_faces, d_weighted, keys[0], vals[0]);
            int bits = 1;
            while (bits < 32 && (1u << bits) < (unsigned)n_verts) bits++;
            for (int shift = 0; shift < bits; shift += 8) {          // stable LSD passes over the digits a vertex index can occupy
                k_rs_hist<<<n_tiles, RS_THREADS, 0, st>>>(keys[cur], n_pairs, shift, n_tiles, hist);
                exclusive_scan_u32(hist, 256 * n_tiles, scratch, st);
                k_rs_scatter<<<n_tiles, RS_THREADS, 0, st>>>(keys[cur], vals[cur], n_pairs, shift, n_tiles, hist, keys[cur ^ 1], vals[cur ^ 1]);
                cur ^= 1;
            }
        }
        ing::k_vertex_normals<<<GV, T, 0, st>>>(keys[cur], vals[cur], n_pairs, d_weighted, n_verts, d_vn);
        TC(cudaMemcpyAsync(h_vnormals, d_vn, sizeof(float) * 3 * (size_t)n_verts, cudaMemcpyDeviceToHost, st));
    }
    TC(cudaStreamSynchronize(st));
    TC(cudaGetLastError());
#undef TA
#undef TC
    tfree();
    for (int k = 0; k < 3; k++) {
        if (h_extent_min) h_extent_min[k] = mn[k];
        if (h_extent_max) h_extent_max[k] = mx[k];
        if (h_center) h_center[k] = cen[k];
    }
    if (h_max_dist) *h_max_dist = ing::ord2f(ext[6]);
    return RT_OK;
}
