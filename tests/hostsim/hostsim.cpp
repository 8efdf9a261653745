// hostsim.cpp - TEST INFRASTRUCTURE ONLY: compiles the per-ray device functions of
// python-raytracer_b200/csrc (rt_core.cuh, rt_bvh.cuh: the RT_HD functions) with the HOST compiler so
// their logic can be checked against the oracle in the GPU-less dev container.  Never loaded by the
// product package; the kernels (rt_kernels.cuh) and the GPU LBVH builder kernels are not part of it.
#include <algorithm>
#include <cstdint>
#include <numeric>
#include <string>
#include <vector>

#include "../../python-raytracer_b200/csrc/rt_bvh.cuh"
#include "../../python-raytracer_b200/csrc/rt_core.cuh"
#include "../../python-raytracer_b200/csrc/rt_scene_build.hpp"

struct HostMesh {
    std::vector<float4> tri, vn;
    std::vector<int4> vidx;
    std::vector<BvhNode4> nodes;
    std::vector<int> order;
};

struct Sim {
    HostScene hs;
    std::vector<HostMesh> meshes;
    std::vector<DevMesh> mesh_hdr;
    std::vector<DevTexture> texs;
    SceneView V;
};

// bottom-up boxes + depths of the binary radix tree (sequential restatement of k_lbvh_leaves_refit / k_lbvh_even)
static void refit(const std::vector<int2>& ch, int c, int depth, const std::vector<float4>& leaf_lo, const std::vector<float4>& leaf_hi,
                  std::vector<float4>& nlo, std::vector<float4>& nhi, std::vector<int>& dep, float4& lo, float4& hi) {
    if (c < 0) { lo = leaf_lo[~c]; hi = leaf_hi[~c]; return; }
    float4 l0, h0, l1, h1;
    refit(ch, ch[c].x, depth + 1, leaf_lo, leaf_hi, nlo, nhi, dep, l0, h0);
    refit(ch, ch[c].y, depth + 1, leaf_lo, leaf_hi, nlo, nhi, dep, l1, h1);
    const int f0 = ch[c].x < 0 ? ~ch[c].x : as_int(l0.w), f1 = ch[c].y < 0 ? ~ch[c].y : as_int(l1.w);
    const int n0 = ch[c].x < 0 ? 1 : as_int(h0.w), n1 = ch[c].y < 0 ? 1 : as_int(h1.w);
    lo = f4(fminf(l0.x, l1.x), fminf(l0.y, l1.y), fminf(l0.z, l1.z), as_float(std::min(f0, f1)));
    hi = f4(fmaxf(h0.x, h1.x), fmaxf(h0.y, h1.y), fmaxf(h0.z, h1.z), as_float(n0 + n1));
    nlo[c] = lo; nhi[c] = hi; dep[c] = depth;
}

static void build_mesh(const RtMesh& m, HostMesh& hm, DevMesh& g) {
    const int n = m.n_faces;
    float3 lo, hi; float pad;
    rtb::mesh_bounds(m, lo, hi, pad);
    rtb::mesh_header(m, g);
    float3 ext = hi - lo;
    float3 inv_ext = f3(ext.x > 0 ? 1.0f / ext.x : 0.0f, ext.y > 0 ? 1.0f / ext.y : 0.0f, ext.z > 0 ? 1.0f / ext.z : 0.0f);
    std::vector<uint64_t> keys(n);
    hm.order.resize(n);
    for (int i = 0; i < n; i++) {
        float3 a = vert_at(m.verts, m.faces[3 * i]), b = vert_at(m.verts, m.faces[3 * i + 1]), c = vert_at(m.verts, m.faces[3 * i + 2]);
        keys[i] = morton63(tri_centroid(a, b, c), lo, inv_ext);
        hm.order[i] = i;
    }
    std::stable_sort(hm.order.begin(), hm.order.end(), [&](int a, int b) { return keys[a] < keys[b]; });
    std::vector<uint64_t> sk(n);
    for (int s = 0; s < n; s++) sk[s] = keys[hm.order[s]];
    auto tri_box = [&](int f, float4& blo, float4& bhi) {
        float3 a = vert_at(m.verts, m.faces[3 * f]), b = vert_at(m.verts, m.faces[3 * f + 1]), d = vert_at(m.verts, m.faces[3 * f + 2]);
        blo = f4(fminf(a.x, fminf(b.x, d.x)), fminf(a.y, fminf(b.y, d.y)), fminf(a.z, fminf(b.z, d.z)), 0.0f);
        bhi = f4(fmaxf(a.x, fmaxf(b.x, d.x)), fmaxf(a.y, fmaxf(b.y, d.y)), fmaxf(a.z, fmaxf(b.z, d.z)), 0.0f);
    };
    std::vector<int2> ch(std::max(n - 1, 0));
    std::vector<float4> leaf_lo(n), leaf_hi(n), nlo(std::max(n - 1, 0)), nhi(std::max(n - 1, 0));
    const bool ploc = RT_PLOC_RADIUS > 0 && n >= 2;
    if (ploc) {                                     // PLOC: the tree decides the triangle slots (depth-first order), not the sort alone
        std::vector<float4> mlo(n), mhi(n);
        for (int p = 0; p < n; p++) tri_box(hm.order[p], mlo[p], mhi[p]);
        std::vector<int> slot_of(n), order2(n);
        ploc_build_host(n, mlo.data(), mhi.data(), RT_PLOC_RADIUS, ch.data(), nlo.data(), nhi.data(), slot_of.data());
        for (int p = 0; p < n; p++) order2[slot_of[p]] = hm.order[p];
        hm.order.swap(order2);
    }
    hm.tri.resize(3 * (size_t)n); hm.vidx.resize(n);
    for (int s = 0; s < n; s++) {
        int f = hm.order[s];
        int i0 = m.faces[3 * f], i1 = m.faces[3 * f + 1], i2 = m.faces[3 * f + 2];
        float3 a = vert_at(m.verts, i0), b = vert_at(m.verts, i1), c = vert_at(m.verts, i2);
        hm.tri[3 * s] = f4(a.x, a.y, a.z, as_float(f));
        hm.tri[3 * s + 1] = f4(b.x, b.y, b.z, 0);
        hm.tri[3 * s + 2] = f4(c.x, c.y, c.z, 0);
        hm.vidx[s].x = i0; hm.vidx[s].y = i1; hm.vidx[s].z = i2; hm.vidx[s].w = f;
        tri_box(f, leaf_lo[s], leaf_hi[s]);
    }
    if (m.vnormals && !m.flat_shaded) {
        hm.vn.resize(m.n_verts);
        for (int k = 0; k < m.n_verts; k++) hm.vn[k] = rtb::h4(m.vnormals + 3 * (size_t)k);
    }
    if (n >= 2) {
        if (!ploc) {
            for (int i = 0; i < n - 1; i++) lbvh_node(sk.data(), n, i, ch[i].x, ch[i].y);
            std::vector<int> dep(n - 1, 0);
            float4 rl, rh;
            refit(ch, 0, 0, leaf_lo, leaf_hi, nlo, nhi, dep, rl, rh);
        }
        std::vector<unsigned> remap(n - 1, 0);
        unsigned n4 = 0;
        std::vector<unsigned> wide(n - 1, 0);
        wide_mark_host(ch.data(), nlo.data(), nhi.data(), n - 1, RT_LEAF_TRIS, wide.data());
        for (int i = 0; i < n - 1; i++) if (wide[i]) remap[i] = n4++;
        hm.nodes.resize(n4);
        for (int i = 0; i < n - 1; i++)
            if (wide[i])
                hm.nodes[remap[i]] = make_wide_node(ch.data(), i, remap.data(), leaf_lo.data(), leaf_hi.data(), nlo.data(), nhi.data(),
                                                    nullptr, nullptr, pad, RT_LEAF_TRIS);
    }
    g.tri = hm.tri.data(); g.nodes = n >= 2 ? hm.nodes.data() : nullptr; g.face_vidx = hm.vidx.data();
    g.vnormals = hm.vn.empty() ? nullptr : hm.vn.data();
}

static int make_sim(const RtSceneDesc* d, Sim& sim) {
    std::string err;
    int rc = rtb::convert_scene(d, sim.hs, err);
    if (rc) return rc;
    sim.meshes.resize(d->n_meshes); sim.mesh_hdr.resize(d->n_meshes);
    for (int i = 0; i < d->n_meshes; i++) build_mesh(d->meshes[i], sim.meshes[i], sim.mesh_hdr[i]);
    sim.texs.resize(d->n_textures);
    for (int i = 0; i < d->n_textures; i++) { sim.texs[i].rgb = d->textures[i].rgb; sim.texs[i].width = d->textures[i].width; sim.texs[i].height = d->textures[i].height; }
    SceneView& V = sim.V;
    V = sim.hs.view;
    V.objects = sim.hs.objs.data(); V.materials = sim.hs.mats.data(); V.lights = sim.hs.lights.data();
    V.xforms = (const DevXform*)d->xforms; V.meshes = sim.mesh_hdr.data(); V.textures = sim.texs.data();
    V.col_x = sim.hs.col_x.data(); V.row_y = sim.hs.row_y.data(); V.dof_origin = sim.hs.dof_origin.data();
    V.ray_origin = sim.hs.ray_origin.data(); V.times = sim.hs.times.data();
    V.tlas_nodes = sim.hs.tlas_nodes.empty() ? nullptr : sim.hs.tlas_nodes.data();
    V.lin_objs = sim.hs.lin_objs.data(); V.n_lin = (int)sim.hs.lin_objs.size();
    V.child_lo = sim.hs.has_nodes ? sim.hs.child_lo.data() : nullptr; V.child_hi = sim.hs.has_nodes ? sim.hs.child_hi.data() : nullptr;
    V.sep_mesh = sim.hs.sep_mesh.data(); V.n_sep_mesh = (int)sim.hs.sep_mesh.size();
    rtb::fill_const_meshes(V, sim.hs.objs, sim.mesh_hdr.data());
    return 0;
}

static void add_tally(RtCounters* c, const Tally& t, uint64_t primary) {
    if (!c) return;
    c->rays_primary += primary; c->rays_shadow += t.shadow; c->rays_reflect += t.reflect; c->rays_refract += t.refract;
    c->bvh_nodes += t.nodes; c->tri_tests += t.tris; c->prim_tests += t.prims; c->shaded_hits += t.shaded;
    c->mesh_hits_smooth += t.smooth; c->error_flags |= t.err;
}

template <int HN>
static void render_px(const Sim& sim, const SceneView& V, const int32_t* cols, const int32_t* rows, int64_t n, float* out, RtCounters* ctr) {
#pragma omp parallel
    {
        Tally tally{}; uint64_t primary = 0;
        int stack[RT_STACK_DEPTH], spill[RT_STACK_SPILL];
#pragma omp for schedule(dynamic, 16)
        for (int64_t i = 0; i < n; i++) {
            float3 colour = f3(0, 0, 0);
            for (int d = 0; d < V.D; d++)
                for (int s = 0; s < V.S; s++) {
                    Ray ray = generate_ray(V, cols[i], rows[i], d, s);
                    for (int t = 0; t < V.T; t++) {
                        PathState ps; ps.ray = ray; ps.fold = fold_identity(); ps.depth_left = V.max_depth; ps.in_shape = 0;
                        TraceCtx cx{V, V.times[t], tally, stack, 1, spill};
                        float3 sample;
                        primary++;
                        while (trace_level<HN>(cx, ps, sample)) {}
                        colour = colour + sample;
                    }
                }
            float nn = (float)(V.D * V.S * V.T);
            out[3 * i] = colour.x / nn; out[3 * i + 1] = colour.y / nn; out[3 * i + 2] = colour.z / nn;
        }
#pragma omp critical
        add_tally(ctr, tally, primary);
    }
}

extern "C" int hostsim_render_pixels(const RtSceneDesc* d, const int32_t* cols, const int32_t* rows, int64_t n, float* out,
                                     int mesh_mode, RtCounters* ctr) {
    Sim sim;
    int rc = make_sim(d, sim);
    if (rc) return rc;
    if (ctr) *ctr = RtCounters{};
    SceneView V = sim.V; V.mesh_mode = mesh_mode;
    if (sim.hs.has_nodes) render_px<RT_V_NODES_MESH>(sim, V, cols, rows, n, out, ctr); else render_px<RT_V_LEAN>(sim, V, cols, rows, n, out, ctr);
    return 0;
}

// the primary ray of (col, row, dof sample d, aa sample s): origin xyz, direction xyz
extern "C" int hostsim_primary_ray(const RtSceneDesc* d, int col, int row, int dd, int s, float* out6) {
    Sim sim;
    int rc = make_sim(d, sim);
    if (rc) return rc;
    Ray r = generate_ray(sim.V, col, row, dd, s);
    out6[0] = r.o.x; out6[1] = r.o.y; out6[2] = r.o.z; out6[3] = r.d.x; out6[4] = r.d.y; out6[5] = r.d.z;
    return 0;
}

// explicit rays through the same functions (the host twins of rt_trace_rays / rt_occluded)
extern "C" int hostsim_trace_rays(const RtSceneDesc* d, int64_t n, const float* org, const float* dir, const double* times,
                                  float* rgb, float* hit_t, int32_t* hit_geom, float* hit_np /* n xyz, p xyz, mat */, int mesh_mode) {
    Sim sim;
    int rc = make_sim(d, sim);
    if (rc) return rc;
    SceneView V = sim.V; V.mesh_mode = mesh_mode;
#pragma omp parallel for schedule(dynamic, 64)
    for (int64_t i = 0; i < n; i++) {
        int stack[RT_STACK_DEPTH], spill[RT_STACK_SPILL];
        Tally tally{};
        PathState ps;
        ps.ray.o = f3(org[3 * i], org[3 * i + 1], org[3 * i + 2]); ps.ray.d = f3(dir[3 * i], dir[3 * i + 1], dir[3 * i + 2]);
        TraceCtx cx{V, (float)times[i], tally, stack, 1, spill};
        Hit h;
        const bool hit = sim.hs.has_nodes ? closest_hit<RT_V_NODES_MESH>(cx, ps.ray, h) : closest_hit<RT_V_LEAN>(cx, ps.ray, h);
        hit_t[i] = hit ? h.t : INFINITY; hit_geom[i] = hit ? h.geom : -1;
        if (hit_np && hit) { float* q = hit_np + 7 * i; q[0] = h.n.x; q[1] = h.n.y; q[2] = h.n.z; q[3] = h.p.x; q[4] = h.p.y; q[5] = h.p.z; q[6] = (float)h.mat; }
        ps.fold = fold_identity(); ps.depth_left = V.max_depth; ps.in_shape = 0;
        float3 sample;
        while (sim.hs.has_nodes ? trace_level<RT_V_NODES_MESH>(cx, ps, sample) : trace_level<RT_V_LEAN>(cx, ps, sample)) {}
        rgb[3 * i] = sample.x; rgb[3 * i + 1] = sample.y; rgb[3 * i + 2] = sample.z;
    }
    return 0;
}

extern "C" int hostsim_occluded(const RtSceneDesc* d, int64_t n, const float* org, const float* dir, const double* times,
                                const double* tmax, int32_t* occ, int mesh_mode) {
    Sim sim;
    int rc = make_sim(d, sim);
    if (rc) return rc;
    SceneView V = sim.V; V.mesh_mode = mesh_mode;
#pragma omp parallel for schedule(dynamic, 64)
    for (int64_t i = 0; i < n; i++) {
        int stack[RT_STACK_DEPTH], spill[RT_STACK_SPILL];
        Tally tally{};
        Ray r; r.o = f3(org[3 * i], org[3 * i + 1], org[3 * i + 2]); r.d = f3(dir[3 * i], dir[3 * i + 1], dir[3 * i + 2]);
        TraceCtx cx{V, (float)times[i], tally, stack, 1, spill};
        occ[i] = (sim.hs.has_nodes ? occluded<RT_V_NODES_MESH>(cx, r, (float)tmax[i]) : occluded<RT_V_LEAN>(cx, r, (float)tmax[i])) ? 1 : 0;
    }
    return 0;
}

// diagnostics: LBVH nodes visited by the first cast_ray level (closest hit + its shadow rays) of every (pixel, sample)
// ray, to estimate how evenly the lanes of a warp are loaded (scripts/warp_balance.py)
extern "C" int hostsim_node_counts(const RtSceneDesc* d, const int32_t* cols, const int32_t* rows, int64_t n, uint32_t* out) {
    Sim sim;
    int rc = make_sim(d, sim);
    if (rc) return rc;
    const SceneView& V = sim.V;
    const int N = V.D * V.S * V.T;
#pragma omp parallel for schedule(dynamic, 16)
    for (int64_t i = 0; i < n; i++) {
        int stack[RT_STACK_DEPTH], spill[RT_STACK_SPILL];
        for (int slot = 0; slot < N; slot++) {
            const int t = slot % V.T, ds = slot / V.T, s = ds % V.S, dd = ds / V.S;
            Tally tally{};
            PathState ps; ps.ray = generate_ray(V, cols[i], rows[i], dd, s); ps.fold = fold_identity(); ps.depth_left = V.max_depth; ps.in_shape = 0;
            TraceCtx cx{V, V.times[t], tally, stack, 1, spill};
            float3 sample;
            if (sim.hs.has_nodes) trace_level<RT_V_NODES_MESH>(cx, ps, sample); else trace_level<RT_V_LEAN>(cx, ps, sample);
            out[i * N + slot] = tally.nodes;
        }
    }
    return 0;
}

// diagnostics: per (pixel, sample slot, light) of the FIRST level: kind of the hit object (-1: miss), occluded, mesh nodes visited,
// triangles tested, object-BVH nodes + primitive tests, of that one shadow ray (scripts/shadow_stats.py)
extern "C" int hostsim_shadow_stats(const RtSceneDesc* d, const int32_t* cols, const int32_t* rows, int64_t n, int32_t* out) {
    Sim sim;
    int rc = make_sim(d, sim);
    if (rc) return rc;
    const SceneView& V = sim.V;
    const int N = V.D * V.S * V.T, L = V.n_lights;
#pragma omp parallel for schedule(dynamic, 16)
    for (int64_t i = 0; i < n; i++) {
        int stack[RT_STACK_DEPTH], spill[RT_STACK_SPILL];
        for (int slot = 0; slot < N; slot++) {
            const int t = slot % V.T, ds = slot / V.T, s = ds % V.S, dd = ds / V.S;
            Tally tally{};
            Ray ray = generate_ray(V, cols[i], rows[i], dd, s);
            TraceCtx cx{V, V.times[t], tally, stack, 1, spill};
            Hit h;
            const bool hit = sim.hs.has_nodes ? closest_hit<RT_V_NODES_MESH>(cx, ray, h) : closest_hit<RT_V_LEAN>(cx, ray, h);
            for (int li = 0; li < L; li++) {
                int32_t* o = out + ((i * N + slot) * L + li) * 5;
                o[0] = hit ? V.objects[h.geom].kind : -1; o[1] = o[2] = o[3] = o[4] = 0;
                if (!hit) continue;
                Tally t2{};
                TraceCtx c2{V, V.times[t], t2, stack, 1, spill};
                o[1] = (sim.hs.has_nodes ? light_occluded<RT_V_NODES_MESH>(c2, h.p, li) : light_occluded<RT_V_LEAN>(c2, h.p, li)) ? 1 : 0;
                o[2] = (int)t2.nodes; o[3] = (int)t2.tris; o[4] = (int)t2.prims;
            }
        }
    }
    return 0;
}

// returns the number of 4-wide nodes (>= 0) or -(status); nodes_out holds up to n_faces-1 nodes of 32 floats
extern "C" int hostsim_bvh(const RtSceneDesc* d, int mesh, float* nodes_out, int32_t* order_out) {
    Sim sim;
    int rc = make_sim(d, sim);
    if (rc) return -rc;
    const HostMesh& hm = sim.meshes[mesh];
    if (nodes_out && !hm.nodes.empty()) memcpy(nodes_out, hm.nodes.data(), sizeof(BvhNode4) * hm.nodes.size());
    if (order_out) memcpy(order_out, hm.order.data(), sizeof(int) * hm.order.size());
    return (int)hm.nodes.size();
}

// pixel mapping of a render job: (col,row,valid) of every local pixel slot, in compact-buffer order
extern "C" long long hostsim_job_pixels(const RtRenderJob* job, int32_t* cols, int32_t* rows, int32_t* valid) {
    PixelMap pm; long long nt;
    if (!rtb::make_pixel_map(job, pm, nt)) return -1;
    if (cols)
        for (long long lp = 0; lp < pm.n_local_pixels; lp++) {
            int c, r;
            valid[lp] = pixel_of(pm, (unsigned)lp, c, r) ? 1 : 0;
            cols[lp] = c; rows[lp] = r;
        }
    return pm.n_local_pixels;
}
