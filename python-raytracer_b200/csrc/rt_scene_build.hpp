// rt_scene_build.hpp - host-side conversion RtSceneDesc -> device record arrays and sampling tables.
// Pure host C++ (no CUDA calls): b200rt.cu uploads the vectors to HBM; tests/hostsim points a SceneView
// at them directly to exercise the same per-ray functions on the CPU.
#pragma once
#include <algorithm>
#include <cmath>
#include <cstring>
#include <string>
#include <vector>

#include "rt_bvh.cuh"
#include "rt_core.cuh"

struct HostScene {
    std::vector<DevObject> objs;
    std::vector<DevMaterial> mats;
    std::vector<DevLight> lights;
    std::vector<float> col_x, row_y, times;
    std::vector<float4> dof_origin, ray_origin;
    std::vector<BvhNode4> tlas_nodes;  // object BVH over the bounded top-level objects (empty: none)
    std::vector<int> lin_objs;         // top-level objects tested one by one (no finite bounds)
    std::vector<int> sep_mesh;         // top-level meshes, traversed after the walk over the other objects
    std::vector<float4> child_lo, child_hi;   // per object: padded bounds in the parent's space (SceneView::child_lo/hi)
    bool has_nodes = false, has_specular = false, has_mesh = false;
    SceneView view{};          // scalars filled; pointers are the caller's business
};

namespace rtb {
inline float3 h3(const float* p) { return f3(p[0], p[1], p[2]); }
inline float4 h4(const float* p, float w = 0.0f) { return f4(p[0], p[1], p[2], w); }

// scene.py:118-138: golden-angle disc points in WORLD x/y around `origin` (float32 results)
inline void sunflower(int n, float3 origin, double radius, std::vector<float4>& out) {
    const double phi = (1.0 + std::sqrt(5.0)) / 2.0;
    const double stride = 2.0 * 3.14159265358979323846 / phi;
    for (int k = 1; k <= n; k++) {
        double r = radius * std::sqrt(k - 0.5) / std::sqrt(n - 0.5);
        double th = k * stride;
        out.push_back(f4((float)(r * std::cos(th) + (double)origin.x), (float)(r * std::sin(th) + (double)origin.y), origin.z, 0.0f));
    }
}

inline int nest_depth(const RtSceneDesc* d, int oi, int depth) {
    const RtObject& o = d->objects[oi];
    if (o.kind != RT_KIND_NODE) return depth;
    int m = depth + 1;
    for (int k = 0; k < o.n_children; k++) m = std::max(m, nest_depth(d, o.first_child + k, depth + 1));
    return m;
}

inline void mesh_bounds(const RtMesh& m, float3& lo, float3& hi, float& pad);

// ---- object BVH (host-built: a scene has tens..thousands of objects; the per-mesh LBVH is the GPU-built one) ----
// Conservative world-space bounds of everything an object can return from intersect()/shadow_intersect()
// over the frame's motion-time range: every accepted hit point lies ON a primitive's surface.
struct ObjBound { float3 lo, hi; bool finite; };
inline float3 min3(float3 a, float3 b) { return f3(fminf(a.x, b.x), fminf(a.y, b.y), fminf(a.z, b.z)); }
inline float3 max3(float3 a, float3 b) { return f3(fmaxf(a.x, b.x), fmaxf(a.y, b.y), fmaxf(a.z, b.z)); }
inline bool finite3(float3 a) { return std::isfinite(a.x) && std::isfinite(a.y) && std::isfinite(a.z); }

inline float half_area(float3 lo, float3 hi) { float3 e = hi - lo; return e.x * e.y + e.y * e.z + e.z * e.x; }

inline ObjBound object_bound(const RtSceneDesc* d, int oi, float tlo, float thi) {
    const RtObject& o = d->objects[oi];
    ObjBound b; b.finite = false; b.lo = f3(INFINITY, INFINITY, INFINITY); b.hi = f3(-INFINITY, -INFINITY, -INFINITY);
    const float3 sp = h3(o.speed);
    const float ts[2] = {tlo, thi};
    switch (o.kind) {
        case RT_KIND_SPHERE:                                   // centre moves linearly (simple_geometry.py:21-24)
            for (int k = 0; k < 2; k++) {
                float3 c = h3(o.a) + sp * ts[k]; float r = fabsf(o.radius);
                b.lo = min3(b.lo, c - f3(r, r, r)); b.hi = max3(b.hi, c + f3(r, r, r));
            }
            b.finite = true; break;
        case RT_KIND_BOX:                                      // corners sorted like AAInterval (helperclasses.py:64-65)
            for (int k = 0; k < 2; k++) {
                float3 p = h3(o.a) + sp * ts[k], q = h3(o.b) + sp * ts[k];
                b.lo = min3(b.lo, min3(p, q)); b.hi = max3(b.hi, max3(p, q));
            }
            b.finite = true; break;
        case RT_KIND_MESH: {                                   // meshes do not move unless the mesh_motion extension is on
            const RtMesh& m = d->meshes[o.mesh];
            if (m.n_faces <= 0 || m.n_verts <= 0) break;
            float pad; float3 lo, hi; mesh_bounds(m, lo, hi, pad);
            for (int k = 0; k < 2; k++) {
                const float3 sh = d->mesh_motion ? sp * ts[k] : f3(0, 0, 0);
                b.lo = min3(b.lo, lo + sh); b.hi = max3(b.hi, hi + sh);
            }
            b.finite = true; break;
        }
        case RT_KIND_NODE: {                                   // hierarchy.py:42-78: hits of the children, mapped by M
            const int c0 = o.first_child, nc = o.n_children;
            float3 lo = b.lo, hi = b.hi;
            if (o.csg == RT_CSG_INTERSECTION) {
                // hierarchy.py:52-57: a hit of child k is kept iff every OTHER child contains it, so every kept hit lies on child
                // k's surface and inside every other child: within the bound of EACH child.  hierarchy.py:87-93: the node shadows
                // iff ALL children do, so a shadowing ray touches each child's bound (not necessarily at one common point: the
                // intersection of the bounds would be wrong there).  Any ONE finite child bound is therefore valid for both
                // queries; the smallest is taken, and unbounded children (planes) do not make the node unbounded:
                // `intersection(plane, box)` - the walls of scenes/NovelScene1.json:83-160 - is bounded by its box.
                float best = INFINITY; bool any = false;
                for (int k = 0; k < nc; k++) {
                    ObjBound cb = object_bound(d, c0 + k, tlo, thi);
                    if (!cb.finite) continue;
                    const float a = half_area(cb.lo, cb.hi);
                    if (!any || a < best) { best = a; lo = cb.lo; hi = cb.hi; any = true; }
                }
                if (!any) break;                                   // also nc == 0: all([]) hits nothing that has an extent
            } else {
                int use = 0;
                if (o.csg == RT_CSG_UNION) use = nc;
                else if (o.csg == RT_CSG_DIFFERENCE && nc >= 2) use = 1;   // A-hits, and B-hits inside A (is_inside(A) => within A's bound)
                if (use == 0) break;
                for (int k = 0; k < use; k++) {
                    ObjBound cb = object_bound(d, c0 + k, tlo, thi);
                    if (!cb.finite) return b;
                    lo = min3(lo, cb.lo); hi = max3(hi, cb.hi);
                }
            }
            const float* M = d->xforms[o.xform].M;
            for (int c = 0; c < 8; c++) {
                float3 p = f3((c & 1) ? hi.x : lo.x, (c & 2) ? hi.y : lo.y, (c & 4) ? hi.z : lo.z);
                float3 q = xform_point(M, p);
                b.lo = min3(b.lo, q); b.hi = max3(b.hi, q);
            }
            b.finite = true; break;
        }
        default: break;                                        // plane: unbounded
    }
    if (b.finite && !(finite3(b.lo) && finite3(b.hi))) b.finite = false;
    return b;
}

struct TlasItem { float3 lo, hi; int obj; int flags; };

// binary tree under construction (same shape as the GPU LBVH builder's arrays, so the 4-wide collapse is shared)
struct BinTree {
    std::vector<int2> children; std::vector<float4> node_lo, node_hi; std::vector<int> node_flags, depth;
    std::vector<float4> leaf_lo, leaf_hi; std::vector<int> leaf_flags;     // indexed by object index
};

// top-down sweep SAH; returns the child link of the subtree over items [b, e) (>= 0 node index, < 0 ~object)
inline int tlas_build(std::vector<TlasItem>& it, int b, int e, BinTree& bt, int depth, float3& lo, float3& hi, int& flags) {
    if (e - b == 1) { lo = it[b].lo; hi = it[b].hi; flags = it[b].flags; return ~it[b].obj; }
    const int idx = (int)bt.children.size();
    bt.children.emplace_back(); bt.node_lo.emplace_back(); bt.node_hi.emplace_back(); bt.node_flags.push_back(0); bt.depth.push_back(depth);
    const int n = e - b;
    int best_axis = 0, best_split = n / 2; float best_cost = INFINITY;
    std::vector<float> right(n);
    for (int ax = 0; ax < 3; ax++) {
        std::stable_sort(it.begin() + b, it.begin() + e, [ax](const TlasItem& p, const TlasItem& q) {
            return comp(p.lo, ax) + comp(p.hi, ax) < comp(q.lo, ax) + comp(q.hi, ax); });
        float3 l = f3(INFINITY, INFINITY, INFINITY), h = f3(-INFINITY, -INFINITY, -INFINITY);
        for (int k = n - 1; k >= 1; k--) { l = min3(l, it[b + k].lo); h = max3(h, it[b + k].hi); right[k] = half_area(l, h) * (n - k); }
        l = f3(INFINITY, INFINITY, INFINITY); h = f3(-INFINITY, -INFINITY, -INFINITY);
        for (int k = 1; k < n; k++) {
            l = min3(l, it[b + k - 1].lo); h = max3(h, it[b + k - 1].hi);
            float cost = half_area(l, h) * k + right[k];
            if (cost < best_cost) { best_cost = cost; best_axis = ax; best_split = k; }
        }
    }
    const int ax = best_axis;
    std::stable_sort(it.begin() + b, it.begin() + e, [ax](const TlasItem& p, const TlasItem& q) {
        return comp(p.lo, ax) + comp(p.hi, ax) < comp(q.lo, ax) + comp(q.hi, ax); });
    float3 l0, h0, l1, h1; int f0, f1;
    const int c0 = tlas_build(it, b, b + best_split, bt, depth + 1, l0, h0, f0);
    const int c1 = tlas_build(it, b + best_split, e, bt, depth + 1, l1, h1, f1);
    lo = min3(l0, l1); hi = max3(h0, h1); flags = f0 | f1;
    bt.children[idx].x = c0; bt.children[idx].y = c1;
    bt.node_lo[idx] = f4(lo.x, lo.y, lo.z, 0.0f); bt.node_hi[idx] = f4(hi.x, hi.y, hi.z, 0.0f); bt.node_flags[idx] = flags;
    return idx;
}

inline float bound_pad(const ObjBound& b) {                  // fp32 slack of the exact tests (hit points a few ulp off the surface)
    const float m = fmaxf(fmaxf(fabsf(b.lo.x), fabsf(b.hi.x)), fmaxf(fmaxf(fabsf(b.lo.y), fabsf(b.hi.y)), fmaxf(fabsf(b.lo.z), fabsf(b.hi.z))));
    return 1e-4f * m + 1e-5f;
}

// splits the top-level objects into the object BVH (finite bounds) and the linear list (the rest, scene order)
inline void build_tlas(const RtSceneDesc* d, HostScene& hs) {
    hs.tlas_nodes.clear(); hs.lin_objs.clear();
    float tlo = INFINITY, thi = -INFINITY;
    for (int t = 0; t < d->n_times; t++) { float v = (float)d->motion_times[t]; tlo = fminf(tlo, v); thi = fmaxf(thi, v); }
    // bounds of the nodes' children, each in its parent's space
    hs.child_lo.assign((size_t)d->n_objects, f4(-INFINITY, -INFINITY, -INFINITY, 0.0f));
    hs.child_hi.assign((size_t)d->n_objects, f4(INFINITY, INFINITY, INFINITY, 0.0f));
    for (int i = d->n_top; i < d->n_objects && std::isfinite(tlo) && std::isfinite(thi); i++) {
        const ObjBound b = object_bound(d, i, tlo, thi);
        if (!b.finite || !finite3(b.lo) || !finite3(b.hi)) continue;
        const float pad = bound_pad(b);
        hs.child_lo[i] = f4(b.lo.x - pad, b.lo.y - pad, b.lo.z - pad, 0.0f);
        hs.child_hi[i] = f4(b.hi.x + pad, b.hi.y + pad, b.hi.z + pad, 0.0f);
    }
    hs.sep_mesh.clear();
    std::vector<TlasItem> items;
    for (int i = 0; i < d->n_top; i++) {
        if (d->objects[i].kind == RT_KIND_MESH) { hs.sep_mesh.push_back(i); continue; }
        ObjBound b = (std::isfinite(tlo) && std::isfinite(thi)) ? object_bound(d, i, tlo, thi) : ObjBound{f3(0, 0, 0), f3(0, 0, 0), false};
        if (!b.finite) { hs.lin_objs.push_back(i); continue; }
        const float pad = bound_pad(b);
        TlasItem ti; ti.lo = b.lo - f3(pad, pad, pad); ti.hi = b.hi + f3(pad, pad, pad); ti.obj = i;
        const int kind = d->objects[i].kind;                  // mesh.py:121-153 and hierarchy.py:94-107 ignore t_max
        ti.flags = (kind == RT_KIND_MESH || kind == RT_KIND_NODE) ? RT_TLAS_NO_TMAX : 0;
        items.push_back(ti);
    }
    if (items.size() < 2) {                                   // nothing to gain: plain loop in scene order over everything but the separate meshes
        hs.lin_objs.clear();
        for (int i = 0; i < d->n_top; i++)
            if (d->objects[i].kind != RT_KIND_MESH) hs.lin_objs.push_back(i);
        return;
    }
    BinTree bt;
    bt.leaf_lo.assign(d->n_top, f4(0, 0, 0, 0)); bt.leaf_hi.assign(d->n_top, f4(0, 0, 0, 0)); bt.leaf_flags.assign(d->n_top, 0);
    for (const TlasItem& ti : items) {
        bt.leaf_lo[ti.obj] = f4(ti.lo.x, ti.lo.y, ti.lo.z, 0.0f); bt.leaf_hi[ti.obj] = f4(ti.hi.x, ti.hi.y, ti.hi.z, 0.0f);
        bt.leaf_flags[ti.obj] = ti.flags;
    }
    float3 lo, hi; int flags;
    tlas_build(items, 0, (int)items.size(), bt, 0, lo, hi, flags);
    // 4-wide collapse (rt_bvh.cuh: largest-area child opened first), nodes numbered in index order
    const int nb = (int)bt.children.size();
    std::vector<unsigned> remap(nb, 0), wide(nb, 0);
    wide_mark_host(bt.children.data(), bt.node_lo.data(), bt.node_hi.data(), nb, 1, wide.data());
    unsigned n4 = 0;
    for (int i = 0; i < nb; i++) if (wide[i]) remap[i] = n4++;
    hs.tlas_nodes.resize(n4);
    for (int i = 0; i < nb; i++)
        if (wide[i])
            hs.tlas_nodes[remap[i]] = make_wide_node(bt.children.data(), i, remap.data(), bt.leaf_lo.data(), bt.leaf_hi.data(),
                                                     bt.node_lo.data(), bt.node_hi.data(), bt.leaf_flags.data(), bt.node_flags.data(), 0.0f);
}

// returns RT_OK or an RtStatus with `err` set
inline int convert_scene(const RtSceneDesc* d, HostScene& hs, std::string& err) {
    auto bad = [&](int code, const char* m) { err = m; return code; };
    if (d->width <= 0 || d->height <= 0 || d->dof_samples <= 0 || d->aa_samples <= 0 || d->n_times <= 0)
        return bad(RT_ERR_INVALID, "resolution and sample counts must be positive");
    for (int m = 0; m < d->n_meshes; m++)
        if (d->meshes[m].n_faces >= (1 << RT_LEAF_SLOT_BITS)) return bad(RT_ERR_LIMIT, "a mesh may hold fewer than 2^28 faces (leaf link encoding)");
    if (d->max_depth < 0 || d->max_depth > 255) return bad(RT_ERR_LIMIT, "max_depth must be in [0, 255]");
    if (d->n_materials > 4093 || d->n_objects >= (1 << 20)) return bad(RT_ERR_LIMIT, "at most 4093 materials and 2^20 - 1 objects (packed hit record)");
    hs.objs.resize(d->n_objects);
    for (int i = 0; i < d->n_objects; i++) {
        const RtObject& o = d->objects[i];
        DevObject& g = hs.objs[i];
        memset(&g, 0, sizeof g);
        if (o.kind < 0 || o.kind > RT_KIND_NODE) return bad(RT_ERR_INVALID, "object kind out of range");
        if (o.kind == RT_KIND_NODE && (o.xform < 0 || o.xform >= d->n_xforms || o.first_child < 0 || o.n_children < 0 ||
                                       o.first_child + o.n_children > d->n_objects))
            return bad(RT_ERR_INVALID, "node child range / xform index out of range");
        if (o.kind == RT_KIND_MESH && (o.mesh < 0 || o.mesh >= d->n_meshes)) return bad(RT_ERR_INVALID, "mesh index out of range");
        if (o.tex >= d->n_textures) return bad(RT_ERR_INVALID, "texture index out of range");
        if ((o.n_mats > 0 && (o.mat0 < 0 || o.mat0 >= d->n_materials)) || (o.n_mats > 1 && (o.mat1 < 0 || o.mat1 >= d->n_materials)))
            return bad(RT_ERR_INVALID, "material index out of range");
        g.a = h4(o.a, o.radius); g.b = h4(o.b); g.speed = h4(o.speed);
        g.kind = o.kind; g.n_mats = o.n_mats; g.mat0 = o.mat0; g.mat1 = o.mat1;
        g.tex = o.tex; g.mesh = o.mesh; g.first_child = o.first_child; g.n_children = o.n_children;
        g.csg = o.csg; g.xform = o.xform;
        g.ax_w = h4(o.ax_w); g.ax_h = h4(o.ax_h);
        g.tex_scale = o.tex_scale;
        if (o.kind == RT_KIND_NODE) hs.has_nodes = true;
        if (o.kind == RT_KIND_MESH) hs.has_mesh = true;
    }
    // moving meshes (extension) are handled by the GENERAL kernels only (the HAS_NODES instantiations: CSG / transform nodes), so
    // that the lean kernels of node-free scenes carry neither the code nor the registers for it
    if (d->mesh_motion && hs.has_mesh) hs.has_nodes = true;
    if (d->n_top < 0 || d->n_top > d->n_objects) return bad(RT_ERR_INVALID, "n_top out of range");
    for (int i = 0; i < d->n_top; i++)
        if (nest_depth(d, i, 0) > RT_MAX_NEST) return bad(RT_ERR_LIMIT, "CSG/transform nesting deeper than RT_MAX_NEST");
    hs.mats.resize(d->n_materials);
    for (int i = 0; i < d->n_materials; i++) {
        const RtMaterial& m = d->materials[i];
        hs.mats[i].diffuse = h4(m.diffuse, (float)m.hardness);
        hs.mats[i].specular = h4(m.specular, (float)m.tint);
        hs.mats[i].refr_index = (float)m.refr_index;
        hs.mats[i].inv_refr = (float)(1.0 / m.refr_index);          // scene.py:194
        hs.mats[i].one_minus_tint = (float)(1.0 - m.tint);          // scene.py:104
        hs.mats[i].type = m.type;
        if (m.type != RT_MAT_DIFFUSE) hs.has_specular = true;
    }
    hs.lights.resize(d->n_lights);
    for (int i = 0; i < d->n_lights; i++) {
        const RtLight& l = d->lights[i];
        hs.lights[i].vector = h4(l.vector, as_float(l.type));
        hs.lights[i].colour_power = f4(l.colour[0] * l.power, l.colour[1] * l.power, l.colour[2] * l.power, 0.0f);   // scene.py:183
        float3 nd = normalize(-h3(l.vector));                        // scene.py:173
        hs.lights[i].dir_norm = f4(nd.x, nd.y, nd.z, 0.0f);
    }
    // sampling tables (scene.py:41-43,57,60,64), accumulated in float64 exactly like the reference's loops
    const double dx = (d->right - d->left) / d->width, dy = (d->top - d->bottom) / d->height;
    hs.col_x.resize(d->width); hs.row_y.resize(d->height);
    { double x = d->left + 0.5 * dx; for (int c = 0; c < d->width; c++) { hs.col_x[c] = (float)x; x += dx; } }
    { double y = d->bottom + 0.5 * dy; for (int r = 0; r < d->height; r++) { hs.row_y[r] = (float)y; y += dy; } }
    sunflower(d->dof_samples, h3(d->cam_pos), d->aperture, hs.dof_origin);
    for (int k = 0; k < d->dof_samples; k++) sunflower(d->aa_samples, f3(hs.dof_origin[k]), 2 * (dx + dy), hs.ray_origin);
    hs.times.resize(d->n_times);
    for (int t = 0; t < d->n_times; t++) hs.times[t] = (float)d->motion_times[t];
    SceneView& V = hs.view;
    V.n_objects = d->n_objects; V.n_top = d->n_top; V.n_materials = d->n_materials; V.n_lights = d->n_lights;
    V.cam_pos = h3(d->cam_pos); V.cam_u = h3(d->cam_u); V.cam_v = h3(d->cam_v); V.cam_w = h3(d->cam_w);
    V.focal_length = (float)d->focal_length;
    V.noise_scale = (float)(0.1 * (dx + dy));
    V.width = d->width; V.height = d->height; V.D = d->dof_samples; V.S = d->aa_samples; V.T = d->n_times;
    V.jitter = d->jitter;
    V.seed_lo = (uint32_t)d->jitter_seed; V.seed_hi = (uint32_t)(d->jitter_seed >> 32);
    V.ambient = h3(d->ambient);
    V.max_depth = d->max_depth;
    V.mesh_mode = 0;
    build_tlas(d, hs);
    {   // padded union of everything the object BVH and the separate meshes can report (SceneView::world_lo/hi)
        float3 lo = f3(INFINITY, INFINITY, INFINITY), hi = f3(-INFINITY, -INFINITY, -INFINITY);
        bool any = false, ok = true;
        for (const BvhNode4& nd : hs.tlas_nodes) {
            const float *lx = &nd.lox.x, *ly = &nd.loy.x, *lz = &nd.loz.x, *hx = &nd.hix.x, *hy = &nd.hiy.x, *hz = &nd.hiz.x;
            const int* lk = &nd.link.x;
            for (int k = 0; k < 4; k++)
                if (lk[k] != RT_LINK_NONE) { lo = min3(lo, f3(lx[k], ly[k], lz[k])); hi = max3(hi, f3(hx[k], hy[k], hz[k])); any = true; }
        }
        for (int oi : hs.sep_mesh) {
            const RtMesh& m = d->meshes[d->objects[oi].mesh];
            if (m.n_faces <= 0) continue;
            float3 glo, ghi;
            if (m.gate == RT_GATE_SPHERE) { const float r = fabsf(m.gate_radius); glo = h3(m.gate_center) - f3(r, r, r); ghi = h3(m.gate_center) + f3(r, r, r); }
            else { glo = min3(h3(m.gate_min), h3(m.gate_max)); ghi = max3(h3(m.gate_min), h3(m.gate_max)); }
            if (d->mesh_motion)                                  // the gate volume moves with the mesh: sweep it over the motion times
                for (int t = 0; t < d->n_times; t++) {
                    const float3 sh = h3(d->objects[oi].speed) * (float)d->motion_times[t];
                    lo = min3(lo, glo + sh); hi = max3(hi, ghi + sh);
                }
            lo = min3(lo, glo); hi = max3(hi, ghi); any = true;
        }
        ok = any && finite3(lo) && finite3(hi);
        V.has_world = ok ? 1 : 0;
        if (ok) {
            ObjBound wb{lo, hi, true};
            const float pad = 4.0f * bound_pad(wb);
            V.world_lo = lo - f3(pad, pad, pad); V.world_hi = hi + f3(pad, pad, pad);
        } else { V.world_lo = f3(-INFINITY, -INFINITY, -INFINITY); V.world_hi = f3(INFINITY, INFINITY, INFINITY); }
    }
    V.tlas_flagged = 0;
    for (const BvhNode4& nd : hs.tlas_nodes) V.tlas_flagged |= (nd.flags.x | nd.flags.y | nd.flags.z | nd.flags.w) & RT_TLAS_NO_TMAX;
    V.mesh_motion = d->mesh_motion ? 1 : 0;
    V.any_motion = 0;
    for (int i = 0; i < d->n_objects; i++)
        if (d->objects[i].speed[0] != 0.0f || d->objects[i].speed[1] != 0.0f || d->objects[i].speed[2] != 0.0f) V.any_motion = 1;
    V.tlas_nodes = nullptr; V.lin_objs = nullptr; V.n_lin = (int)hs.lin_objs.size();
    V.sep_mesh = nullptr; V.n_sep_mesh = (int)hs.sep_mesh.size();
    // the parameter-bank copies (SceneConst); the mesh headers follow once the LBVHs exist (fill_const_meshes)
    SceneConst& c = V.c;
    memset(&c, 0, sizeof c);
    for (int i = 0; i < d->n_lights && i < RT_CONST_LIGHTS; i++) {
        c.light_vec[i] = hs.lights[i].vector; c.light_colour[i] = hs.lights[i].colour_power; c.light_dirn[i] = hs.lights[i].dir_norm;
    }
    c.n_lin_planes = -1;
    if ((int)hs.lin_objs.size() <= RT_CONST_LIN) {
        bool planes = true;
        for (int oi : hs.lin_objs) planes = planes && hs.objs[oi].kind == RT_KIND_PLANE;
        if (planes) {
            c.n_lin_planes = (int)hs.lin_objs.size();
            for (int k = 0; k < c.n_lin_planes; k++) {
                const DevObject& o = hs.objs[hs.lin_objs[k]];
                c.lin_obj[k] = hs.lin_objs[k]; c.lin_point[k] = o.a; c.lin_normal[k] = o.b; c.lin_speed[k] = o.speed;
            }
        }
    }
    // the lean kernels take the top-level mesh headers from the parameter bank: more meshes than it holds -> the general kernels
    if ((int)hs.sep_mesh.size() > RT_CONST_MESH) hs.has_nodes = true;
    c.n_sep_mesh = (int)hs.sep_mesh.size() <= RT_CONST_MESH ? (int)hs.sep_mesh.size() : -1;
    for (int k = 0; k < c.n_sep_mesh; k++) c.sep_obj[k] = hs.sep_mesh[k];
    return RT_OK;
}

// SceneConst::sep_mesh: copies of the top-level meshes' headers, once their pointers are final (device LBVH built / host arrays laid out)
inline void fill_const_meshes(SceneView& V, const std::vector<DevObject>& objs, const DevMesh* headers) {
    for (int k = 0; k < V.c.n_sep_mesh; k++) V.c.sep_mesh[k] = headers[objs[V.c.sep_obj[k]].mesh];
}

// mesh bounds + the padding used for the conservative LBVH boxes
inline void mesh_bounds(const RtMesh& m, float3& lo, float3& hi, float& pad) {
    lo = f3(INFINITY, INFINITY, INFINITY); hi = f3(-INFINITY, -INFINITY, -INFINITY);
    float maxabs = 0.0f;
    for (int k = 0; k < m.n_verts; k++) {
        float3 v = h3(m.verts + 3 * (size_t)k);
        lo = f3(fminf(lo.x, v.x), fminf(lo.y, v.y), fminf(lo.z, v.z));
        hi = f3(fmaxf(hi.x, v.x), fmaxf(hi.y, v.y), fmaxf(hi.z, v.z));
        maxabs = fmaxf(maxabs, fmaxf(fabsf(v.x), fmaxf(fabsf(v.y), fabsf(v.z))));
    }
    // the reference's inclusive edge tests accept points a few ulp outside the triangle: keep node boxes conservative
    pad = 1e-5f * maxabs + 1e-7f;
}
inline void mesh_header(const RtMesh& m, DevMesh& g) {
    memset(&g, 0, sizeof g);
    g.n_tris = m.n_faces; g.flat_shaded = m.flat_shaded; g.gate = m.gate; g.gate_radius = m.gate_radius;
    g.gate_min = h4(m.gate_min); g.gate_max = h4(m.gate_max); g.gate_center = h4(m.gate_center);
}

// RtRenderJob -> PixelMap (tile interleave across ranks, include/b200rt.h); false on a malformed job
inline bool make_pixel_map(const RtRenderJob* job, PixelMap& pm, long long& n_local_tiles) {
    if (!job || job->ncols < 0 || job->nrows < 0 || job->tile_w <= 0 || job->tile_h <= 0 || job->world <= 0 ||
        job->rank < 0 || job->rank >= job->world)
        return false;
    pm.col0 = job->col0; pm.ncols = job->ncols; pm.row0 = job->row0; pm.nrows = job->nrows;
    pm.tile_w = job->tile_w; pm.tile_h = job->tile_h; pm.rank = job->rank; pm.world = job->world;
    pm.n_tile_rows = std::max(1, (job->nrows + job->tile_h - 1) / job->tile_h);
    pm.tile_px = job->tile_w * job->tile_h;
    const long long n_tile_cols = (job->ncols + job->tile_w - 1) / job->tile_w;
    const long long n_tiles = (job->ncols == 0 || job->nrows == 0) ? 0 : n_tile_cols * pm.n_tile_rows;
    n_local_tiles = n_tiles > job->rank ? (n_tiles - job->rank + job->world - 1) / job->world : 0;
    pm.n_local_pixels = n_local_tiles * pm.tile_px;
    if (pm.n_local_pixels >= (1ll << 31) || n_tiles * (long long)std::max(1, job->world) >= (1ll << 31)) return false;   // 32-bit pixel slots
    pm.fd_tile_px = make_fastdiv((unsigned)pm.tile_px); pm.fd_tile_rows = make_fastdiv((unsigned)pm.n_tile_rows);
    pm.fd_tile_w = make_fastdiv((unsigned)pm.tile_w);
    pm.fd_N = pm.fd_T = pm.fd_S = make_fastdiv(1);
    return true;
}
}  // namespace rtb
