/* rt_oracle.c - Tier-1 oracle: CPU restatement of the reference's render path in plain C.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing in the product package may import, link or execute this
 * file; only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs use it, as the checker / reported baseline, never as the thing shipped.
 *
 * It follows the reference function by function (citations: file:line under
 * /root/reference/provided), brute force and recursive exactly like the Python, and models
 * its mixed arithmetic: PyGLM vectors are float32 (every vector op rounds to float32, no
 * FMA contraction: compile with -ffp-contract=off), while everything PyGLM hands back to
 * Python (dot, length, components) continues in float64.
 *
 * Pinning: this restatement is checked against the committed tier-0 fixtures in
 * tests/golden/ (the UNMODIFIED reference run through oracle/shims, which itself
 * reproduces the author's 5 stored deterministic renders bit-exactly; see
 * oracle/make_golden.py and tests/test_oracle_golden.py).
 */
#define _GNU_SOURCE
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif
#include "../include/b200rt.h"

/* ---------------------------------------------------------------- float32 vector maths (glm shim semantics) */
typedef struct { float x, y, z; } v3;

static inline v3 V3(float x, float y, float z) { v3 r = {x, y, z}; return r; }
static inline v3 v3p(const float* p) { return V3(p[0], p[1], p[2]); }
static inline v3 vadd(v3 a, v3 b) { return V3(a.x + b.x, a.y + b.y, a.z + b.z); }
static inline v3 vsub(v3 a, v3 b) { return V3(a.x - b.x, a.y - b.y, a.z - b.z); }
static inline v3 vmul(v3 a, v3 b) { return V3(a.x * b.x, a.y * b.y, a.z * b.z); }
static inline v3 vscale(v3 a, float s) { return V3(a.x * s, a.y * s, a.z * s); }
static inline v3 vscaled(v3 a, double s) { return vscale(a, (float)s); } /* vec3 * python float */
static inline v3 vneg(v3 a) { return V3(-a.x, -a.y, -a.z); }
static inline float vdot(v3 a, v3 b) { float t = a.x * b.x + a.y * b.y; return t + a.z * b.z; }
static inline v3 vcross(v3 a, v3 b) {
    return V3(a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y);
}
static inline float vlength(v3 a) { return sqrtf(vdot(a, a)); }
static inline v3 vnormalize(v3 a) { float inv = 1.0f / sqrtf(vdot(a, a)); return vscale(a, inv); }
static inline int veq(v3 a, v3 b) { return a.x == b.x && a.y == b.y && a.z == b.z; }

/* GLSL reflect / refract as GLM implements them */
static inline v3 vreflect(v3 i, v3 n) {
    float d = vdot(n, i);
    v3 t = vscale(n, d);
    t = vscale(t, 2.0f);
    return vsub(i, t);
}
static inline v3 vrefract(v3 i, v3 n, float eta) {
    float d = vdot(n, i);
    float k = 1.0f - (eta * eta) * (1.0f - d * d);
    if (k >= 0.0f) {
        float s = eta * d + sqrtf(k);
        return vsub(vscale(i, eta), vscale(n, s));
    }
    return V3(0, 0, 0);
}

/* mat4 (column-major m[col*4+row]) * vec4, GLM generic order: (m0*x + m1*y) + (m2*z + m3*w) */
static inline void m4mulv4(const float* m, float x, float y, float z, float w, float* out) {
    for (int i = 0; i < 4; i++) {
        float a0 = m[0 * 4 + i] * x + m[1 * 4 + i] * y;
        float a1 = m[2 * 4 + i] * z + m[3 * 4 + i] * w;
        out[i] = a0 + a1;
    }
}
static inline v3 xform_point(const float* m, v3 p) { float o[4]; m4mulv4(m, p.x, p.y, p.z, 1.0f, o); return V3(o[0], o[1], o[2]); }
static inline v3 xform_dir(const float* m, v3 d) { float o[4]; m4mulv4(m, d.x, d.y, d.z, 0.0f, o); return V3(o[0], o[1], o[2]); }
/* hierarchy.py:76: normalize(transpose(Minv) * vec4(n, 0)).xyz  (vec4 normalised INCLUDING w) */
static inline v3 xform_normal(const float* minv, v3 n) {
    float t[16];
    for (int c = 0; c < 4; c++) for (int r = 0; r < 4; r++) t[c * 4 + r] = minv[r * 4 + c];
    float o[4];
    m4mulv4(t, n.x, n.y, n.z, 0.0f, o);
    float d0 = o[0] * o[0] + o[1] * o[1];
    float d1 = o[2] * o[2] + o[3] * o[3];
    float inv = 1.0f / sqrtf(d0 + d1);
    return V3(o[0] * inv, o[1] * inv, o[2] * inv);
}

/* ---------------------------------------------------------------- Philox4x32-10 (oracle/philox.py) */
static void philox4x32_10(uint32_t c[4], uint32_t k0, uint32_t k1) {
    for (int r = 0; r < 10; r++) {
        uint64_t p0 = (uint64_t)0xD2511F53u * c[0];
        uint64_t p1 = (uint64_t)0xCD9E8D57u * c[2];
        uint32_t n0 = (uint32_t)(p1 >> 32) ^ c[1] ^ k0;
        uint32_t n1 = (uint32_t)p1;
        uint32_t n2 = (uint32_t)(p0 >> 32) ^ c[3] ^ k1;
        uint32_t n3 = (uint32_t)p0;
        c[0] = n0; c[1] = n1; c[2] = n2; c[3] = n3;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
}

/* ---------------------------------------------------------------- records */
typedef struct { v3 o, d; } Ray;                                  /* helperclasses.py:13-16 */
typedef struct { double t; v3 n, p; int mat; int geom; } Hit;     /* geometry/__init__.py:15-22 */
typedef struct { Hit* v; int n, cap; } HitList;

static void hl_push(HitList* l, Hit h) {
    if (l->n == l->cap) {
        l->cap = l->cap ? l->cap * 2 : 8;
        l->v = (Hit*)realloc(l->v, sizeof(Hit) * (size_t)l->cap);
    }
    l->v[l->n++] = h;
}

typedef struct {
    const RtSceneDesc* s;
    double time;          /* scene.current_time, scene.py:68 */
    RtCounters c;
    /* stability probe (oracle_render_pixels_perturbed; NOT reference behaviour, 0 = off): every ray handed to cast_ray
     * and every closest-hit position is moved by a random vector of relative length `perturb` (a few float32 ulp) before
     * it is used, so that a pixel whose reference value hinges on sub-ulp coincidences shows up as unstable */
    double perturb;
    uint32_t pstate[4];
} Ctx;

#define EPSILON 1e-4          /* geometry/__init__.py:12 */
#define SHADOW_EPS 1e-4       /* geometry/__init__.py:39 */
#define SHADOW_EPS_SPHERE 1e-3 /* simple_geometry.py:13 */

static inline v3 moved(const RtObject* o, const float* base, double time) {
    /* `center + speed * current_time` (simple_geometry.py:21-24); speed None == zero speed */
    return vadd(v3p(base), vscaled(v3p(o->speed), time));
}

/* ---------------------------------------------------------------- Sphere, simple_geometry.py:20-80 */
static int sphere_abc(const RtObject* o, const Ray* r, double time, v3* center, double* a, double* b, double* c) {
    *center = moved(o, o->a, time);
    v3 oc = vsub(r->o, *center);
    *a = (double)vdot(r->d, r->d);
    *b = 2.0 * (double)vdot(r->d, oc);
    *c = (double)vdot(oc, oc) - o->radius_d * o->radius_d;
    return 0;
}
static void sphere_intersect(Ctx* cx, int oi, const Ray* r, HitList* out) {
    const RtObject* o = &cx->s->objects[oi];
    v3 center; double a, b, c;
    cx->c.prim_tests++;
    sphere_abc(o, r, cx->time, &center, &a, &b, &c);
    double disc = b * b - 4 * a * c;
    if (disc < 0) return;
    double sq = sqrt(disc);
    double ts[2] = {(-b - sq) / (2 * a), (-b + sq) / (2 * a)};
    for (int k = 0; k < 2; k++) {
        double t = ts[k];
        if (t > 0) {
            Hit h;
            h.t = t;
            h.p = vadd(r->o, vscaled(r->d, t));
            h.n = vnormalize(vsub(h.p, center));
            h.mat = o->n_mats > 0 ? o->mat0 : -2; /* -2: IndexError in the reference */
            h.geom = oi;
            hl_push(out, h);
        }
    }
}
static int sphere_shadow(Ctx* cx, int oi, const Ray* r, double tmax) {
    const RtObject* o = &cx->s->objects[oi];
    v3 center; double a, b, c;
    cx->c.prim_tests++;
    sphere_abc(o, r, cx->time, &center, &a, &b, &c);
    double disc = b * b - 4 * a * c;
    if (disc < 0) return 0;
    double sq = sqrt(disc);
    double t = (-b - sq) / (2 * a);
    if (SHADOW_EPS_SPHERE < t && t < tmax) return 1;
    t = (-b + sq) / (2 * a);
    if (SHADOW_EPS_SPHERE < t && t < tmax) return 1;
    return 0;
}
static int sphere_inside(Ctx* cx, int oi, v3 p) {
    const RtObject* o = &cx->s->objects[oi];
    v3 center = moved(o, o->a, cx->time);
    return (double)vlength(vsub(p, center)) < o->radius_d;
}

/* ---------------------------------------------------------------- Plane, simple_geometry.py:105-173 */
static int plane_get_material(Ctx* cx, int oi, v3 point) {
    const RtObject* o = &cx->s->objects[oi];
    v3 position = moved(o, o->a, cx->time);
    v3 n = v3p(o->b);
    if (o->n_mats == 1) return o->mat0;
    if (o->n_mats == 0) return -2;
    point = vsub(point, vscaled(n, (double)vdot(vsub(point, position), n)));
    double x = (double)vdot(vsub(point, position), v3p(o->ax_w));
    double z = (double)vdot(vsub(point, position), v3p(o->ax_h));
    double dx = floor((double)position.x - x);
    double dz = floor((double)position.z - z);
    double m = fmod(dx + dz, 2.0);
    if (m < 0) m += 2.0;
    return m == 0.0 ? o->mat0 : o->mat1;
}
static void plane_intersect(Ctx* cx, int oi, const Ray* r, HitList* out) {
    const RtObject* o = &cx->s->objects[oi];
    cx->c.prim_tests++;
    v3 point = moved(o, o->a, cx->time);
    v3 n = v3p(o->b);
    double denom = (double)vdot(r->d, n);
    if (fabs(denom) > EPSILON) {
        double t = (double)vdot(vsub(point, r->o), n) / denom;
        if (t >= 0) {
            Hit h;
            h.t = t;
            h.p = vadd(r->o, vscaled(r->d, t));
            h.n = n;
            h.mat = plane_get_material(cx, oi, h.p);
            h.geom = oi;
            hl_push(out, h);
        }
    }
}
static int plane_shadow(Ctx* cx, int oi, const Ray* r, double tmax) {
    const RtObject* o = &cx->s->objects[oi];
    cx->c.prim_tests++;
    v3 point = moved(o, o->a, cx->time);
    v3 n = v3p(o->b);
    double denom = (double)vdot(r->d, n);
    if (fabs(denom) > EPSILON) {
        double t = (double)vdot(vsub(point, r->o), n) / denom;
        return SHADOW_EPS < t && t < tmax;
    }
    return 0; /* None */
}
static double py_mod(double a, double m) { /* Python float % positive int */
    double r = fmod(a, m);
    if (r != 0.0 && r < 0.0) r += m;
    return r;
}
static v3 texel(const RtTexture* t, long i, long j) {
    if (i < 0) i = 0;                             /* reference would raise IndexError; clamp instead */
    if (j < 0) j = 0;
    if (i >= t->width) i = t->width - 1;
    if (j >= t->height) j = t->height - 1;
    const uint8_t* px = t->rgb + ((size_t)j * (size_t)t->width + (size_t)i) * 3;
    return V3((float)(px[0] / 255.0), (float)(px[1] / 255.0), (float)(px[2] / 255.0));
}
static v3 plane_get_diffuse(Ctx* cx, int oi, v3 point, int* err) {
    const RtObject* o = &cx->s->objects[oi];
    if (o->tex >= 0) {
        const RtTexture* tx = &cx->s->textures[o->tex];
        v3 position = moved(o, o->a, cx->time);
        v3 n = v3p(o->b);
        /* projection uses the UN-moved self.point, simple_geometry.py:163 */
        point = vsub(point, vscaled(n, (double)vdot(vsub(point, v3p(o->a)), n)));
        double fi = py_mod((double)vdot(vsub(point, position), v3p(o->ax_w)) * 1000.0 / o->tex_scale, (double)tx->width);
        double fj = py_mod((double)vdot(vsub(point, position), v3p(o->ax_h)) * 1000.0 / o->tex_scale, (double)tx->height);
        return texel(tx, (long)fi, (long)fj);
    }
    int m = plane_get_material(cx, oi, point);
    if (m < 0) { *err = 1; return V3(0, 0, 0); }
    return v3p(cx->s->materials[m].diffuse);
}

/* ---------------------------------------------------------------- AABB, simple_geometry.py:188-355 */
typedef struct { double start, end; int label; } Interval;      /* helperclasses.py:62-66 */
static inline Interval mk_interval(double t1, double t2, int label) {
    Interval iv; iv.start = t1 < t2 ? t1 : t2; iv.end = t1 > t2 ? t1 : t2; iv.label = label;
    /* Python min/max return the first argument on ties/NaN-free input; equal values are identical */
    return iv;
}
/* slab intervals shared by AABB.intersect, AABB.shadow_intersect, BoundingAABB.intersect.
 * returns 0 when a zero direction component puts the origin outside (early `return []`) */
static int slab_intervals(v3 mn, v3 mx, const Ray* r, Interval* entry, Interval* exit_) {
    Interval iv[3];
    const float o[3] = {r->o.x, r->o.y, r->o.z}, d[3] = {r->d.x, r->d.y, r->d.z};
    const float lo[3] = {mn.x, mn.y, mn.z}, hi[3] = {mx.x, mx.y, mx.z};
    for (int k = 0; k < 3; k++) {
        if (d[k] == 0) {
            if (!(lo[k] < o[k] && o[k] < hi[k])) return 0;
            iv[k] = mk_interval(-INFINITY, INFINITY, k);
        } else {
            double t1 = ((double)lo[k] - (double)o[k]) / (double)d[k];
            double t2 = ((double)hi[k] - (double)o[k]) / (double)d[k];
            iv[k] = mk_interval(t1, t2, k);
        }
    }
    /* max(..., key=start): first maximal ; min(..., key=end): first minimal */
    Interval e = iv[0];
    if (iv[1].start > e.start) e = iv[1];
    if (iv[2].start > e.start) e = iv[2];
    Interval x = iv[0];
    if (iv[1].end < x.end) x = iv[1];
    if (iv[2].end < x.end) x = iv[2];
    *entry = e; *exit_ = x;
    return 1;
}
static void box_bounds(Ctx* cx, const RtObject* o, v3* mn, v3* mx) {
    *mn = moved(o, o->a, cx->time);
    *mx = moved(o, o->b, cx->time);
}
static void box_intersect(Ctx* cx, int oi, const Ray* r, HitList* out) {
    const RtObject* o = &cx->s->objects[oi];
    cx->c.prim_tests++;
    v3 mn, mx; box_bounds(cx, o, &mn, &mx);
    Interval e, x;
    if (!slab_intervals(mn, mx, r, &e, &x)) return;
    if (e.start > x.end || e.start < 0) return;
    const float d[3] = {r->d.x, r->d.y, r->d.z};
    double ts[2] = {e.start, x.end};
    for (int k = 0; k < 2; k++) {
        float nn[3] = {0, 0, 0};
        if (d[e.label] < 0) nn[e.label] = 1.0f; else if (d[e.label] > 0) nn[e.label] = -1.0f;
        Hit h;
        h.t = ts[k];
        h.n = V3(nn[0], nn[1], nn[2]);
        h.p = vadd(r->o, vscaled(r->d, ts[k]));
        h.mat = o->n_mats > 0 ? o->mat0 : -2;
        h.geom = oi;
        hl_push(out, h);
    }
}
static int box_shadow(Ctx* cx, int oi, const Ray* r, double tmax) {
    const RtObject* o = &cx->s->objects[oi];
    cx->c.prim_tests++;
    v3 mn, mx; box_bounds(cx, o, &mn, &mx);
    Interval e, x;
    if (!slab_intervals(mn, mx, r, &e, &x)) return 0;
    if (e.start > x.end) return 0;
    return SHADOW_EPS < e.start && e.start < tmax;
}
static int box_inside(Ctx* cx, int oi, v3 p) {
    const RtObject* o = &cx->s->objects[oi];
    v3 mn, mx; box_bounds(cx, o, &mn, &mx);
    return mn.x < p.x && p.x < mx.x && mn.y < p.y && p.y < mx.y && mn.z < p.z && p.z < mx.z;
}
static v3 box_get_diffuse(Ctx* cx, int oi, v3 point, int* err) {
    const RtObject* o = &cx->s->objects[oi];
    if (o->tex >= 0) {
        const RtTexture* tx = &cx->s->textures[o->tex];
        v3 mn, mx; box_bounds(cx, o, &mn, &mx);
        double W = tx->width, Hh = tx->height;
        double x = ((double)point.x - mn.x) / ((double)mx.x - mn.x);
        double y = ((double)point.y - mn.y) / ((double)mx.y - mn.y);
        double z = ((double)point.z - mn.z) / ((double)mx.z - mn.z);
        double i, j;
        if (fabs((double)point.x - mn.x) < EPSILON) { i = z * W; j = (1 - y) * Hh; }
        else if (fabs((double)point.x - mx.x) < EPSILON) { i = (1 - z) * W; j = (1 - y) * Hh; }
        else if (fabs((double)point.y - mn.y) < EPSILON) { i = x * W; j = (1 - z) * Hh; }
        else if (fabs((double)point.y - mx.y) < EPSILON) { i = x * W; j = z * Hh; }
        else if (fabs((double)point.z - mn.z) < EPSILON) { i = (1 - x) * W; j = (1 - y) * Hh; }
        else if (fabs((double)point.z - mx.z) < EPSILON) { i = x * W; j = (1 - y) * Hh; }
        else { i = 0; j = 0; }
        i = fmin(fmax(0, i), W - 1);
        j = fmin(fmax(0, j), Hh - 1);
        return texel(tx, (long)i, (long)j);  /* PIL truncates float coordinates */
    }
    if (o->n_mats < 1) { *err = 1; return V3(0, 0, 0); }
    return v3p(cx->s->materials[o->mat0].diffuse);
}

/* ---------------------------------------------------------------- bounding volumes, bounding_volumes.py */
/* `shift`: (0,0,0) in the reference (meshes stand still, mesh.py:72-153).  EXTENSION (RtSceneDesc.mesh_motion, not reference
 * behaviour): speed * current_time, added to the bounding volume and to every vertex the way simple_geometry.py:21-24 moves
 * the primitives (float32 vec3 + vec3 * float). */
static inline v3 mesh_shift(const Ctx* cx, const RtObject* o) {
    return cx->s->mesh_motion ? vscaled(v3p(o->speed), cx->time) : V3(0, 0, 0);
}
static int gate_intersect(const RtMesh* m, const Ray* r, v3 shift) {
    if (m->gate == RT_GATE_SPHERE) {  /* bounding_volumes.py:18-37 */
        v3 center = vadd(v3p(m->gate_center), shift);
        v3 oc = vsub(r->o, center);
        double a = (double)vdot(r->d, r->d);
        double b = 2.0 * (double)vdot(r->d, oc);
        double c = (double)vdot(oc, oc) - (double)m->gate_radius * (double)m->gate_radius;
        double disc = b * b - 4 * a * c;
        if (disc < 0) return 0;
        double t = (-b - sqrt(disc)) / (2 * a);
        if (t > 0) return 1;
        t = (-b + sqrt(disc)) / (2 * a);
        return t > 0;
    }
    Interval e, x;   /* bounding_volumes.py:49-83 */
    if (!slab_intervals(vadd(v3p(m->gate_min), shift), vadd(v3p(m->gate_max), shift), r, &e, &x)) return 0;
    if (e.start > x.end || e.start < 0) return 0;
    return 1;
}

/* ---------------------------------------------------------------- Mesh, mesh.py:72-153 */
static void bary_f32(v3 p, v3 a, v3 b, v3 c, float* l) {   /* oracle/shims/igl.py */
    v3 v0 = vsub(b, a), v1 = vsub(c, a), v2 = vsub(p, a);
    float d00 = vdot(v0, v0), d01 = vdot(v0, v1), d11 = vdot(v1, v1), d20 = vdot(v2, v0), d21 = vdot(v2, v1);
    float denom = d00 * d11 - d01 * d01;
    l[1] = (d11 * d20 - d01 * d21) / denom;
    l[2] = (d00 * d21 - d01 * d20) / denom;
    l[0] = 1.0f - (l[1] + l[2]);
}
static void mesh_intersect(Ctx* cx, int oi, const Ray* r, HitList* out) {
    const RtObject* o = &cx->s->objects[oi];
    const RtMesh* m = &cx->s->meshes[o->mesh];
    cx->c.prim_tests++;
    const v3 shift = mesh_shift(cx, o);
    if (!gate_intersect(m, r, shift)) return;
    for (int f = 0; f < m->n_faces; f++) {
        const int32_t* fc = m->faces + 3 * (size_t)f;
        v3 v0 = v3p(m->verts + 3 * (size_t)fc[0]), v1 = v3p(m->verts + 3 * (size_t)fc[1]), v2 = v3p(m->verts + 3 * (size_t)fc[2]);
        if (cx->s->mesh_motion) { v0 = vadd(v0, shift); v1 = vadd(v1, shift); v2 = vadd(v2, shift); }
        cx->c.tri_tests++;
        v3 e1 = vsub(v1, v0), e2 = vsub(v2, v0);
        v3 normal = vnormalize(vcross(e1, e2));
        double denom = (double)vdot(r->d, normal);
        if (fabs(denom) < EPSILON) continue;
        double time = (double)vdot(vsub(v0, r->o), normal) / denom;
        if (time < 0) continue;
        v3 point = vadd(r->o, vscaled(r->d, time));
        float b0 = vdot(vcross(vsub(v1, v0), vsub(point, v0)), normal);
        float b1 = vdot(vcross(vsub(v2, v1), vsub(point, v1)), normal);
        float b2 = vdot(vcross(vsub(v0, v2), vsub(point, v2)), normal);
        if (b0 >= 0 && b1 >= 0 && b2 >= 0) {
            if (!m->flat_shaded) {
                float l[3];
                bary_f32(point, v0, v1, v2, l);
                v3 n0 = v3p(m->vnormals + 3 * (size_t)fc[0]), n1 = v3p(m->vnormals + 3 * (size_t)fc[1]), n2 = v3p(m->vnormals + 3 * (size_t)fc[2]);
                normal = vnormalize(vadd(vadd(vscale(n0, l[0]), vscale(n1, l[1])), vscale(n2, l[2])));
            }
            Hit h;
            h.t = time; h.n = normal; h.p = point;
            h.mat = o->n_mats > 0 ? o->mat0 : -2;
            h.geom = oi;
            hl_push(out, h);
        }
    }
}
static int mesh_shadow(Ctx* cx, int oi, const Ray* r, double tmax) {
    (void)tmax; /* ignored by the reference, mesh.py:121-153 */
    const RtObject* o = &cx->s->objects[oi];
    const RtMesh* m = &cx->s->meshes[o->mesh];
    cx->c.prim_tests++;
    const v3 shift = mesh_shift(cx, o);
    if (!gate_intersect(m, r, shift)) return 0;
    for (int f = 0; f < m->n_faces; f++) {
        const int32_t* fc = m->faces + 3 * (size_t)f;
        v3 v0 = v3p(m->verts + 3 * (size_t)fc[0]), v1 = v3p(m->verts + 3 * (size_t)fc[1]), v2 = v3p(m->verts + 3 * (size_t)fc[2]);
        if (cx->s->mesh_motion) { v0 = vadd(v0, shift); v1 = vadd(v1, shift); v2 = vadd(v2, shift); }
        cx->c.tri_tests++;
        v3 normal = vcross(vsub(v1, v0), vsub(v2, v0));   /* un-normalised */
        double denom = (double)vdot(r->d, normal);
        if (fabs(denom) < EPSILON) continue;
        double time = (double)vdot(vsub(v0, r->o), normal) / denom;
        if (time < SHADOW_EPS) continue;
        v3 point = vadd(r->o, vscaled(r->d, time));
        v3 c0 = vcross(vsub(v1, v0), vsub(point, v0));
        v3 c1 = vcross(vsub(v2, v1), vsub(point, v1));
        v3 c2 = vcross(vsub(v0, v2), vsub(point, v2));
        if (vdot(c0, normal) >= 0 && vdot(c1, normal) >= 0 && vdot(c2, normal) >= 0) return 1;
    }
    return 0;
}

/* ---------------------------------------------------------------- dispatch + Hierarchy, hierarchy.py:42-138 */
static void obj_intersect(Ctx* cx, int oi, const Ray* r, HitList* out);
static int obj_shadow(Ctx* cx, int oi, const Ray* r, double tmax);
static int obj_inside(Ctx* cx, int oi, v3 p);
static int obj_get_material(Ctx* cx, int oi, v3 p);

static Ray node_local_ray(const RtXform* xf, const Ray* r) {
    Ray m;
    m.o = xform_point(xf->Minv, r->o);
    m.d = xform_dir(xf->Minv, r->d);
    return m;
}
static void node_intersect(Ctx* cx, int oi, const Ray* r, HitList* out) {
    const RtObject* o = &cx->s->objects[oi];
    const RtXform* xf = &cx->s->xforms[o->xform];
    Ray mr = node_local_ray(xf, r);
    HitList acc = {0, 0, 0};
    int c0 = o->first_child, nc = o->n_children;
    if (o->csg == RT_CSG_UNION) {
        for (int k = 0; k < nc; k++) obj_intersect(cx, c0 + k, &mr, &acc);
    } else if (o->csg == RT_CSG_INTERSECTION) {
        for (int k = 0; k < nc; k++) {
            HitList ch = {0, 0, 0};
            obj_intersect(cx, c0 + k, &mr, &ch);
            for (int i = 0; i < ch.n; i++) {
                int all = 1;
                for (int j = 0; j < nc; j++) {     /* all([...]) builds the whole list: no short-circuit needed for the result */
                    if (j == k) continue;
                    if (!obj_inside(cx, c0 + j, ch.v[i].p)) { all = 0; }
                }
                if (all) hl_push(&acc, ch.v[i]);
            }
            free(ch.v);
        }
    } else if (o->csg == RT_CSG_DIFFERENCE && nc >= 2) {
        HitList h1 = {0, 0, 0}, h2 = {0, 0, 0};
        obj_intersect(cx, c0, &mr, &h1);
        obj_intersect(cx, c0 + 1, &mr, &h2);
        for (int i = 0; i < h1.n; i++)
            if (!obj_inside(cx, c0 + 1, h1.v[i].p)) hl_push(&acc, h1.v[i]);
        for (int i = 0; i < h2.n; i++)
            if (obj_inside(cx, c0, h2.v[i].p)) {
                Hit h = h2.v[i];
                h.mat = obj_get_material(cx, c0, h.p);
                h.n = vneg(h.n);
                hl_push(&acc, h);
            }
        free(h1.v); free(h2.v);
    }
    for (int i = 0; i < acc.n; i++) {
        Hit h = acc.v[i];
        if (h.mat == -1) h.mat = o->n_mats > 0 ? o->mat0 : -2;   /* `mat is None` */
        h.p = xform_point(xf->M, h.p);
        h.n = xform_normal(xf->Minv, h.n);
        hl_push(out, h);
    }
    free(acc.v);
}
static int node_shadow(Ctx* cx, int oi, const Ray* r, double tmax) {
    const RtObject* o = &cx->s->objects[oi];
    const RtXform* xf = &cx->s->xforms[o->xform];
    Ray mr = node_local_ray(xf, r);
    int c0 = o->first_child, nc = o->n_children;
    if (o->csg == RT_CSG_UNION) {
        for (int k = 0; k < nc; k++) if (obj_shadow(cx, c0 + k, &mr, tmax)) return 1;
        return 0;
    } else if (o->csg == RT_CSG_INTERSECTION) {
        for (int k = 0; k < nc; k++) if (!obj_shadow(cx, c0 + k, &mr, tmax)) return 0;
        return 1;
    } else if (o->csg == RT_CSG_DIFFERENCE && nc >= 2) {
        HitList h1 = {0, 0, 0}, h2 = {0, 0, 0};
        int res = 0;
        obj_intersect(cx, c0, &mr, &h1);
        obj_intersect(cx, c0 + 1, &mr, &h2);
        for (int i = 0; i < h1.n && !res; i++)
            if (h1.v[i].t > SHADOW_EPS && !obj_inside(cx, c0 + 1, h1.v[i].p)) res = 1;
        for (int i = 0; i < h2.n && !res; i++)
            if (h2.v[i].t > SHADOW_EPS && obj_inside(cx, c0, h2.v[i].p)) res = 1;
        free(h1.v); free(h2.v);
        return res;
    }
    return 0;
}
static int node_inside(Ctx* cx, int oi, v3 p) {
    const RtObject* o = &cx->s->objects[oi];
    const RtXform* xf = &cx->s->xforms[o->xform];
    v3 q = xform_point(xf->Minv, p);
    int c0 = o->first_child, nc = o->n_children;
    if (o->csg == RT_CSG_UNION) {
        for (int k = 0; k < nc; k++) if (obj_inside(cx, c0 + k, q)) return 1;
        return 0;
    } else if (o->csg == RT_CSG_INTERSECTION) {
        for (int k = 0; k < nc; k++) if (!obj_inside(cx, c0 + k, q)) return 0;
        return 1;
    } else if (o->csg == RT_CSG_DIFFERENCE && nc >= 2) {
        int a = obj_inside(cx, c0, q), b = obj_inside(cx, c0 + 1, q);
        return a && !b;
    }
    return 0;
}
static int node_get_material(Ctx* cx, int oi, v3 p) {
    const RtObject* o = &cx->s->objects[oi];
    const RtXform* xf = &cx->s->xforms[o->xform];
    v3 q = xform_point(xf->Minv, p);
    for (int k = 0; k < o->n_children; k++)
        if (obj_inside(cx, o->first_child + k, q)) return obj_get_material(cx, o->first_child + k, q);
    return -1; /* None */
}

static void obj_intersect(Ctx* cx, int oi, const Ray* r, HitList* out) {
    switch (cx->s->objects[oi].kind) {
        case RT_KIND_SPHERE: sphere_intersect(cx, oi, r, out); break;
        case RT_KIND_PLANE: plane_intersect(cx, oi, r, out); break;
        case RT_KIND_BOX: box_intersect(cx, oi, r, out); break;
        case RT_KIND_MESH: mesh_intersect(cx, oi, r, out); break;
        case RT_KIND_NODE: node_intersect(cx, oi, r, out); break;
    }
}
static int obj_shadow(Ctx* cx, int oi, const Ray* r, double tmax) {
    switch (cx->s->objects[oi].kind) {
        case RT_KIND_SPHERE: return sphere_shadow(cx, oi, r, tmax);
        case RT_KIND_PLANE: return plane_shadow(cx, oi, r, tmax);
        case RT_KIND_BOX: return box_shadow(cx, oi, r, tmax);
        case RT_KIND_MESH: return mesh_shadow(cx, oi, r, tmax);
        case RT_KIND_NODE: return node_shadow(cx, oi, r, tmax);
    }
    return 0;
}
static int obj_inside(Ctx* cx, int oi, v3 p) {
    switch (cx->s->objects[oi].kind) {
        case RT_KIND_SPHERE: return sphere_inside(cx, oi, p);
        case RT_KIND_BOX: return box_inside(cx, oi, p);
        case RT_KIND_NODE: return node_inside(cx, oi, p);
        default: return 0;   /* Plane, Mesh: base class, geometry/__init__.py:53 */
    }
}
static int obj_get_material(Ctx* cx, int oi, v3 p) {
    const RtObject* o = &cx->s->objects[oi];
    switch (o->kind) {
        case RT_KIND_PLANE: return plane_get_material(cx, oi, p);
        case RT_KIND_NODE: return node_get_material(cx, oi, p);
        default: return o->n_mats > 0 ? o->mat0 : -2;  /* geometry/__init__.py:56-57 */
    }
}

/* ---------------------------------------------------------------- shading, scene.py:81-209 */
static v3 cast_ray(Ctx* cx, const Ray* ray, int max_recursion, int in_shape);

/* stability probe only: v + perturb * max(1, |v|_inf) * (random vector in [-1,1]^3) */
static v3 probe_jiggle(Ctx* cx, v3 v, int relative_to_length) {
    cx->pstate[0]++;
    uint32_t c[4] = {cx->pstate[0], cx->pstate[1], 0x9e3779b9u, 0};
    philox4x32_10(c, cx->pstate[2], cx->pstate[3]);
    double m = fmax(fabs((double)v.x), fmax(fabs((double)v.y), fabs((double)v.z)));
    double scale = cx->perturb * (relative_to_length ? m : fmax(1.0, m));
    return V3((float)((double)v.x + scale * (c[0] * 0x1p-31 - 1.0)), (float)((double)v.y + scale * (c[1] * 0x1p-31 - 1.0)),
              (float)((double)v.z + scale * (c[2] * 0x1p-31 - 1.0)));
}

static v3 regular_lighting(Ctx* cx, const Ray* ray, const Hit* hit) {
    const RtSceneDesc* s = cx->s;
    v3 colour = V3(0, 0, 0);
    int err = 0;
    const RtObject* g = &s->objects[hit->geom];
    const RtMaterial* mat = &s->materials[hit->mat];
    v3 diffuse;
    if (g->kind == RT_KIND_BOX) diffuse = box_get_diffuse(cx, hit->geom, hit->p, &err);
    else if (g->kind == RT_KIND_PLANE) diffuse = plane_get_diffuse(cx, hit->geom, hit->p, &err);
    else diffuse = v3p(mat->diffuse);
    if (err) cx->c.error_flags |= 1ull << RT_ERR_NO_MATERIAL;
    cx->c.shaded_hits++;

    for (int li = 0; li < s->n_lights; li++) {
        const RtLight* L = &s->lights[li];
        Ray sr; double tmax;
        v3 lv = v3p(L->vector);
        if (L->type == RT_LIGHT_POINT) { sr.o = hit->p; sr.d = vsub(lv, hit->p); tmax = 1.0; }
        else { sr.o = hit->p; sr.d = vneg(lv); tmax = INFINITY; }
        cx->c.rays_shadow++;
        int skip = 0;
        for (int oi = 0; oi < s->n_top; oi++)
            if (obj_shadow(cx, oi, &sr, tmax)) { skip = 1; break; }
        if (skip) continue;
        v3 light_dir = L->type == RT_LIGHT_POINT ? vnormalize(vsub(lv, hit->p)) : vnormalize(vneg(lv));
        v3 normal = hit->n;
        double nl = (double)vdot(normal, light_dir);
        v3 lambert = vscaled(diffuse, nl > 0.0 ? nl : 0.0);
        v3 half_vect = vnormalize(vsub(light_dir, ray->d));
        double nh = (double)vdot(normal, half_vect);
        double sp = pow(nh > 0.0 ? nh : 0.0, mat->hardness);
        v3 specular = vscaled(v3p(mat->specular), sp);
        colour = vadd(colour, vmul(vscale(v3p(L->colour), L->power), vadd(lambert, specular)));
    }
    colour = vadd(colour, vmul(v3p(s->ambient), diffuse));
    return colour;
}

static v3 compute_refraction(Ctx* cx, const Ray* ray, Hit* hit, int in_shape, int max_recursion) {
    const RtMaterial* mat = &cx->s->materials[hit->mat];
    float eta;
    if (in_shape) { eta = (float)mat->refr_index; hit->n = vneg(hit->n); }
    else eta = (float)(1.0 / mat->refr_index);
    v3 rd = vrefract(ray->d, hit->n, eta);
    if (veq(rd, V3(0, 0, 0))) return V3(0, 0, 0);   /* total internal reflection -> black, scene.py:198-199 */
    Ray rr;
    rr.o = vadd(hit->p, vscaled(rd, 0.0001));
    rr.d = rd;
    if (max_recursion - 1 > 0) cx->c.rays_refract++;   /* a depth-0 call returns black without tracing */
    return cast_ray(cx, &rr, max_recursion - 1, !in_shape);
}

static v3 cast_ray(Ctx* cx, const Ray* ray_in, int max_recursion, int in_shape) {
    if (max_recursion == 0) return V3(0, 0, 0);
    const RtSceneDesc* s = cx->s;
    Ray jiggled;
    const Ray* ray = ray_in;
    if (cx->perturb > 0) { jiggled.o = probe_jiggle(cx, ray_in->o, 0); jiggled.d = probe_jiggle(cx, ray_in->d, 1); ray = &jiggled; }
    HitList hl = {0, 0, 0};
    for (int oi = 0; oi < s->n_top; oi++) obj_intersect(cx, oi, ray, &hl);
    if (hl.n == 0) { free(hl.v); return V3(0, 0, 0); }
    Hit first = hl.v[0];
    for (int i = 1; i < hl.n; i++) if (hl.v[i].t < first.t) first = hl.v[i];
    free(hl.v);
    if (cx->perturb > 0) first.p = probe_jiggle(cx, first.p, 0);
    if (first.mat < 0) {   /* reference raises IndexError */
        cx->c.error_flags |= 1ull << RT_ERR_NO_MATERIAL;
        return V3(0, 0, 0);
    }
    const RtMaterial* mat = &s->materials[first.mat];
    v3 colour;
    if (mat->type == RT_MAT_MIRROR) {
        v3 rd = vreflect(ray->d, first.n);
        Ray rr; rr.o = vadd(first.p, vscaled(rd, 0.01)); rr.d = rd;
        if (max_recursion - 1 > 0) cx->c.rays_reflect++;
        v3 reflection = cast_ray(cx, &rr, max_recursion - 1, 0);
        colour = regular_lighting(cx, ray, &first);
        colour = vadd(vscaled(colour, mat->tint), vscaled(reflection, 1.0 - mat->tint));
    } else if (mat->type == RT_MAT_REFRACTIVE) {
        v3 refraction = compute_refraction(cx, ray, &first, in_shape, max_recursion);
        colour = regular_lighting(cx, ray, &first);
        colour = vadd(vscaled(colour, mat->tint), vscaled(refraction, 1.0 - mat->tint));
    } else {
        colour = regular_lighting(cx, ray, &first);
    }
    float cxr = fmaxf(0.0f, fminf(1.0f, colour.x));
    float cyr = fmaxf(0.0f, fminf(1.0f, colour.y));
    float czr = fmaxf(0.0f, fminf(1.0f, colour.z));
    return V3(cxr, cyr, czr);
}

/* ---------------------------------------------------------------- sampling, scene.py:35-79,118-138 */
static void sunflower(int n, v3 origin, double radius, v3* out) {
    const double phi = (1.0 + sqrt(5.0)) / 2.0;
    const double stride = 2.0 * M_PI / phi;
    for (int k = 1; k <= n; k++) {
        double r = radius * sqrt(k - 0.5) / sqrt(n - 0.5);
        double th = k * stride;
        double x = r * cos(th) + (double)origin.x;
        double y = r * sin(th) + (double)origin.y;
        out[k - 1] = V3((float)x, (float)y, origin.z);
    }
}

/* x (or y) of pixel centre `idx`, accumulated exactly as scene.py:42,48,75,77 do */
static double accum_coord(double start_edge, double step, int first, int idx) {
    double v = start_edge + (0.5 + first) * step;
    for (int i = first; i < idx; i++) v += step;
    return v;
}

static void render_pixel(Ctx* cx, int col, int row, int strip_col0, double* rgb_out, v3* dof_pts, v3* aa_pts) {
    const RtSceneDesc* s = cx->s;
    double dx = (s->right - s->left) / s->width;
    double dy = (s->top - s->bottom) / s->height;
    double x = accum_coord(s->left, dx, strip_col0, col);
    double y = accum_coord(s->bottom, dy, 0, row);
    v3 pos = v3p(s->cam_pos), u = v3p(s->cam_u), v = v3p(s->cam_v), w = v3p(s->cam_w);
    v3 colour = V3(0, 0, 0);
    v3 base_dir = vnormalize(vsub(vadd(vscaled(u, x), vscaled(v, y)), vscaled(w, s->cam_d)));
    v3 focal_point = vadd(pos, vscaled(base_dir, s->focal_length));
    int D = s->dof_samples, S = s->aa_samples, T = s->n_times;
    sunflower(D, pos, s->aperture, dof_pts);
    for (int d = 0; d < D; d++) {
        v3 dof_origin = dof_pts[d];
        v3 dof_dir = vnormalize(vsub(focal_point, dof_origin));
        sunflower(S, dof_origin, 2 * (dx + dy), aa_pts);
        for (int a = 0; a < S; a++) {
            Ray ray; ray.o = aa_pts[a]; ray.d = dof_dir;
            if (s->jitter) {
                uint64_t idx = (((uint64_t)col * (uint64_t)s->height + (uint64_t)row) * (uint64_t)D + (uint64_t)d) * (uint64_t)S + (uint64_t)a;
                uint32_t c[4] = {(uint32_t)idx, (uint32_t)(idx >> 32), 0, 0};
                philox4x32_10(c, (uint32_t)s->jitter_seed, (uint32_t)(s->jitter_seed >> 32));
                double r1 = c[0] * 0x1p-32, r2 = c[1] * 0x1p-32, r3 = c[2] * 0x1p-32;
                v3 noise = vscaled(vnormalize(V3((float)r1, (float)r2, (float)r3)), 0.1 * (dx + dy));
                ray.o = vadd(ray.o, noise);
            }
            for (int t = 0; t < T; t++) {
                cx->time = s->motion_times[t];
                cx->c.rays_primary++;
                colour = vadd(colour, cast_ray(cx, &ray, s->max_depth, 0));
            }
        }
    }
    float n = (float)(S * D * T);
    rgb_out[0] = (double)(colour.x / n);
    rgb_out[1] = (double)(colour.y / n);
    rgb_out[2] = (double)(colour.z / n);
}

static void counters_add(RtCounters* a, const RtCounters* b) {
    a->rays_primary += b->rays_primary; a->rays_shadow += b->rays_shadow;
    a->rays_reflect += b->rays_reflect; a->rays_refract += b->rays_refract;
    a->bvh_nodes += b->bvh_nodes; a->tri_tests += b->tri_tests; a->prim_tests += b->prim_tests;
    a->shaded_hits += b->shaded_hits; a->mesh_hits_smooth += b->mesh_hits_smooth;
    a->queued_rays += b->queued_rays; a->error_flags |= b->error_flags;
}

/* Render an explicit pixel list. strip_col0[i] is the first column of the strip the pixel
 * would belong to in the reference's render(subimage, tasks) (it seeds the x accumulation,
 * scene.py:42); pass 0 for a full-frame render. out is (n, 3) float64. */
int oracle_render_pixels(const RtSceneDesc* s, const int32_t* cols, const int32_t* rows, const int32_t* strip_col0,
                         int64_t n, double* out, RtCounters* counters, int n_threads) {
    if (!s || s->abi_version != RT_ABI_VERSION) return RT_ERR_INVALID;
    RtCounters total; memset(&total, 0, sizeof total);
#ifdef _OPENMP
    if (n_threads > 0) omp_set_num_threads(n_threads);
#else
    (void)n_threads;
#endif
#pragma omp parallel
    {
        Ctx cx; memset(&cx, 0, sizeof cx); cx.s = s;
        v3* dof_pts = (v3*)malloc(sizeof(v3) * (size_t)(s->dof_samples > 0 ? s->dof_samples : 1));
        v3* aa_pts = (v3*)malloc(sizeof(v3) * (size_t)(s->aa_samples > 0 ? s->aa_samples : 1));
#pragma omp for schedule(dynamic, 16)
        for (int64_t i = 0; i < n; i++)
            render_pixel(&cx, cols[i], rows[i], strip_col0 ? strip_col0[i] : 0, out + 3 * i, dof_pts, aa_pts);
#pragma omp critical
        counters_add(&total, &cx.c);
        free(dof_pts); free(aa_pts);
    }
    if (counters) *counters = total;
    return RT_OK;
}

/* Stability probe (tests only; NOT reference behaviour): oracle_render_pixels with every cast_ray input ray and every
 * closest-hit position moved by a random vector of relative size `eps` (see Ctx.perturb).  Deterministic in (seed, pixel). */
int oracle_render_pixels_perturbed(const RtSceneDesc* s, const int32_t* cols, const int32_t* rows, int64_t n, double* out,
                                   double eps, uint64_t seed, int n_threads) {
    if (!s || s->abi_version != RT_ABI_VERSION) return RT_ERR_INVALID;
#ifdef _OPENMP
    if (n_threads > 0) omp_set_num_threads(n_threads);
#else
    (void)n_threads;
#endif
#pragma omp parallel
    {
        Ctx cx; memset(&cx, 0, sizeof cx); cx.s = s; cx.perturb = eps;
        v3* dof_pts = (v3*)malloc(sizeof(v3) * (size_t)(s->dof_samples > 0 ? s->dof_samples : 1));
        v3* aa_pts = (v3*)malloc(sizeof(v3) * (size_t)(s->aa_samples > 0 ? s->aa_samples : 1));
#pragma omp for schedule(dynamic, 16)
        for (int64_t i = 0; i < n; i++) {
            cx.pstate[0] = 0; cx.pstate[1] = (uint32_t)(cols[i] * 65599u + rows[i]);
            cx.pstate[2] = (uint32_t)seed; cx.pstate[3] = (uint32_t)(seed >> 32);
            render_pixel(&cx, cols[i], rows[i], 0, out + 3 * i, dof_pts, aa_pts);
        }
        free(dof_pts); free(aa_pts);
    }
    return RT_OK;
}

/* What the FIRST sample of a pixel sees (dof 0, aa 0, time 0, no jitter): out[4*i..] = leaf object index of the closest hit
 * (-1: none), material index, bit li set when light li is shadowed at the hit (scene.py:160-164), material type (RtMatType).
 * Tests use it to say WHERE a residue pixel lies: silhouettes / material (checker) edges / shadow edges are the places where
 * these values change between neighbouring pixels; mirror / refractive materials mark the reflection and refraction regions. */
int oracle_pixel_features(const RtSceneDesc* s, const int32_t* cols, const int32_t* rows, int64_t n, int32_t* out, int n_threads) {
    if (!s || s->abi_version != RT_ABI_VERSION) return RT_ERR_INVALID;
#ifdef _OPENMP
    if (n_threads > 0) omp_set_num_threads(n_threads);
#else
    (void)n_threads;
#endif
    double dx = (s->right - s->left) / s->width, dy = (s->top - s->bottom) / s->height;
#pragma omp parallel
    {
        Ctx cx; memset(&cx, 0, sizeof cx); cx.s = s;
#pragma omp for schedule(dynamic, 64)
        for (int64_t i = 0; i < n; i++) {
            double x = accum_coord(s->left, dx, 0, cols[i]), y = accum_coord(s->bottom, dy, 0, rows[i]);
            v3 pos = v3p(s->cam_pos), u = v3p(s->cam_u), v = v3p(s->cam_v), w = v3p(s->cam_w);
            v3 base_dir = vnormalize(vsub(vadd(vscaled(u, x), vscaled(v, y)), vscaled(w, s->cam_d)));
            v3 focal_point = vadd(pos, vscaled(base_dir, s->focal_length));
            v3 dof0, aa0;
            sunflower(1, pos, s->dof_samples == 1 ? s->aperture : 0.0, &dof0);     /* centre of the lens when there are several */
            if (s->dof_samples != 1) dof0 = pos;
            aa0 = dof0;
            Ray ray; ray.o = aa0; ray.d = vnormalize(vsub(focal_point, dof0));
            cx.time = s->motion_times[0];
            HitList hl = {0, 0, 0};
            for (int oi = 0; oi < s->n_top; oi++) obj_intersect(&cx, oi, &ray, &hl);
            int32_t* o = out + 4 * i;
            o[0] = -1; o[1] = -1; o[2] = 0; o[3] = 0;
            if (hl.n > 0) {
                Hit first = hl.v[0];
                for (int k = 1; k < hl.n; k++) if (hl.v[k].t < first.t) first = hl.v[k];
                o[0] = first.geom; o[1] = first.mat;
                if (first.mat >= 0) o[3] = s->materials[first.mat].type;
                for (int li = 0; li < s->n_lights && li < 31; li++) {
                    const RtLight* L = &s->lights[li];
                    Ray sr; double tmax; v3 lv = v3p(L->vector);
                    if (L->type == RT_LIGHT_POINT) { sr.o = first.p; sr.d = vsub(lv, first.p); tmax = 1.0; }
                    else { sr.o = first.p; sr.d = vneg(lv); tmax = INFINITY; }
                    for (int oi = 0; oi < s->n_top; oi++)
                        if (obj_shadow(&cx, oi, &sr, tmax)) { o[2] |= 1 << li; break; }
                }
            }
            free(hl.v);
        }
    }
    return RT_OK;
}

/* Scene.render over the rectangle [col0,col0+ncols) x [row0,row0+nrows); out (ncols,nrows,3) float64,
 * axis 0 = column, axis 1 = row bottom->top (scene.py:39,73). */
int oracle_render(const RtSceneDesc* s, int col0, int ncols, int row0, int nrows, double* out,
                  RtCounters* counters, int n_threads) {
    int64_t n = (int64_t)ncols * nrows;
    int32_t* cols = (int32_t*)malloc(sizeof(int32_t) * (size_t)(n > 0 ? n : 1));
    int32_t* rows = (int32_t*)malloc(sizeof(int32_t) * (size_t)(n > 0 ? n : 1));
    int32_t* sc0 = (int32_t*)malloc(sizeof(int32_t) * (size_t)(n > 0 ? n : 1));
    for (int c = 0; c < ncols; c++)
        for (int r = 0; r < nrows; r++) {
            cols[(int64_t)c * nrows + r] = col0 + c;
            rows[(int64_t)c * nrows + r] = row0 + r;
            sc0[(int64_t)c * nrows + r] = col0;
        }
    int rc = oracle_render_pixels(s, cols, rows, sc0, n, out, counters, n_threads);
    free(cols); free(rows); free(sc0);
    return rc;
}

/* Trace explicit rays (unit tests): origins/dirs (n,3) float32, times (n) float64 -> rgb (n,3) float32,
 * first-hit record hit_t (n) float64 (inf = miss), hit_n/hit_p (n,3) float32, hit_geom/hit_mat (n) int32 */
int oracle_trace_rays(const RtSceneDesc* s, int64_t n, const float* origins, const float* dirs, const double* times,
                      float* rgb, double* hit_t, float* hit_n, float* hit_p, int32_t* hit_geom, int32_t* hit_mat) {
    if (!s || s->abi_version != RT_ABI_VERSION) return RT_ERR_INVALID;
    Ctx cx; memset(&cx, 0, sizeof cx); cx.s = s;
    for (int64_t i = 0; i < n; i++) {
        Ray r; r.o = v3p(origins + 3 * i); r.d = v3p(dirs + 3 * i);
        cx.time = times ? times[i] : 0.0;
        if (rgb) { v3 c = cast_ray(&cx, &r, s->max_depth, 0); rgb[3 * i] = c.x; rgb[3 * i + 1] = c.y; rgb[3 * i + 2] = c.z; }
        if (hit_t) {
            HitList hl = {0, 0, 0};
            for (int oi = 0; oi < s->n_top; oi++) obj_intersect(&cx, oi, &r, &hl);
            Hit first; first.t = INFINITY; first.n = V3(0, 0, 0); first.p = V3(0, 0, 0); first.mat = -1; first.geom = -1;
            for (int k = 0; k < hl.n; k++) if (hl.v[k].t < first.t) first = hl.v[k];
            free(hl.v);
            hit_t[i] = first.t;
            if (hit_n) { hit_n[3 * i] = first.n.x; hit_n[3 * i + 1] = first.n.y; hit_n[3 * i + 2] = first.n.z; }
            if (hit_p) { hit_p[3 * i] = first.p.x; hit_p[3 * i + 1] = first.p.y; hit_p[3 * i + 2] = first.p.z; }
            if (hit_geom) hit_geom[i] = first.geom;
            if (hit_mat) hit_mat[i] = first.mat;
        }
    }
    return RT_OK;
}

/* Any-hit shadow query of scene.py:161-164 for explicit rays: occluded (n) int32 */
int oracle_occluded(const RtSceneDesc* s, int64_t n, const float* origins, const float* dirs, const double* times,
                    const double* tmax, int32_t* occluded) {
    if (!s || s->abi_version != RT_ABI_VERSION) return RT_ERR_INVALID;
    Ctx cx; memset(&cx, 0, sizeof cx); cx.s = s;
    for (int64_t i = 0; i < n; i++) {
        Ray r; r.o = v3p(origins + 3 * i); r.d = v3p(dirs + 3 * i);
        cx.time = times ? times[i] : 0.0;
        int occ = 0;
        for (int oi = 0; oi < s->n_top && !occ; oi++) occ = obj_shadow(&cx, oi, &r, tmax ? tmax[i] : INFINITY);
        occluded[i] = occ;
    }
    return RT_OK;
}

int oracle_philox(uint64_t idx, uint64_t seed, uint32_t out[4]) {
    uint32_t c[4] = {(uint32_t)idx, (uint32_t)(idx >> 32), 0, 0};
    philox4x32_10(c, (uint32_t)seed, (uint32_t)(seed >> 32));
    memcpy(out, c, sizeof c);
    return 0;
}

int oracle_num_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}
