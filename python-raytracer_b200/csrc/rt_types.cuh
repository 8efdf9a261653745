// rt_types.cuh - device-resident scene layout (what rt_scene_create builds in HBM).
#pragma once
#include "rt_math.cuh"
#include "../../include/b200rt.h"

#define RT_MAX_NEST 8        // CSG / transform nesting depth supported (NovelScene uses 5)
#define RT_BLOCK 128         // threads per block of every ray kernel (the traversal stack layout depends on it)
#ifndef RT_STACK_DEPTH
#define RT_STACK_DEPTH 32    // BVH traversal stack entries per ray held in shared memory (object BVH + mesh LBVH share it)
#endif
#define RT_STACK_ROWS RT_STACK_DEPTH
#ifndef RT_STACK_SPILL
#define RT_STACK_SPILL 88
#endif
// RT_STACK_SPILL: further entries in thread-local memory; beyond both the render reports RT_ERR_LIMIT
#define RT_TLAS_NO_TMAX 1    // object-BVH child flag
#define RT_MODE_BRUTE_MESH 1
#define RT_MODE_LINEAR_TOP 2

// One geometry object, 16-byte aligned so fields load as 128-bit vectors (warp-uniform index
// -> one broadcast transaction).  Meaning of a/b per kind as in RtObject (include/b200rt.h).
struct __attribute__((aligned(16))) DevObject {
    float4 a;        // sphere: centre, w = radius | plane: point | box: minpos
    float4 b;        // plane: normal | box: maxpos
    float4 speed;    // xyz
    int kind, n_mats, mat0, mat1;
    int tex, mesh, first_child, n_children;
    int csg, xform, pad0, pad1;
    float4 ax_w;     // plane width axis
    float4 ax_h;     // plane height axis
    double tex_scale;
    double pad2;
};

struct __attribute__((aligned(16))) DevMaterial {
    float4 diffuse;   // w = hardness
    float4 specular;  // w = tint
    float refr_index; // eta inside
    float inv_refr;   // (float)(1.0 / refr_index)
    float one_minus_tint;
    int type;
};

struct __attribute__((aligned(16))) DevLight {
    float4 vector;    // w = type
    float4 colour_power;  // colour * (float)power, scene.py:183
    float4 dir_norm;  // directional: normalize(-vector) (scene.py:173), precomputed
};

struct __attribute__((aligned(16))) DevXform { float M[16]; float Minv[16]; };   // columns are read as float4

// Binary LBVH node as the builder emits it before the 4-wide collapse (build-time only), 64 B:
//   q0 = (c0.min.xyz, c0.max.x)  q1 = (c0.max.yz, c1.min.xy)  q2 = (c1.min.z, c1.max.xyz)  q3 = (bits child0, bits child1, -, -)
struct __attribute__((aligned(16))) BvhNode { float4 q0, q1, q2, q3; };

// Traversal node: 4-wide, 128 B = one cache line, boxes of the (up to) four children in SoA so the four slab
// tests vectorise over registers; link >= 0: node index, < 0: ~leaf (triangle slot / object index),
// RT_LINK_NONE: empty slot, box [+inf, -inf] (never hit).  The walk addresses the planes by byte offset (RayInv).  flags: per-child bits (object BVH only).
#define RT_LINK_NONE 0x7fffffff
// mesh LBVH leaves hold 1..RT_LEAF_TRIS consecutive triangle slots: link = ~(first_slot | (count - 1) << RT_LEAF_SLOT_BITS)
#ifndef RT_LEAF_TRIS
#define RT_LEAF_TRIS 2
#endif
#define RT_LEAF_SLOT_BITS 28
struct __attribute__((aligned(16))) BvhNode4 { float4 lox, loy, loz, hix, hiy, hiz; int4 link; int4 flags; };

struct DevMesh {
    const float4* tri;       // 3 float4 per triangle slot (Morton order): v0|v1|v2, w of v0 = bits(original face index)
    const BvhNode4* nodes;   // 4-wide nodes (root = 0); NULL when n_tris < 2
    const int4* face_vidx;   // per triangle slot: vertex indices (for smooth normals) + original face
    const float4* vnormals;  // per vertex, NULL when flat shaded
    int n_tris;
    int flat_shaded;
    int gate;
    float gate_radius;
    float4 gate_min, gate_max, gate_center;
};

struct DevTexture { const uint8_t* rgb; int width, height; };

// Copies of the small per-scene records that EVERY ray reads, held in the kernel parameter bank (SceneView is passed by value):
// constant-bank operands / LDC instead of global loads through the L1 data pipe, which is what bounds the walkers.
#define RT_CONST_LIGHTS 4      // the first lights
#define RT_CONST_LIN 4         // the linearly tested top-level objects when they are all planes (the usual case: ground, walls)
#define RT_CONST_MESH 8        // the headers of the top-level meshes (scenes with more run on the general kernels)
struct SceneConst {
    int n_lin_planes;          // == n_lin when every linearly tested object is a plane and n_lin <= RT_CONST_LIN, else -1
    int n_sep_mesh;            // == n_sep_mesh when <= RT_CONST_MESH, else -1
    int lin_obj[RT_CONST_LIN];
    int sep_obj[RT_CONST_MESH];
    float4 light_vec[RT_CONST_LIGHTS], light_colour[RT_CONST_LIGHTS], light_dirn[RT_CONST_LIGHTS];    // DevLight fields
    float4 lin_point[RT_CONST_LIN], lin_normal[RT_CONST_LIN], lin_speed[RT_CONST_LIN];
    DevMesh sep_mesh[RT_CONST_MESH];
};

struct SceneView {
    const DevObject* objects; int n_objects, n_top;
    const DevMaterial* materials; int n_materials;
    const DevLight* lights; int n_lights;
    const DevXform* xforms;
    const DevMesh* meshes;
    const DevTexture* textures;
    float3 ambient;
    int max_depth;
    int mesh_mode;    // test aids: bit 0 = brute force over all faces (mesh.py:77), bit 1 = linear loop over the top-level objects
    // object BVH over the bounded top-level objects (replaces the `for obj in self.objects` loops of scene.py:87,160
    // for everything that has a finite extent); same 4-wide node layout as the mesh LBVH, link < 0: ~object index,
    // flags = per-child bits (RT_TLAS_NO_TMAX: the subtree holds an object whose shadow test ignores t_max).
    const BvhNode4* tlas_nodes;  // NULL: no object BVH (fewer than 2 bounded objects) -> plain loop over n_top
    int tlas_flagged;            // some child of the object BVH carries RT_TLAS_NO_TMAX (otherwise the walk skips the flag rows)
    int any_motion;              // some object has a non-zero speed (otherwise primitive tests skip the speed loads)
    int mesh_motion;             // EXTENSION (RtSceneDesc::mesh_motion): meshes translate by speed * time like the primitives do
    const int* lin_objs;         // top-level objects tested one by one, scene order: no finite bounds (planes, nodes holding
    int n_lin;                   // planes), or everything but the separate meshes when there is no object BVH
    // top-level meshes are kept OUT of that walk and traversed after it (gate test each, scene order), so that the lanes
    // of a warp enter the (long) LBVH traversal together instead of at whatever step their walk happens to reach the mesh
    const int* sep_mesh;
    int n_sep_mesh;
    // padded union of the object BVH's root boxes and the separate meshes' gate volumes: a ray that misses it needs neither
    // walk (most rays of an open scene: ground-plane hits whose shadow rays leave upwards).  Lives in the kernel parameter bank.
    int has_world;
    float3 world_lo, world_hi;
    // conservative bounds of every object IN ITS PARENT'S SPACE (children of CSG / transform nodes; top-level objects and
    // unbounded ones carry [-inf, +inf]): a ray or point outside them cannot hit / be inside that child, so whole subtrees
    // of a node are skipped without transforming the ray into them.  NULL when the scene has no nodes.
    const float4* child_lo;
    const float4* child_hi;
    // sampling tables (scene.py:41-43,118-138 precomputed once per scene)
    const float* col_x;          // [W]  (float)x of pixel-centre column
    const float* row_y;          // [H]
    const float4* dof_origin;    // [D]
    const float4* ray_origin;    // [D*S] AA point around each DOF point
    const float* times;          // [T] (float)motion_times
    float3 cam_pos, cam_u, cam_v, cam_w;
    float focal_length, noise_scale;
    int width, height, D, S, T, jitter;
    uint32_t seed_lo, seed_hi;
    SceneConst c;
};

struct Ray { float3 o, d; };

// geometry/__init__.py:15-22
struct Hit {
    float t;
    float3 n, p;
    int mat;    // material index; -1 = None (hierarchy.py:73), -2 = reference would raise IndexError
    int geom;   // leaf object index (scene.py:143 dispatches on the hit's geometry)
};

// per-thread tallies, reduced per block into RtCounters
struct Tally {
    unsigned shadow, reflect, refract, nodes, tris, prims, shaded, smooth, err;
};
