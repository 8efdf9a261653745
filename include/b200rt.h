/* b200rt.h - C ABI of the B200-native ray-intersection + shading path.
 *
 * Drop-in boundary for the hot path of SpacewaIker/python-raytracer
 * (citations are file:line under the reference's `provided/`):
 *
 *   reference caller           main.py:25-30      load_scene(infile) ; full_scene.render([subimage, tasks])
 *   reference entry point      scene.py:35-79     Scene.render(subimage, tasks) -> ndarray (W_sub, H, 3)
 *   reference inner path       scene.py:81-209    cast_ray / _compute_regular_lighting / _compute_refraction
 *                              geometry/ *.py     intersect / shadow_intersect / is_inside / get_material / get_diffuse
 *   reference fan-in           glue.py:17-26      concatenate strips, rot90, *255 truncation
 *
 * The reference has no FFI of its own (pure Python); the binding a maintainer adds is the
 * ctypes stub shown in INTEGRATION.md.  Every entry point returns 0 on success or a
 * non-zero RtStatus; nothing throws across the ABI; there is no hidden global state besides
 * a thread-local last-error string.  `stream` arguments are `cudaStream_t` passed as void*.
 *
 * The SAME scene description (`RtSceneDesc`, host pointers) is consumed by the CPU oracle
 * (`oracle/rt_oracle.c`, test infrastructure) so that both sides see identical inputs.
 */
#ifndef B200RT_H
#define B200RT_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RT_ABI_VERSION 3

/* ---- enums ------------------------------------------------------------------------- */
enum RtKind { RT_KIND_SPHERE = 0, RT_KIND_PLANE = 1, RT_KIND_BOX = 2, RT_KIND_MESH = 3, RT_KIND_NODE = 4 };
enum RtMatType { RT_MAT_DIFFUSE = 0, RT_MAT_MIRROR = 1, RT_MAT_REFRACTIVE = 2 };      /* scene.py:97,105 */
enum RtLightType { RT_LIGHT_POINT = 0, RT_LIGHT_DIRECTIONAL = 1 };                   /* scene.py:153,156 */
enum RtCsg { RT_CSG_UNION = 0, RT_CSG_INTERSECTION = 1, RT_CSG_DIFFERENCE = 2, RT_CSG_NONE = 3 }; /* hierarchy.py:49-58 */
enum RtGate { RT_GATE_AABB = 0, RT_GATE_SPHERE = 1 };                                 /* mesh.py:48-51 */

enum RtStatus {
    RT_OK = 0,
    RT_ERR_INVALID = 1,      /* bad argument / inconsistent description */
    RT_ERR_CUDA = 2,         /* a CUDA runtime call failed; see rt_last_error() */
    RT_ERR_NOMEM = 3,
    RT_ERR_NO_MATERIAL = 4,  /* a shaded hit had no material (reference: IndexError, geometry/__init__.py:57) */
    RT_ERR_LIMIT = 5         /* scene exceeds a compiled limit (CSG nesting depth, lights) */
};

/* ---- scene records (plain data, host pointers) ------------------------------------- */

/* helperclasses.py:28-37 ; scene_parser.py:129-142 */
typedef struct RtMaterial {
    float diffuse[3];
    float specular[3];
    int32_t type;       /* RtMatType */
    int32_t _pad;
    double hardness;    /* exponent of scene.py:181 (Python number) */
    double tint;        /* scene.py:104,108 (Python float) */
    double refr_index;  /* scene.py:191,194 (Python float) */
} RtMaterial;           /* 56 bytes */

/* helperclasses.py:50-56 ; scene_parser.py:107-122 */
typedef struct RtLight {
    int32_t type;       /* RtLightType */
    float colour[3];
    float vector[3];    /* point: position ; directional: direction */
    float power;
} RtLight;              /* 32 bytes */

/* One geometry object (leaf primitive or CSG/transform node).  Top-level objects come
 * first, in scene order (scene.py:87 iterates them in that order); the children of a node
 * are contiguous at [first_child, first_child + n_children).
 *   sphere  simple_geometry.py:12-83    a = center, radius
 *   plane   simple_geometry.py:86-176   a = point, b = normal, ax_w/ax_h = texture/checker axes, tex, tex_scale
 *   box     simple_geometry.py:179-355  a = minpos, b = maxpos, tex
 *   mesh    mesh.py:16-156              mesh = index into meshes[]
 *   node    hierarchy.py:11-146         csg, xform = index into xforms[], children
 */
typedef struct RtObject {
    int32_t kind;        /* RtKind */
    int32_t n_mats;      /* len(materials) after set_fallback_material (hierarchy.py:21-28) */
    int32_t mat0, mat1;  /* materials[0], materials[1] as indices into materials[] (-1 = absent) */
    int32_t tex;         /* texture index or -1 */
    int32_t mesh;        /* mesh index or -1 */
    int32_t first_child; /* node */
    int32_t n_children;  /* node */
    int32_t csg;         /* RtCsg (node) */
    int32_t xform;       /* index into xforms[] (node) */
    float radius;        /* sphere */
    float _pad0;
    float a[3];
    float b[3];
    float speed[3];      /* (0,0,0) when the reference's speed is None */
    float ax_w[3];       /* plane width_axis  simple_geometry.py:93-102 */
    float ax_h[3];       /* plane height_axis simple_geometry.py:103 */
    float _pad1;
    double tex_scale;    /* plane texture_scale (simple_geometry.py:158-160), 1.0 default */
    double radius_d;     /* sphere radius as the Python float the parser read (used by the oracle) */
} RtObject;              /* 128 bytes */

/* hierarchy.py:30-37: M = T*Rx*Ry*Rz*S, Minv = inverse(M); column-major (m[col*4+row]) like GLM */
typedef struct RtXform {
    float M[16];
    float Minv[16];
} RtXform;

/* mesh.py:17-51: vertices already (v + translate) * scale; per-vertex normals already
 * area-weighted and normalised (NULL when flat shaded); gate = the bounding volume chosen at mesh.py:48-51 */
typedef struct RtMesh {
    int32_t n_verts;
    int32_t n_faces;
    const float* verts;      /* n_verts * 3 */
    const int32_t* faces;    /* n_faces * 3 */
    const float* vnormals;   /* n_verts * 3 or NULL */
    int32_t flat_shaded;
    int32_t gate;            /* RtGate */
    float gate_min[3];
    float gate_max[3];
    float gate_center[3];
    float gate_radius;
    float _pad[2];
} RtMesh;

/* PIL image as the reference samples it: getpixel((i, j)) = rgb[(j*width + i)*3 ...] */
typedef struct RtTexture {
    int32_t width, height;
    const uint8_t* rgb;
} RtTexture;

/* The whole scene: what scene_parser.load_scene() hands to scene.Scene (scene_parser.py:158) */
typedef struct RtSceneDesc {
    int32_t abi_version;     /* RT_ABI_VERSION */
    int32_t width, height;   /* helperclasses.py:75-79 */
    /* camera basis and frustum, helperclasses.py:81-94 */
    float cam_pos[3], cam_u[3], cam_v[3], cam_w[3];
    double cam_d, left, right, bottom, top;
    /* lens and sampling, helperclasses.py:96-108 ; scene.py:28-29 */
    double focal_length, aperture;
    int32_t dof_samples, aa_samples, jitter;
    int32_t n_times;
    const double* motion_times;
    float ambient[3];
    int32_t max_depth;       /* cast_ray's max_recursion default, scene.py:81 (10) */
    uint64_t jitter_seed;    /* Philox4x32-10 key; counter = ((col*H+row)*D+dof)*S+aa */
    int32_t n_lights;
    int32_t n_materials;
    int32_t n_objects;       /* all objects, top-level first */
    int32_t n_top;           /* len(scene.objects) */
    int32_t n_xforms;
    int32_t n_meshes;
    int32_t n_textures;
    int32_t mesh_motion;     /* EXTENSION, 0 = reference behaviour: the reference stores a mesh's `speed` (scene_parser.py:255) and never
                              * uses it (mesh.py:72-153: meshes stand still).  Non-zero: mesh objects translate by speed * time exactly
                              * like spheres / planes / boxes do (simple_geometry.py:21-24), tested against the oracle extended the same way */
    const RtLight* lights;
    const RtMaterial* materials;
    const RtObject* objects;
    const RtXform* xforms;
    const RtMesh* meshes;
    const RtTexture* textures;
} RtSceneDesc;

/* Which pixels one call renders.  The region [col0,col0+ncols) x [row0,row0+nrows) is cut
 * into tile_w x tile_h tiles numbered column-major (tile = tcol * n_tile_rows + trow);
 * this call renders the tiles with tile % world == rank (scene.py:36-37 strips generalised
 * to interleaved tiles for load balance; glue.py:19-22 is the gather).  With world == 1 and
 * direct != 0 the output is written straight in the reference layout (ncols, nrows, 3);
 * otherwise it is the compact tile buffer (n_local_tiles, tile_h, tile_w, 3) that
 * rt_assemble_tiles() scatters into the reference layout. */
typedef struct RtRenderJob {
    int32_t col0, ncols, row0, nrows;
    int32_t tile_w, tile_h;
    int32_t rank, world;
    int32_t direct;          /* 0 = compact tile buffer; 1 = reference layout (world must be 1); RT_DIRECT_SHARED_FRAME = reference
                              * layout of the WHOLE region with world > 1: every rank writes its own tiles into one shared frame */
    int32_t mesh_mode;       /* test aids, bit field: 0 = accelerated; bit 0 = brute force over all faces (mesh.py:77);
                              * bit 1 = plain in-order loop over the top-level objects (scene.py:87,160) instead of the object BVH */
    int64_t max_batch_rays;  /* 0 = default */
} RtRenderJob;

#define RT_DIRECT_SHARED_FRAME 2

/* Device-side tallies of one rt_render call */
typedef struct RtCounters {
    uint64_t rays_primary;
    uint64_t rays_shadow;
    uint64_t rays_reflect;
    uint64_t rays_refract;
    uint64_t bvh_nodes;      /* BVH nodes fetched */
    uint64_t tri_tests;      /* triangles tested */
    uint64_t prim_tests;     /* analytic primitive tests */
    uint64_t shaded_hits;
    uint64_t mesh_hits_smooth;
    uint64_t queued_rays;    /* secondary rays written to / read from the compacted queues */
    uint64_t kernel_launches;/* launches issued by this call */
    uint64_t error_flags;    /* bit (RtStatus) set by kernels */
} RtCounters;

/* Per-kernel-class device timing (CUDA events on the launch stream) + the tallies of each class for the last
 * rt_render call; used by bench.py for the roofline of the dominant kernel.  One cast_ray level is three launches
 * (closest hit, shadow rays, shading); level 0 (the generated rays) is timed per kernel, deeper levels as a whole. */
typedef struct RtProfile {
    double ms_hit, ms_shadow, ms_shade;            /* level 0: k_hit / k_shadow / k_shade, summed since profiling was enabled */
    double ms_secondary, ms_resolve;               /* all kernels of levels >= 1 ; k_resolve */
    uint64_t n_hit, n_shadow, n_shade, n_secondary, n_resolve;   /* timed launches (n_secondary: levels) */
    RtCounters hit, shadow, shade;                 /* tallies of the level-0 kernels of the last rt_render */
    RtCounters secondary;                          /* tallies of all deeper-level kernels of the last rt_render */
    /* the queued levels in detail (their sum is ms_secondary): level 1 per kernel, levels >= 2 as a whole, k_tail */
    double ms_l1_hit, ms_l1_shadow, ms_l1_shade, ms_deep, ms_tail;
    uint64_t n_l1_hit, n_deep, n_tail;
} RtProfile;

/* Roofline denominators MEASURED on the current device by hand-written micro-benchmarks (csrc/rt_peaks.cuh): FP32 FMA
 * throughput of the CUDA cores, warp-instruction issue rate, L1 data-pipe wavefront rate (128 B each); *_nominal = SMs x
 * per-SM rate x the device's reported SM clock, for comparison. */
typedef struct RtPeaks {
    double fp32_fma_tflops, issue_ginst, lsu_gwavefronts;
    double fp32_fma_tflops_nominal, issue_ginst_nominal, lsu_gwavefronts_nominal;
    double sm_clock_mhz;
    int32_t sm_count, _pad;
} RtPeaks;

typedef struct RtScene RtScene;  /* opaque, device-resident */

/* ---- entry points -------------------------------------------------------------------- */

/* Replaces scene.Scene.__init__ + the per-object state built by scene_parser.load_scene
 * (scene_parser.py:158-161): copies the description to HBM in SoA form and precomputes the
 * pixel-centre and sunflower tables of scene.py:41-43,118-138. */
int rt_scene_create(const RtSceneDesc* desc, RtScene** out);

/* Replaces the per-ray brute-force face loop set-up (mesh.py:72-77): builds one LBVH per
 * mesh on the GPU (Morton codes, radix sort, Karras hierarchy, bottom-up refit). */
int rt_scene_build_bvh(RtScene* scene, void* stream);

/* Replaces Scene.render(subimage, tasks) (scene.py:35-79).  d_out is DEVICE memory:
 * job->direct ? (ncols, nrows, 3) : (n_local_tiles, tile_h, tile_w, 3), float32.
 * counters may be NULL; when given, the call synchronises the stream before returning.
 * The job rectangle must lie inside the frame (RT_ERR_INVALID otherwise).  ONE render call may be in flight per scene
 * handle: the wavefront buffers (samples, hit records, ray queues, counters) belong to the handle and are re-used and
 * re-grown by every call; use one handle per stream for concurrent renders.  The caller's current device is preserved.
 * Frames of more than one wavefront batch (64 Mi rays) are worked on two streams inside the call - level 0 of every batch on
 * `stream`, the deeper levels of the batch before it on a high-priority side stream of the handle - and joined before the call
 * returns its work to `stream`: to the caller the call is ordered on `stream` like a single kernel. */
int rt_render(RtScene* scene, const RtRenderJob* job, float* d_out, RtCounters* counters, void* stream);

/* Same, HOST output buffer (pageable or pinned): device staging + D2H copy inside the call;
 * synchronous.  This is the end-to-end call Scene.render() uses.  A page-locked destination (cudaHostAlloc / cudaHostRegister)
 * receives the part of the frame every batch completes while the later batches still render. */
int rt_render_host(RtScene* scene, const RtRenderJob* job, float* h_out, RtCounters* counters, void* stream);

/* rt_render for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid
 * it).  No counterpart in the reference.  Same arguments, same output (bit for bit)
 * and same tallies as rt_render, but the call allocates nothing - no cudaMalloc / cudaFree / allocating library routine: the
 * wavefront scratch (sample buffer, hit planes, occlusion planes, both ray queues: 216 bytes per ray) is a file-scope device array
 * of RT_STATIC_MAX_RAYS rays = 13.5 MiB that every GPU which loads the library holds until the process exits, and the frame is
 * worked through in batches of at most RT_STATIC_MAX_RAYS rays (any frame size; slower than rt_render for big frames).
 * RT_ERR_LIMIT when the scene's samples per pixel (dof * aa * times) exceed RT_STATIC_MAX_RAYS or its lights exceed
 * RT_STATIC_MAX_LIGHTS.  job->max_batch_rays is ignored.  ALL calls on one device share the array, whatever the scene handle
 * or stream: calls that overlap on the GPU race (order them on one stream or with events).  The array keeps the last call's
 * contents; nothing in it needs clearing (every record read was written by an earlier kernel of the same batch). */
#define RT_STATIC_MAX_RAYS 65536
#define RT_STATIC_MAX_LIGHTS 8
int rt_render_static(RtScene* scene, const RtRenderJob* job, float* d_out, RtCounters* counters, void* stream);

/* Number of pixels / tiles a job writes for (rank, world): sizes the compact buffer. */
int rt_job_extent(const RtRenderJob* job, int64_t* n_local_tiles, int64_t* n_floats);

/* Replaces glue.py:19-22: scatter `world` compact tile buffers (concatenated in rank order,
 * each rank's buffer starting at rank_offsets_floats[r]) into the reference layout (ncols, nrows, 3). */
int rt_assemble_tiles(const RtRenderJob* job, const float* d_gathered, const int64_t* rank_offsets_floats,
                      float* d_out, void* stream);

/* Fan-in without a gather (render.nu:10-15 + glue.py:19-22 fused into the render): one process per GPU of an NVLink/NVSwitch
 * box.  Rank 0 allocates the frame with rt_shared_alloc and hands the 64-byte handle to the other ranks (any transport:
 * torch.distributed broadcast); they map it with rt_shared_open (CUDA IPC, peer access enabled on first use) and pass the
 * mapped pointer as d_out of rt_render with job.direct = RT_DIRECT_SHARED_FRAME: k_resolve then stores every finished pixel of
 * their tiles straight into rank 0's HBM.  The caller orders the readers after the writers (one stream-ordered barrier). */
typedef struct RtIpcHandle { unsigned char bytes[64]; } RtIpcHandle;
int rt_shared_alloc(int64_t bytes, void** d_ptr, RtIpcHandle* handle);
int rt_shared_free(void* d_ptr);
int rt_shared_open(const RtIpcHandle* handle, void** d_ptr);
int rt_shared_close(void* d_ptr);

/* Replaces main.py:31-33 / glue.py:23-25: rot90 + (image*255) truncation, (ncols,nrows,3) f32 -> (nrows,ncols,3) u8 */
int rt_finalize_u8(const float* d_image, int32_t ncols, int32_t nrows, uint8_t* d_out_u8, void* stream);

/* Scene.render() + main.py:31-33 in one call, HOST output already in PNG layout: renders the whole region of `job`
 * (world must be 1), converts on the device (rot90 + *255 truncation) and copies (nrows, ncols, 3) uint8 to h_out_u8:
 * a quarter of the float image's bytes cross PCIe.  Synchronous. */
int rt_render_u8_host(RtScene* scene, const RtRenderJob* job, uint8_t* h_out_u8, RtCounters* counters, void* stream);

int rt_scene_destroy(RtScene* scene);

/* ---- mesh ingestion at scale: the per-mesh load-time work of Mesh.__init__ (mesh.py:17-51) ---- */
typedef struct RtObj RtObj;
/* Replaces igl.read_obj (mesh.py:20) for large files: parses the `v`, `vn` and `f` rows on the host (doubles, first index of
 * every corner token, 1-based or negative, polygons fan-triangulated); returns the counts, rt_obj_copy fills caller arrays
 * (verts n_verts*3 f64, norms n_norms*3 f64, faces n_faces*3 i32; any may be NULL), rt_obj_free releases the handle. */
int rt_obj_load(const char* path, RtObj** out, int64_t* n_verts, int64_t* n_norms, int64_t* n_faces);
int rt_obj_copy(const RtObj* obj, double* verts, double* norms, int32_t* faces);
int rt_obj_free(RtObj* obj);
/* Replaces the vertex loop of mesh.py:23-24 ((v + translate) * scale in fp32), Mesh._compute_normals (mesh.py:53-70:
 * area-weighted face normals accumulated per vertex IN FACE ORDER, then normalised) and the extent / centre / farthest
 * vertex of mesh.py:29-43, all on the device.  Pointers are HOST buffers: raw vertices as read from the OBJ (n_verts*3
 * f64), faces (n_faces*3 i32); outputs verts and vnormals (n_verts*3 f32; vnormals only when want_normals), extent min /
 * max / centre (3 f32 each) and the bounding-sphere radius.  Synchronous.  Results are bit-identical to the host-side
 * numpy loader of the Python package (tests/test_gpu_parity.py). */
int rt_mesh_ingest(int32_t n_verts, const double* h_raw_verts, int32_t n_faces, const int32_t* h_faces,
                   const float* translate, float scale, int32_t want_normals, float* h_verts, float* h_vnormals,
                   float* h_extent_min, float* h_extent_max, float* h_center, float* h_max_dist, void* stream);

/* Introspection used by tests and the bench */
int rt_scene_bvh_stats(const RtScene* scene, int32_t mesh, int64_t* n_nodes, int64_t* n_leaf_tris, int64_t* bytes);
/* Copies the LBVH of `mesh` back to host buffers for validation: 4-wide nodes (n_nodes*32 floats: 6 float4 child boxes in
 * SoA, int4 links, int4 flags), leaf order (n_faces int32) */
int rt_scene_bvh_download(const RtScene* scene, int32_t mesh, float* h_nodes, int32_t* h_order);
/* Unit-test aids: cast_ray (scene.py:81-116) and the any-hit loop (scene.py:160-164) for explicit rays.
 * All pointers are HOST buffers: origins/dirs (n,3) f32, times (n) f32; outputs rgb (n,3), first-hit t (n, inf = miss),
 * normal/position (n,3), leaf object index and material index (n).  Outputs may be NULL. */
int rt_trace_rays(RtScene* scene, int64_t n, const float* h_origins, const float* h_dirs, const float* h_times,
                  float* h_rgb, float* h_hit_t, float* h_hit_n, float* h_hit_p, int32_t* h_hit_geom, int32_t* h_hit_mat,
                  int32_t mesh_mode, void* stream);
int rt_occluded(RtScene* scene, int64_t n, const float* h_origins, const float* h_dirs, const float* h_times,
                const float* h_tmax, int32_t* h_occluded, int32_t mesh_mode, void* stream);
/* enable != 0: start recording an event pair around every kernel of rt_render (and reset the sums); 0: stop */
int rt_scene_profile(RtScene* scene, int32_t enable);
/* synchronises the recorded events and returns the sums */
int rt_scene_profile_read(RtScene* scene, RtProfile* out);
/* runs the three micro-benchmarks on `stream` of the current device (a few ms each); synchronous */
int rt_measure_peaks(RtPeaks* out, void* stream);
const char* rt_last_error(void);
int rt_abi_version(void);

#ifdef __cplusplus
}
#endif
#endif /* B200RT_H */
