// rt_bvh.cuh - LBVH construction on the GPU (replaces the per-ray brute-force face loop set-up of
// mesh.py:72-77): 63-bit Morton codes of triangle centroids -> LSD radix sort (8 x 8 bits, hand
// written: per-tile histograms, one scan, stable scatter by warp match + per-round block offsets)
// -> Karras 2012 radix-tree hierarchy -> bottom-up AABB refit with atomic arrival counters ->
// collapse of every other level into 4-wide 128-byte traversal nodes (even-depth binary nodes survive; their
// children are the grandchildren, or a leaf child as is; compact indices by a prefix sum, so the layout is
// deterministic) + triangles re-laid in Morton order.
#pragma once
#include "rt_types.cuh"

// ---- Morton codes -------------------------------------------------------------------------------
RT_HD uint64_t expand21(uint64_t v) {   // spread the low 21 bits to every third bit
    v &= 0x1fffffull;
    v = (v | v << 32) & 0x1f00000000ffffull;
    v = (v | v << 16) & 0x1f0000ff0000ffull;
    v = (v | v << 8) & 0x100f00f00f00f00full;
    v = (v | v << 4) & 0x10c30c30c30c30c3ull;
    v = (v | v << 2) & 0x1249249249249249ull;
    return v;
}
// explicitly rounded single operations: nvcc would contract `centroid * (1/3) - lo` into one FMA, the host restatement
// (tests/hostsim, -ffp-contract=off) cannot, and keys one ulp apart order duplicate triangles differently
#if defined(__CUDA_ARCH__)
#define RT_SUB_RN(a, b) __fsub_rn((a), (b))
#define RT_MUL_RN(a, b) __fmul_rn((a), (b))
#define RT_ADD_RN(a, b) __fadd_rn((a), (b))
#else
#define RT_SUB_RN(a, b) ((a) - (b))
#define RT_MUL_RN(a, b) ((a) * (b))
#define RT_ADD_RN(a, b) ((a) + (b))
#endif
RT_HD float3 tri_centroid(float3 a, float3 b, float3 c) {
    const float k = 1.0f / 3.0f;
    return f3(RT_MUL_RN(RT_ADD_RN(RT_ADD_RN(a.x, b.x), c.x), k), RT_MUL_RN(RT_ADD_RN(RT_ADD_RN(a.y, b.y), c.y), k),
              RT_MUL_RN(RT_ADD_RN(RT_ADD_RN(a.z, b.z), c.z), k));
}
RT_HD uint64_t morton63(float3 c, float3 lo, float3 inv_ext) {
    float x = fminf(fmaxf(RT_MUL_RN(RT_SUB_RN(c.x, lo.x), inv_ext.x), 0.0f), 1.0f);
    float y = fminf(fmaxf(RT_MUL_RN(RT_SUB_RN(c.y, lo.y), inv_ext.y), 0.0f), 1.0f);
    float z = fminf(fmaxf(RT_MUL_RN(RT_SUB_RN(c.z, lo.z), inv_ext.z), 0.0f), 1.0f);
    uint64_t xi = (uint64_t)fminf(x * 2097152.0f, 2097151.0f);
    uint64_t yi = (uint64_t)fminf(y * 2097152.0f, 2097151.0f);
    uint64_t zi = (uint64_t)fminf(z * 2097152.0f, 2097151.0f);
    return (expand21(xi) << 2) | (expand21(yi) << 1) | expand21(zi);
}

RT_HD float3 vert_at(const float* verts, int i) { return f3(verts[3 * (size_t)i], verts[3 * (size_t)i + 1], verts[3 * (size_t)i + 2]); }

// ---- Karras 2012: node i of the binary radix tree over n sorted keys ------------------------------------
RT_HD int clz64(uint64_t x) {
#if defined(__CUDA_ARCH__)
    return __clzll((long long)x);
#else
    return x ? __builtin_clzll(x) : 64;
#endif
}
RT_HD int clz32(uint32_t x) {
#if defined(__CUDA_ARCH__)
    return __clz((int)x);
#else
    return x ? __builtin_clz(x) : 32;
#endif
}
// common-prefix length of keys i and j; identical keys fall back to the index (unique) so the tree stays binary
RT_HD int lbvh_delta(const uint64_t* keys, int n, int i, int j) {
    if (j < 0 || j >= n) return -1;
    uint64_t a = keys[i], b = keys[j];
    if (a == b) return 64 + clz32((uint32_t)i ^ (uint32_t)j);
    return clz64(a ^ b);
}
// children: >= 0 internal node index, < 0 leaf ~slot
RT_HD void lbvh_node(const uint64_t* keys, int n, int i, int& left, int& right) {
    int d = (lbvh_delta(keys, n, i, i + 1) - lbvh_delta(keys, n, i, i - 1)) >= 0 ? 1 : -1;
    int dmin = lbvh_delta(keys, n, i, i - d);
    int lmax = 2;
    while (lbvh_delta(keys, n, i, i + lmax * d) > dmin) lmax *= 2;
    int l = 0;
    for (int t = lmax / 2; t >= 1; t /= 2)
        if (lbvh_delta(keys, n, i, i + (l + t) * d) > dmin) l += t;
    int j = i + l * d;
    int dnode = lbvh_delta(keys, n, i, j);
    int s = 0;
    for (int t = (l + 1) / 2;; t = (t + 1) / 2) {
        if (lbvh_delta(keys, n, i, i + (s + t) * d) > dnode) s += t;
        if (t <= 1) break;
    }
    int gamma = i + s * d + (d < 0 ? d : 0);
    int lo = i < j ? i : j, hi = i < j ? j : i;
    left = (lo == gamma) ? ~gamma : gamma;
    right = (hi == gamma + 1) ? ~(gamma + 1) : gamma + 1;
}

// ---- 4-wide collapse (shared by the GPU emit kernel, the host object-BVH builder and tests/hostsim) ----------
// The up-to-4 children of the traversal node rooted at binary node i: start from its two children and keep opening the
// internal child with the LARGEST surface area (ties: the earlier slot) into its own two children, in place, until four
// slots are used or only leaves are left.  Compared with taking the grandchildren blindly (depth parity) this fills
// 3.8 instead of 3.0 of 4 slots on the 1M-triangle mesh and opens the boxes a ray is most likely to hit first.
RT_HD float box_half_area(const float4& lo, const float4& hi) {
    const float dx = hi.x - lo.x, dy = hi.y - lo.y, dz = hi.z - lo.z;
    return dx * dy + dy * dz + dz * dx;
}
// leaf_tris > 1 (mesh LBVH only): a binary node whose subtree holds at most that many triangles is not opened but becomes
// ONE leaf slot over its (consecutive, Morton-ordered) triangle slots; the refit leaves (first slot, count) in the .w of
// the node's box corners.
RT_HD bool wide_openable(const float4* node_hi, int c, int leaf_tris) { return c >= 0 && (leaf_tris <= 1 || as_int(node_hi[c].w) > leaf_tris); }
RT_HD int wide_children(const int2* children, const float4* node_lo, const float4* node_hi, int i, int leaf_tris, int out[4]) {
    const int2 ch = children[i];
    out[0] = ch.x; out[1] = ch.y;
    int n = 2;
    while (n < 4) {
        int pick = -1; float best = -1.0f;
        for (int k = 0; k < n; k++)
            if (wide_openable(node_hi, out[k], leaf_tris)) { const float a = box_half_area(node_lo[out[k]], node_hi[out[k]]); if (a > best) { best = a; pick = k; } }
        if (pick < 0) break;
        const int2 g = children[out[pick]];
        for (int k = n; k > pick + 1; k--) out[k] = out[k - 1];
        out[pick] = g.x; out[pick + 1] = g.y;
        n++;
    }
    return n;
}
// which binary nodes become traversal nodes (host form: breadth-first from the root; the GPU does it in rounds, k_lbvh_mark)
inline void wide_mark_host(const int2* children, const float4* node_lo, const float4* node_hi, int n_internal, int leaf_tris, unsigned* is_wide) {
    for (int i = 0; i < n_internal; i++) is_wide[i] = 0;
    if (n_internal <= 0) return;
    int* queue = new int[n_internal];
    int qh = 0, qt = 0;
    queue[qt++] = 0; is_wide[0] = 1;
    while (qh < qt) {
        int c[4];
        const int n = wide_children(children, node_lo, node_hi, queue[qh++], leaf_tris, c);
        for (int k = 0; k < n; k++) if (wide_openable(node_hi, c[k], leaf_tris)) { is_wide[c[k]] = 1; queue[qt++] = c[k]; }
    }
    delete[] queue;
}
RT_HD void set4(float4& v, int k, float x) { if (k == 0) v.x = x; else if (k == 1) v.y = x; else if (k == 2) v.z = x; else v.w = x; }
RT_HD void set4i(int4& v, int k, int x) { if (k == 0) v.x = x; else if (k == 1) v.y = x; else if (k == 2) v.z = x; else v.w = x; }
// one traversal node from the binary tree: boxes padded by `pad`; remap: binary index -> compact wide index;
// leaf_flags may be NULL (mesh LBVH)
RT_HD BvhNode4 make_wide_node(const int2* children, int i, const unsigned* remap, const float4* leaf_lo, const float4* leaf_hi,
                              const float4* node_lo, const float4* node_hi, const int* leaf_flags, const int* node_flags, float pad,
                              int leaf_tris = 1) {
    BvhNode4 nd;
    const float inf = RT_INF;                  // empty slots: [+inf, -inf] boxes, missed by every ray (see RT_SLAB4)
    nd.lox = nd.loy = nd.loz = f4(inf, inf, inf, inf);
    nd.hix = nd.hiy = nd.hiz = f4(-inf, -inf, -inf, -inf);
    nd.link.x = nd.link.y = nd.link.z = nd.link.w = RT_LINK_NONE;
    nd.flags.x = nd.flags.y = nd.flags.z = nd.flags.w = 0;
    int c[4];
    const int n = wide_children(children, node_lo, node_hi, i, leaf_tris, c);
    for (int k = 0; k < n; k++) {
        const bool leaf = c[k] < 0;
        const float4 lo = leaf ? leaf_lo[~c[k]] : node_lo[c[k]], hi = leaf ? leaf_hi[~c[k]] : node_hi[c[k]];
        set4(nd.lox, k, lo.x - pad); set4(nd.loy, k, lo.y - pad); set4(nd.loz, k, lo.z - pad);
        set4(nd.hix, k, hi.x + pad); set4(nd.hiy, k, hi.y + pad); set4(nd.hiz, k, hi.z + pad);
        int link = leaf ? c[k] : (int)remap[c[k]];
        if (!leaf && !wide_openable(node_hi, c[k], leaf_tris))      // small subtree folded into one multi-triangle leaf
            link = ~(as_int(node_lo[c[k]].w) | ((as_int(node_hi[c[k]].w) - 1) << RT_LEAF_SLOT_BITS));
        set4i(nd.link, k, link);
        if (leaf_flags) set4i(nd.flags, k, leaf ? leaf_flags[~c[k]] : node_flags[c[k]]);
    }
    return nd;
}

// ---- PLOC: parallel locally-ordered clustering (Meister & Bittner 2018) ------------------------------------------------
// Bottom-up agglomeration over the Morton-ordered leaves instead of the Karras radix tree: every cluster of the current array
// looks RT_PLOC_RADIUS places to either side for the neighbour whose union with it has the smallest surface area, mutual
// nearest neighbours merge, the array is compacted (order kept), repeat until one cluster is left.  The tree follows the
// geometry instead of the bit pattern of the keys: on the 1 M-triangle bench mesh 4.4 % fewer node visits and 12 % fewer
// triangle tests than the radix tree under the same 4-wide collapse.  Every decision is made on explicitly rounded float
// operations (no FMA contraction) with index tie-breaks, so the GPU kernels and the sequential restatement below build the
// SAME tree, bit for bit.  Node numbering: the k-th node created gets index (n - 2 - k): the root is node 0 and children have
// larger indices than their parents.  A cluster code is >= 0 for a node, ~p < 0 for the leaf at Morton position p.
// RT_PLOC_RADIUS 0 selects the radix tree.
#ifndef RT_PLOC_RADIUS
#define RT_PLOC_RADIUS 8
#endif
RT_HD float ploc_cost(const float4& la, const float4& ha, const float4& lb, const float4& hb) {
    const float dx = RT_SUB_RN(fmaxf(ha.x, hb.x), fminf(la.x, lb.x)), dy = RT_SUB_RN(fmaxf(ha.y, hb.y), fminf(la.y, lb.y)),
                dz = RT_SUB_RN(fmaxf(ha.z, hb.z), fminf(la.z, lb.z));
    return RT_ADD_RN(RT_ADD_RN(RT_MUL_RN(dx, dy), RT_MUL_RN(dy, dz)), RT_MUL_RN(dz, dx));
}
// nearest neighbour of cluster i among [i - r, i + r] of the current array (m >= 2 clusters); ties -> the smallest index
RT_HD int ploc_nearest(const float4* lo, const float4* hi, int m, int i, int r) {
    const int j0 = i - r < 0 ? 0 : i - r, j1 = i + r >= m ? m - 1 : i + r;
    const float4 la = lo[i], ha = hi[i];
    int best = -1; float bc = 0.0f;
    for (int j = j0; j <= j1; j++) {
        if (j == i) continue;
        const float c = ploc_cost(la, ha, lo[j], hi[j]);
        if (best < 0 || c < bc) { bc = c; best = j; }
    }
    return best;
}
// .w of a cluster's hi corner: triangles under it (bit pattern)
RT_HD void ploc_union(const float4& la, const float4& ha, const float4& lb, const float4& hb, float4& lo, float4& hi) {
    lo = f4(fminf(la.x, lb.x), fminf(la.y, lb.y), fminf(la.z, lb.z), 0.0f);
    hi = f4(fmaxf(ha.x, hb.x), fmaxf(ha.y, hb.y), fmaxf(ha.z, hb.z), as_float(as_int(ha.w) + as_int(hb.w)));
}
// triangle slots of the two children of node `nd` whose own first slot is known (depth-first order, left child first: the triangles
// under any node are consecutive, which the multi-triangle leaves of the 4-wide collapse need); leaf children get their code
// rewritten from ~(Morton position) to ~slot
RT_HD void ploc_place_children(int2* children, float4* node_lo, const float4* node_hi, int* first, int* slot_of, int nd) {
    int2 ch = children[nd];
    const int f0 = first[nd];
    const int c0 = ch.x < 0 ? 1 : as_int(node_hi[ch.x].w);
    node_lo[nd].w = as_float(f0);
    if (ch.x < 0) { slot_of[~ch.x] = f0; ch.x = ~f0; } else first[ch.x] = f0;
    if (ch.y < 0) { slot_of[~ch.y] = f0 + c0; ch.y = ~(f0 + c0); } else first[ch.y] = f0 + c0;
    children[nd] = ch;
}
#if !defined(__CUDA_ARCH__)
// Sequential restatement (tests/hostsim: the checker of the GPU build).  leaf_lo/hi: boxes in Morton order.  Outputs the binary
// tree (children with leaf codes ~SLOT, boxes with .w = first slot / triangle count like the radix-tree refit leaves them) and
// slot_of[p]: the triangle slot of Morton position p.  Returns the number of clustering rounds.
inline int ploc_build_host(int n, const float4* leaf_lo, const float4* leaf_hi, int radius, int2* children, float4* node_lo, float4* node_hi, int* slot_of) {
    if (n < 2) { if (n == 1) slot_of[0] = 0; return 0; }
    int* cid = new int[n]; int* cid2 = new int[n]; int* nn = new int[n];
    float4* lo = new float4[n]; float4* hi = new float4[n]; float4* lo2 = new float4[n]; float4* hi2 = new float4[n];
    int* round_end = new int[n];                                         // nodes created up to and including round t
    for (int i = 0; i < n; i++) { cid[i] = ~i; lo[i] = leaf_lo[i]; hi[i] = leaf_hi[i]; lo[i].w = 0.0f; hi[i].w = as_float(1); }
    int m = n, created = 0, rounds = 0;
    while (m > 1) {
        for (int i = 0; i < m; i++) nn[i] = ploc_nearest(lo, hi, m, i, radius);
        int w = 0;
        for (int i = 0; i < m; i++) {
            const int j = nn[i];
            const bool mutual = nn[j] == i;
            if (mutual && j < i) continue;                               // merged into the cluster at j
            if (mutual) {
                const int id = n - 2 - created++;
                children[id].x = cid[i]; children[id].y = cid[j];
                ploc_union(lo[i], hi[i], lo[j], hi[j], lo2[w], hi2[w]);
                node_lo[id] = lo2[w]; node_hi[id] = hi2[w];
                cid2[w] = id;
            } else { cid2[w] = cid[i]; lo2[w] = lo[i]; hi2[w] = hi[i]; }
            w++;
        }
        int* t = cid; cid = cid2; cid2 = t;
        float4* tl = lo; lo = lo2; lo2 = tl; tl = hi; hi = hi2; hi2 = tl;
        m = w; round_end[rounds++] = created;
    }
    // triangle slots, top-down: the nodes of a round have their parents in later rounds
    int* first = new int[n - 1];
    first[0] = 0;
    for (int t = rounds - 1; t >= 0; t--)
        for (int k = (t ? round_end[t - 1] : 0); k < round_end[t]; k++) ploc_place_children(children, node_lo, node_hi, first, slot_of, n - 2 - k);
    delete[] cid; delete[] cid2; delete[] nn; delete[] lo; delete[] hi; delete[] lo2; delete[] hi2; delete[] first; delete[] round_end;
    return rounds;
}
#endif

#if defined(__CUDACC__)
// ---- kernels ------------------------------------------------------------------------------------------
__global__ void k_morton(const float* __restrict__ verts, const int* __restrict__ faces, int n, float3 lo, float3 inv_ext,
                         uint64_t* __restrict__ keys, int* __restrict__ vals) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float3 a = vert_at(verts, faces[3 * i]), b = vert_at(verts, faces[3 * i + 1]), c = vert_at(verts, faces[3 * i + 2]);
    keys[i] = morton63(tri_centroid(a, b, c), lo, inv_ext);
    vals[i] = i;
}

#define RS_THREADS 256
#define RS_ROUNDS 16
#define RS_TILE (RS_THREADS * RS_ROUNDS)

// per-tile digit histogram, hist[digit * n_tiles + tile]
__global__ void __launch_bounds__(RS_THREADS) k_rs_hist(const uint64_t* __restrict__ keys, int n, int shift, int n_tiles, unsigned* __restrict__ hist) {
    __shared__ unsigned h[256];
    h[threadIdx.x] = 0;
    __syncthreads();
    const int base = blockIdx.x * RS_TILE;
    for (int r = 0; r < RS_ROUNDS; r++) {
        int i = base + r * RS_THREADS + threadIdx.x;
        if (i < n) atomicAdd(&h[(unsigned)(keys[i] >> shift) & 255u], 1u);
    }
    __syncthreads();
    hist[threadIdx.x * n_tiles + blockIdx.x] = h[threadIdx.x];
}

// exclusive scan of `m` counters in place.  Three launches per level (the single 1024-thread block this replaces spent 336 us on
// the 65 536 per-tile digit counters of a 1 M-triangle sort, 15 times per build: a third of the LBVH set-up):
//   k_scan_reduce : sum of every SCAN_TILE-element tile -> part[tile]
//   (recursion)   : exclusive scan of part[] (one block when it fits, else another level)
//   k_scan_apply  : per-tile exclusive scan (warp shuffles + one shared pass) offset by part[tile]
#define SCAN_THREADS 256
#define SCAN_PER_THREAD 8
#define SCAN_TILE (SCAN_THREADS * SCAN_PER_THREAD)
__device__ __forceinline__ unsigned block_exclusive_scan(unsigned v, unsigned* total) {       // SCAN_THREADS threads
    __shared__ unsigned warp_sum[SCAN_THREADS / 32];
    __shared__ unsigned block_total;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    unsigned inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { unsigned u = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += u; }
    if (lane == 31) warp_sum[warp] = inc;
    __syncthreads();
    if (warp == 0) {
        unsigned w = lane < SCAN_THREADS / 32 ? warp_sum[lane] : 0, winc = w;
#pragma unroll
        for (int o = 1; o < SCAN_THREADS / 32; o <<= 1) { unsigned u = __shfl_up_sync(0xffffffffu, winc, o); if (lane >= o) winc += u; }
        if (lane < SCAN_THREADS / 32) warp_sum[lane] = winc - w;
        if (lane == SCAN_THREADS / 32 - 1) block_total = winc;
    }
    __syncthreads();
    const unsigned r = inc - v + warp_sum[warp];
    if (total) *total = block_total;
    __syncthreads();
    return r;
}
__global__ void __launch_bounds__(SCAN_THREADS) k_scan_reduce(const unsigned* __restrict__ data, int m, unsigned* __restrict__ part) {
    const int base = blockIdx.x * SCAN_TILE;
    unsigned s = 0;
#pragma unroll
    for (int k = 0; k < SCAN_PER_THREAD; k++) { const int i = base + k * SCAN_THREADS + threadIdx.x; if (i < m) s += data[i]; }
    unsigned tot;
    block_exclusive_scan(s, &tot);
    if (threadIdx.x == 0) part[blockIdx.x] = tot;
}
// part == NULL: a single tile, offset 0
__global__ void __launch_bounds__(SCAN_THREADS) k_scan_apply(unsigned* __restrict__ data, int m, const unsigned* __restrict__ part) {
    const int base = blockIdx.x * SCAN_TILE + threadIdx.x * SCAN_PER_THREAD;     // each thread owns SCAN_PER_THREAD consecutive elements
    unsigned v[SCAN_PER_THREAD], s = 0;
#pragma unroll
    for (int k = 0; k < SCAN_PER_THREAD; k++) { v[k] = base + k < m ? data[base + k] : 0; s += v[k]; }
    unsigned run = block_exclusive_scan(s, nullptr) + (part ? part[blockIdx.x] : 0u);
#pragma unroll
    for (int k = 0; k < SCAN_PER_THREAD; k++) { if (base + k < m) data[base + k] = run; run += v[k]; }
}
// `scratch` holds at least scan_scratch_elems(m) unsigned
inline size_t scan_scratch_elems(long long m) {
    size_t tot = 0;
    while (m > SCAN_TILE) { m = (m + SCAN_TILE - 1) / SCAN_TILE; tot += (size_t)m; }
    return tot + 1;
}
inline void exclusive_scan_u32(unsigned* data, int m, unsigned* scratch, cudaStream_t st) {
    if (m <= 0) return;
    if (m <= SCAN_TILE) { k_scan_apply<<<1, SCAN_THREADS, 0, st>>>(data, m, nullptr); return; }
    const int tiles = (m + SCAN_TILE - 1) / SCAN_TILE;
    k_scan_reduce<<<tiles, SCAN_THREADS, 0, st>>>(data, m, scratch);
    exclusive_scan_u32(scratch, tiles, scratch + tiles, st);
    k_scan_apply<<<tiles, SCAN_THREADS, 0, st>>>(data, m, scratch);
}

// stable scatter: rounds walk the tile in key order; inside a round, rank = (#equal digits in lower warps
// of this round) + (#equal digits in lower lanes), on top of the per-digit count of earlier rounds
__global__ void __launch_bounds__(RS_THREADS) k_rs_scatter(const uint64_t* __restrict__ keys, const int* __restrict__ vals, int n, int shift,
                                                           int n_tiles, const unsigned* __restrict__ offs,
                                                           uint64_t* __restrict__ keys_out, int* __restrict__ vals_out) {
    __shared__ unsigned warp_cnt[RS_THREADS / 32][256];
    __shared__ unsigned digit_base[256];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    digit_base[threadIdx.x] = offs[threadIdx.x * n_tiles + blockIdx.x];
    const int base = blockIdx.x * RS_TILE;
    for (int r = 0; r < RS_ROUNDS; r++) {
        for (int w = 0; w < RS_THREADS / 32; w++) warp_cnt[w][threadIdx.x] = 0;
        __syncthreads();
        const int i = base + r * RS_THREADS + threadIdx.x;
        const bool valid = i < n;
        uint64_t key = valid ? keys[i] : 0;
        const unsigned digit = valid ? ((unsigned)(key >> shift) & 255u) : 256u + (unsigned)lane;   // invalid lanes match nobody
        const unsigned peers = __match_any_sync(0xffffffffu, digit);
        const unsigned rank = __popc(peers & ((1u << lane) - 1));
        if (valid && rank == 0) warp_cnt[warp][digit] = __popc(peers);
        __syncthreads();
        {   // thread d turns the per-warp counts of digit d into exclusive offsets and advances the digit base
            unsigned run = digit_base[threadIdx.x];
            for (int w = 0; w < RS_THREADS / 32; w++) { unsigned c = warp_cnt[w][threadIdx.x]; warp_cnt[w][threadIdx.x] = run; run += c; }
            digit_base[threadIdx.x] = run;
        }
        __syncthreads();
        if (valid) {
            const unsigned dst = warp_cnt[warp][digit] + rank;
            keys_out[dst] = key;
            vals_out[dst] = vals[i];
        }
        __syncthreads();
    }
}

__global__ void k_lbvh_hierarchy(const uint64_t* __restrict__ keys, int n, int2* __restrict__ children, int* __restrict__ parent_node,
                                 int* __restrict__ parent_leaf) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n - 1) return;
    int l, r;
    lbvh_node(keys, n, i, l, r);
    children[i] = make_int2(l, r);
    if (l < 0) parent_leaf[~l] = i; else parent_node[l] = i;
    if (r < 0) parent_leaf[~r] = i; else parent_node[r] = i;
    if (i == 0) parent_node[0] = -1;
}

// leaves: triangles in Morton order + their boxes; then walk up, second arrival merges the children
__global__ void k_lbvh_leaves_refit(const float* __restrict__ verts, const int* __restrict__ faces, const int* __restrict__ order, int n,
                                    const int2* __restrict__ children, const int* __restrict__ parent_node, const int* __restrict__ parent_leaf,
                                    float4* __restrict__ tri, int4* __restrict__ face_vidx,
                                    float4* __restrict__ leaf_lo, float4* __restrict__ leaf_hi,
                                    float4* __restrict__ node_lo, float4* __restrict__ node_hi, unsigned* __restrict__ arrive) {
    int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n) return;
    const int f = order[s];
    const int i0 = faces[3 * f], i1 = faces[3 * f + 1], i2 = faces[3 * f + 2];
    float3 a = vert_at(verts, i0), b = vert_at(verts, i1), c = vert_at(verts, i2);
    tri[3 * (size_t)s] = f4(a.x, a.y, a.z, as_float(f));
    tri[3 * (size_t)s + 1] = f4(b.x, b.y, b.z, 0.0f);
    tri[3 * (size_t)s + 2] = f4(c.x, c.y, c.z, 0.0f);
    face_vidx[s] = make_int4(i0, i1, i2, f);
    float3 lo = f3(fminf(a.x, fminf(b.x, c.x)), fminf(a.y, fminf(b.y, c.y)), fminf(a.z, fminf(b.z, c.z)));
    float3 hi = f3(fmaxf(a.x, fmaxf(b.x, c.x)), fmaxf(a.y, fmaxf(b.y, c.y)), fmaxf(a.z, fmaxf(b.z, c.z)));
    leaf_lo[s] = f4(lo.x, lo.y, lo.z, 0.0f);
    leaf_hi[s] = f4(hi.x, hi.y, hi.z, 0.0f);
    if (n < 2) return;
    __threadfence();
    int node = parent_leaf[s];
    while (node >= 0) {
        if (atomicAdd(&arrive[node], 1u) == 0) return;     // first child to arrive stops; the second merges
        __threadfence();
        int2 ch = children[node];
        float4 l0, h0, l1, h1;                             // L1 is not coherent: read the siblings' boxes from L2
        if (ch.x < 0) { l0 = __ldcg(leaf_lo + ~ch.x); h0 = __ldcg(leaf_hi + ~ch.x); } else { l0 = __ldcg(node_lo + ch.x); h0 = __ldcg(node_hi + ch.x); }
        if (ch.y < 0) { l1 = __ldcg(leaf_lo + ~ch.y); h1 = __ldcg(leaf_hi + ~ch.y); } else { l1 = __ldcg(node_lo + ch.y); h1 = __ldcg(node_hi + ch.y); }
        // .w of the corners: first triangle slot and number of triangles under the node (bit patterns)
        const int f0 = ch.x < 0 ? ~ch.x : __float_as_int(l0.w), f1 = ch.y < 0 ? ~ch.y : __float_as_int(l1.w);
        const int n0 = ch.x < 0 ? 1 : __float_as_int(h0.w), n1 = ch.y < 0 ? 1 : __float_as_int(h1.w);
        node_lo[node] = f4(fminf(l0.x, l1.x), fminf(l0.y, l1.y), fminf(l0.z, l1.z), __int_as_float(min(f0, f1)));
        node_hi[node] = f4(fmaxf(h0.x, h1.x), fmaxf(h0.y, h1.y), fmaxf(h0.z, h1.z), __int_as_float(n0 + n1));
        __threadfence();
        node = parent_node[node];
    }
}

// Which binary nodes become traversal nodes, top-down in rounds: a node marked in round r-1 marks the internal ones among
// its wide children in round r (mark = round + 1, 0 = unmarked; the root is marked by the host before round 1).
__global__ void k_lbvh_mark(int n, const int2* __restrict__ children, const float4* __restrict__ node_lo, const float4* __restrict__ node_hi,
                            unsigned* __restrict__ mark, unsigned round, unsigned* __restrict__ changed) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n - 1 || mark[i] != round) return;
    int c[4];
    const int nc = wide_children(children, node_lo, node_hi, i, RT_LEAF_TRIS, c);
    bool any = false;
    for (int k = 0; k < nc; k++) if (wide_openable(node_hi, c[k], RT_LEAF_TRIS)) { mark[c[k]] = round + 1; any = true; }
    if (any) *changed = 1u;
}
// marks -> 0/1 flags (`even` keeps its name from the depth-parity collapse it replaces) and the scan input
__global__ void k_lbvh_flags(int n, const unsigned* __restrict__ mark, unsigned* __restrict__ even, unsigned* __restrict__ remap) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n - 1) return;
    const unsigned e = mark[i] ? 1u : 0u;
    even[i] = e; remap[i] = e;          // remap is turned into compact indices by an exclusive scan
}

// final 128-byte 4-wide traversal nodes (padded boxes of the up-to-4 children in SoA)
__global__ void k_lbvh_emit4(int n, const int2* __restrict__ children, const unsigned* __restrict__ even, const unsigned* __restrict__ remap,
                             const float4* __restrict__ leaf_lo, const float4* __restrict__ leaf_hi,
                             const float4* __restrict__ node_lo, const float4* __restrict__ node_hi, float pad, BvhNode4* __restrict__ out) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n - 1 || !even[i]) return;
    out[remap[i]] = make_wide_node(children, i, remap, leaf_lo, leaf_hi, node_lo, node_hi, nullptr, nullptr, pad, RT_LEAF_TRIS);
}
// ---- PLOC on the device (one round = nearest neighbours, merge / keep flags, two prefix sums, apply; the host reads the
// number of merges of the round back).  Same functions, same order of decisions as ploc_build_host.
__global__ void k_ploc_leaf_boxes(const float* __restrict__ verts, const int* __restrict__ faces, const int* __restrict__ order, int n,
                                  int* __restrict__ cid, float4* __restrict__ lo, float4* __restrict__ hi) {
    int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n) return;
    const int f = order[p];
    float3 a = vert_at(verts, faces[3 * f]), b = vert_at(verts, faces[3 * f + 1]), c = vert_at(verts, faces[3 * f + 2]);
    cid[p] = ~p;
    lo[p] = f4(fminf(a.x, fminf(b.x, c.x)), fminf(a.y, fminf(b.y, c.y)), fminf(a.z, fminf(b.z, c.z)), 0.0f);
    hi[p] = f4(fmaxf(a.x, fmaxf(b.x, c.x)), fmaxf(a.y, fmaxf(b.y, c.y)), fmaxf(a.z, fmaxf(b.z, c.z)), __int_as_float(1));
}
__global__ void k_ploc_nn(const float4* __restrict__ lo, const float4* __restrict__ hi, int m, int r, int* __restrict__ nn) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < m) nn[i] = ploc_nearest(lo, hi, m, i, r);
}
// merge: this cluster absorbs its mutual nearest neighbour (the lower position of the pair); keep: it is still there next round
__global__ void k_ploc_flags(const int* __restrict__ nn, int m, unsigned* __restrict__ merge, unsigned* __restrict__ keep,
                             unsigned* __restrict__ merge_scan, unsigned* __restrict__ keep_scan) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= m) return;
    const int j = nn[i];
    const bool mutual = nn[j] == i;
    const unsigned mg = (mutual && i < j) ? 1u : 0u, kp = (mutual && j < i) ? 0u : 1u;
    merge[i] = mg; keep[i] = kp; merge_scan[i] = mg; keep_scan[i] = kp;
}
__global__ void k_ploc_apply(int n, int m, int created, const int* __restrict__ nn, const unsigned* __restrict__ merge, const unsigned* __restrict__ keep,
                             const unsigned* __restrict__ merge_scan, const unsigned* __restrict__ keep_scan,
                             const int* __restrict__ cid, const float4* __restrict__ lo, const float4* __restrict__ hi,
                             int* __restrict__ cid2, float4* __restrict__ lo2, float4* __restrict__ hi2,
                             int2* __restrict__ children, float4* __restrict__ node_lo, float4* __restrict__ node_hi) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= m || !keep[i]) return;
    const int w = (int)keep_scan[i];
    if (merge[i]) {
        const int j = nn[i];
        const int id = n - 2 - (created + (int)merge_scan[i]);
        float4 l, h;
        ploc_union(lo[i], hi[i], lo[j], hi[j], l, h);
        children[id] = make_int2(cid[i], cid[j]);
        node_lo[id] = l; node_hi[id] = h;
        cid2[w] = id; lo2[w] = l; hi2[w] = h;
    } else { cid2[w] = cid[i]; lo2[w] = lo[i]; hi2[w] = hi[i]; }
}
// triangle slots of the children of the nodes created as numbers [k0, k1) (one clustering round: their parents came later)
__global__ void k_ploc_place(int n, int k0, int k1, int2* __restrict__ children, float4* __restrict__ node_lo, const float4* __restrict__ node_hi,
                             int* __restrict__ first, int* __restrict__ slot_of) {
    int k = k0 + blockIdx.x * blockDim.x + threadIdx.x;
    if (k < k1) ploc_place_children(children, node_lo, node_hi, first, slot_of, n - 2 - k);
}
// triangles, vertex ids, leaf boxes and the slot -> face table, laid out by slot
__global__ void k_ploc_emit_leaves(const float* __restrict__ verts, const int* __restrict__ faces, const int* __restrict__ order_m,
                                   const int* __restrict__ slot_of, int n, float4* __restrict__ tri, int4* __restrict__ face_vidx,
                                   float4* __restrict__ leaf_lo, float4* __restrict__ leaf_hi, int* __restrict__ order_out) {
    int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n) return;
    const int f = order_m[p], s = slot_of[p];
    const int i0 = faces[3 * f], i1 = faces[3 * f + 1], i2 = faces[3 * f + 2];
    float3 a = vert_at(verts, i0), b = vert_at(verts, i1), c = vert_at(verts, i2);
    tri[3 * (size_t)s] = f4(a.x, a.y, a.z, as_float(f));
    tri[3 * (size_t)s + 1] = f4(b.x, b.y, b.z, 0.0f);
    tri[3 * (size_t)s + 2] = f4(c.x, c.y, c.z, 0.0f);
    face_vidx[s] = make_int4(i0, i1, i2, f);
    leaf_lo[s] = f4(fminf(a.x, fminf(b.x, c.x)), fminf(a.y, fminf(b.y, c.y)), fminf(a.z, fminf(b.z, c.z)), 0.0f);
    leaf_hi[s] = f4(fmaxf(a.x, fmaxf(b.x, c.x)), fmaxf(a.y, fmaxf(b.y, c.y)), fmaxf(a.z, fmaxf(b.z, c.z)), 0.0f);
    order_out[s] = f;
}
#endif  // __CUDACC__
