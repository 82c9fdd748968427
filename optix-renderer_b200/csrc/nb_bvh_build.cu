/* nb_bvh_build.cu — K_bvh_build: the BVH built on the device (nb_build_bvh mode NB_BVH_DEVICE).
 *
 * The reference builds its tree on the host (BVHBuildTask, src/utils/bvh.cpp:100-380: binned SAH,
 * exact sweep below 32 primitives); libnorib200_host restates that builder bit for bit and
 * NB_BVH_IMPORT takes its output. This mode needs no host tree at all: a linear BVH over the
 * primitives' Morton order, every stage a data-parallel kernel,
 *   bounds + centroids -> 63-bit Morton keys -> radix sort (cub) -> radix-tree topology
 *   (Karras 2012) -> bottom-up box refit -> subtrees of <= NB_LBVH_LEAF primitives collapsed into
 *   leaves -> pre-order numbering by one more sort (pre-order = ascending first primitive,
 *   larger range first) -> skip links by binary search -> nodes + slots in the layout of
 *   nb_internal.h.
 * The traversal kernels are unchanged. A different tree visits primitives in a different order,
 * so results can differ from the reference's only where its own order matters: two primitives hit
 * at exactly the same t (it keeps the one later in ITS leaf order), or a hit exactly on a box face
 * (boxes here are padded by a few ulps, the reference's are exact). tests/test_bvh_build_gpu.py
 * counts both on the oracle ray sets. */
#include "nb_internal.h"
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>
#include <algorithm>

#define NB_LBVH_LEAF 4u

struct NbBuildPrim { float3 bmin, bmax, centroid; };

__device__ __forceinline__ float3 f3min(float3 a, float3 b) { return make_float3(fminf(a.x, b.x), fminf(a.y, b.y), fminf(a.z, b.z)); }
__device__ __forceinline__ float3 f3max(float3 a, float3 b) { return make_float3(fmaxf(a.x, b.x), fmaxf(a.y, b.y), fmaxf(a.z, b.z)); }
/* order-preserving float <-> uint, for atomicMin/atomicMax on floats */
__device__ __forceinline__ uint32_t nb_f2o(float f) { const uint32_t u = __float_as_uint(f); return (u & 0x80000000u) ? ~u : (u | 0x80000000u); }
__device__ __forceinline__ float nb_o2f(uint32_t u) { return __uint_as_float((u & 0x80000000u) ? (u & 0x7fffffffu) : ~u); }

__device__ __forceinline__ float3 nb_ldcg3(const float3 *p) {
    const float *f = (const float *) p;
    return make_float3(__ldcg(f), __ldcg(f + 1), __ldcg(f + 2));
}

__device__ __forceinline__ void nb_global_prim(const NbDeviceScene &sc, const uint32_t *prim_offset, uint32_t g, uint32_t &sid, uint32_t &local) {
    uint32_t lo = 0, hi = sc.n_shapes;              /* BVH::findShape (bvh.h:105-109) */
    while (hi - lo > 1) { const uint32_t mid = (lo + hi) >> 1; if (prim_offset[mid] <= g) lo = mid; else hi = mid; }
    sid = lo; local = g - prim_offset[lo];
}

/* stage 1: per-primitive bounds (Mesh::getBoundingBox mesh.cpp:198-204, Sphere sphere.cpp:56-58),
 * centroid (mesh.cpp:206-212) and the bounds of all centroids */
__global__ void nb_k_build_bounds(NbDeviceScene sc, const uint32_t *prim_offset, uint32_t n, NbBuildPrim *prims, uint32_t *cbounds, uint32_t *pbounds = nullptr) {
    const uint32_t g = blockIdx.x * blockDim.x + threadIdx.x;
    float3 c = make_float3(0.f, 0.f, 0.f);
    if (g < n) {
        uint32_t sid, local;
        nb_global_prim(sc, prim_offset, g, sid, local);
        const nb_shape sh = sc.shapes[sid];
        NbBuildPrim p;
        if (sh.type == NB_SHAPE_SPHERE) {
            c = make_float3(sh.center[0], sh.center[1], sh.center[2]);
            p.bmin = make_float3(c.x - sh.radius, c.y - sh.radius, c.z - sh.radius);
            p.bmax = make_float3(c.x + sh.radius, c.y + sh.radius, c.z + sh.radius);
        } else {
            const uint32_t *f = sc.faces + ((size_t) sh.face_offset + local) * 3;
            const float *P = sc.positions;
            const size_t i0 = sh.vertex_offset + f[0], i1 = sh.vertex_offset + f[1], i2 = sh.vertex_offset + f[2];
            const float3 p0 = make_float3(P[3 * i0], P[3 * i0 + 1], P[3 * i0 + 2]);
            const float3 p1 = make_float3(P[3 * i1], P[3 * i1 + 1], P[3 * i1 + 2]);
            const float3 p2 = make_float3(P[3 * i2], P[3 * i2 + 1], P[3 * i2 + 2]);
            p.bmin = f3min(p0, f3min(p1, p2)); p.bmax = f3max(p0, f3max(p1, p2));
            c = make_float3((p0.x + p1.x + p2.x) * (1.f / 3.f), (p0.y + p1.y + p2.y) * (1.f / 3.f), (p0.z + p1.z + p2.z) * (1.f / 3.f));
        }
        p.centroid = c;
        prims[g] = p;
        atomicMin(&cbounds[0], nb_f2o(c.x)); atomicMin(&cbounds[1], nb_f2o(c.y)); atomicMin(&cbounds[2], nb_f2o(c.z));
        atomicMax(&cbounds[3], nb_f2o(c.x)); atomicMax(&cbounds[4], nb_f2o(c.y)); atomicMax(&cbounds[5], nb_f2o(c.z));
        if (pbounds) {                                   /* the root's box (NB_BVH_DEVICE_SAH), reduced per warp first */
            uint32_t lo[3] = { nb_f2o(p.bmin.x), nb_f2o(p.bmin.y), nb_f2o(p.bmin.z) }, hi[3] = { nb_f2o(p.bmax.x), nb_f2o(p.bmax.y), nb_f2o(p.bmax.z) };
            const uint32_t live = __activemask();
            for (int a = 0; a < 3; ++a) { lo[a] = __reduce_min_sync(live, lo[a]); hi[a] = __reduce_max_sync(live, hi[a]); }
            if ((threadIdx.x & 31u) == (uint32_t) (__ffs(live) - 1)) for (int a = 0; a < 3; ++a) { atomicMin(&pbounds[a], lo[a]); atomicMax(&pbounds[3 + a], hi[a]); }
        }
    }
}

__device__ __forceinline__ unsigned long long nb_expand21(uint32_t v) {      /* ..cba -> ..c00b00a */
    unsigned long long x = v & 0x1fffffull;
    x = (x | x << 32) & 0x1f00000000ffffull;
    x = (x | x << 16) & 0x1f0000ff0000ffull;
    x = (x | x << 8) & 0x100f00f00f00f00full;
    x = (x | x << 4) & 0x10c30c30c30c30c3ull;
    x = (x | x << 2) & 0x1249249249249249ull;
    return x;
}

/* stage 2: 63-bit Morton key of the centroid inside the centroid bounds */
__global__ void nb_k_build_morton(const NbBuildPrim *prims, const uint32_t *cbounds, uint32_t n, unsigned long long *keys, uint32_t *ids) {
    const uint32_t g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= n) return;
    const float3 lo = make_float3(nb_o2f(cbounds[0]), nb_o2f(cbounds[1]), nb_o2f(cbounds[2]));
    const float3 hi = make_float3(nb_o2f(cbounds[3]), nb_o2f(cbounds[4]), nb_o2f(cbounds[5]));
    const float3 c = prims[g].centroid;
    const float sx = hi.x > lo.x ? 2097151.f / (hi.x - lo.x) : 0.f, sy = hi.y > lo.y ? 2097151.f / (hi.y - lo.y) : 0.f,
                sz = hi.z > lo.z ? 2097151.f / (hi.z - lo.z) : 0.f;
    const uint32_t x = (uint32_t) fminf(fmaxf((c.x - lo.x) * sx, 0.f), 2097151.f);
    const uint32_t y = (uint32_t) fminf(fmaxf((c.y - lo.y) * sy, 0.f), 2097151.f);
    const uint32_t z = (uint32_t) fminf(fmaxf((c.z - lo.z) * sz, 0.f), 2097151.f);
    keys[g] = (nb_expand21(x) << 2) | (nb_expand21(y) << 1) | nb_expand21(z);
    ids[g] = g;
}

/* length of the common prefix of the keys at sorted positions i and j; equal keys are told apart
 * by their position (Karras 2012, section 4) */
__device__ __forceinline__ int nb_delta(const unsigned long long *keys, int n, int i, int j) {
    if (j < 0 || j >= n) return -1;
    const unsigned long long a = keys[i], b = keys[j];
    if (a == b) return 64 + __clz((uint32_t) i ^ (uint32_t) j);
    return __clzll((long long) (a ^ b));
}

/* stage 3: topology of the binary radix tree. Internal node i in [0, n-2]; child ids >= n-1 are
 * leaves (sorted position = id - (n-1)). */
__global__ void nb_k_build_topology(const unsigned long long *keys, int n, uint32_t *left, uint32_t *right, uint32_t *first,
                                    uint32_t *last, uint32_t *parent /* 2n-1 */) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n - 1) return;
    const int d = nb_delta(keys, n, i, i + 1) - nb_delta(keys, n, i, i - 1) >= 0 ? 1 : -1;
    const int dmin = nb_delta(keys, n, i, i - d);
    int lmax = 2;
    while (nb_delta(keys, n, i, i + lmax * d) > dmin) lmax *= 2;
    int l = 0;
    for (int t = lmax >> 1; t >= 1; t >>= 1)
        if (nb_delta(keys, n, i, i + (l + t) * d) > dmin) l += t;
    const int j = i + l * d;
    const int dnode = nb_delta(keys, n, i, j);
    int s = 0;
    for (int t = (l + 1) >> 1; ; t = (t + 1) >> 1) {
        if (nb_delta(keys, n, i, i + (s + t) * d) > dnode) s += t;
        if (t <= 1) break;
    }
    const int gamma = i + s * d + min(d, 0);
    const int lo = min(i, j), hi = max(i, j);
    const uint32_t lc = lo == gamma ? (uint32_t) (n - 1 + gamma) : (uint32_t) gamma;
    const uint32_t rc = hi == gamma + 1 ? (uint32_t) (n - 1 + gamma + 1) : (uint32_t) (gamma + 1);
    left[i] = lc; right[i] = rc; first[i] = (uint32_t) lo; last[i] = (uint32_t) hi;
    parent[lc] = (uint32_t) i; parent[rc] = (uint32_t) i;
    if (i == 0) parent[0] = 0xffffffffu;
}

/* stage 4: boxes bottom-up; the second child to arrive at a node continues upwards */
__global__ void nb_k_build_refit(const NbBuildPrim *prims, const uint32_t *ids, int n, const uint32_t *left, const uint32_t *right,
                                 const uint32_t *parent, uint32_t *visits, float3 *bmin, float3 *bmax /* 2n-1 */) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n) return;
    const NbBuildPrim p = prims[ids[k]];
    uint32_t node = (uint32_t) (n - 1 + k);
    bmin[node] = p.bmin; bmax[node] = p.bmax;
    __threadfence();
    uint32_t up = parent[node];
    while (up != 0xffffffffu) {
        if (atomicAdd(&visits[up], 1u) == 0u) return;          /* first arrival: the sibling will finish the job */
        __threadfence();
        const uint32_t a = left[up], b = right[up];
        /* the children's boxes were written by other threads: read them from L2 */
        const float3 lo = f3min(nb_ldcg3(bmin + a), nb_ldcg3(bmin + b));
        const float3 hi = f3max(nb_ldcg3(bmax + a), nb_ldcg3(bmax + b));
        bmin[up] = lo; bmax[up] = hi;
        __threadfence();
        up = parent[up];
    }
}

/* stage 5: which nodes survive the leaf collapse, and their pre-order sort keys.
 * A subtree of <= NB_LBVH_LEAF primitives becomes one leaf; everything below it disappears. */
__global__ void nb_k_build_keep(int n, const uint32_t *first, const uint32_t *last, const uint32_t *parent,
                                unsigned long long *okeys, uint32_t *oids, uint32_t *n_kept) {
    const int id = blockIdx.x * blockDim.x + threadIdx.x;
    if (id >= 2 * n - 1) return;
    const bool is_prim = id >= n - 1;
    const uint32_t a = is_prim ? (uint32_t) (id - (n - 1)) : first[id], b = is_prim ? a : last[id];
    const uint32_t up = parent[id];
    const bool kept = up == 0xffffffffu || (last[up] - first[up] + 1u) > NB_LBVH_LEAF;
    okeys[id] = kept ? (((unsigned long long) a << 32) | (unsigned long long) (0xffffffffu - b)) : ~0ull;
    oids[id] = (uint32_t) id;
    if (kept) atomicAdd(n_kept, 1u);
}

/* stage 6: node records in pre-order (= sorted order of stage 5) */
__global__ void nb_k_build_emit_nodes(int n, uint32_t n_kept, const unsigned long long *skeys, const uint32_t *sids, const float3 *bmin,
                                      const float3 *bmax, float4 *nodes, uint32_t *leaf_last, const uint32_t *sah_split = nullptr) {
    const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n_kept) return;
    const unsigned long long key = skeys[p];
    const uint32_t a = (uint32_t) (key >> 32), b = 0xffffffffu - (uint32_t) key;
    const uint32_t id = sids[p];
    /* skip link: first node in pre-order whose range starts after this subtree */
    const unsigned long long want = (unsigned long long) (b + 1u) << 32;
    uint32_t lo = p + 1, hi = n_kept;
    while (lo < hi) { const uint32_t mid = (lo + hi) >> 1; if (skeys[mid] < want) lo = mid + 1; else hi = mid; }
    const uint32_t skip = lo >= n_kept ? NB_NODE_END : lo;
    const bool leaf = sah_split ? sah_split[id] == 0xffffffffu /* NB_SAH_LEAF */ : (id >= (uint32_t) (n - 1) || (b - a + 1u) <= NB_LBVH_LEAF);
    if (leaf) leaf_last[b] = 1u;
    float3 mn = bmin[id], mx = bmax[id];
    /* pad by a few ulps of the coordinates' magnitude: never tighter than the exact box */
    const float ex = 4e-7f * fmaxf(fabsf(mn.x), fabsf(mx.x)) + 1e-30f, ey = 4e-7f * fmaxf(fabsf(mn.y), fabsf(mx.y)) + 1e-30f,
                ez = 4e-7f * fmaxf(fabsf(mn.z), fabsf(mx.z)) + 1e-30f;
    mn = make_float3(mn.x - ex, mn.y - ey, mn.z - ez); mx = make_float3(mx.x + ex, mx.y + ey, mx.z + ez);
    uint32_t w1 = a;
    if (!leaf) {
        /* right child = first node after the left child's subtree; both children's (unpadded) boxes
         * give the axis of the ordered traversal */
        const unsigned long long lkey = skeys[p + 1];
        const unsigned long long rwant = (unsigned long long) ((0xffffffffu - (uint32_t) lkey) + 1u) << 32;
        uint32_t rl = p + 2, rh = n_kept;
        while (rl < rh) { const uint32_t mid = (rl + rh) >> 1; if (skeys[mid] < rwant) rl = mid + 1; else rh = mid; }
        const float3 lmn = bmin[sids[p + 1]], lmx = bmax[sids[p + 1]], rmn = bmin[sids[rl]], rmx = bmax[sids[rl]];
        const float l0[3] = { lmn.x, lmn.y, lmn.z }, l1[3] = { lmx.x, lmx.y, lmx.z }, r0[3] = { rmn.x, rmn.y, rmn.z }, r1[3] = { rmx.x, rmx.y, rmx.z };
        w1 = nb_inner_word(rl, l0, l1, r0, r1);
    }
    nodes[2 * (size_t) p] = make_float4(__uint_as_float(skip), __uint_as_float(w1), mn.x, mn.y);
    nodes[2 * (size_t) p + 1] = make_float4(mn.z, mx.x, mx.y, mx.z);
}

/* stage 7: primitive slots in Morton order */
__global__ void nb_k_build_emit_slots(NbDeviceScene sc, const uint32_t *prim_offset, const uint32_t *ids, const uint32_t *leaf_last,
                                      uint32_t n, float4 *slots, uint32_t *slot_tag, uint32_t *slot_prim) {
    const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n) return;
    uint32_t sid, local;
    nb_global_prim(sc, prim_offset, ids[k], sid, local);
    const nb_shape sh = sc.shapes[sid];
    float4 a = make_float4(0.f, 0.f, 0.f, 0.f), b = a, c = a;
    uint32_t tag = sid | (leaf_last[k] ? NB_SLOT_LAST : 0u);
    if (sh.type == NB_SHAPE_SPHERE) {
        a = make_float4(sh.center[0], sh.center[1], sh.center[2], sh.radius);
        tag |= NB_SLOT_SPHERE;
    } else {
        const uint32_t *f = sc.faces + ((size_t) sh.face_offset + local) * 3;
        const float *P = sc.positions;
        const size_t i0 = sh.vertex_offset + f[0], i1 = sh.vertex_offset + f[1], i2 = sh.vertex_offset + f[2];
        /* edge1 = p1 - p0, edge2 = p2 - p0 (mesh.cpp:107), rounded once each */
        a = make_float4(P[3 * i0], P[3 * i0 + 1], P[3 * i0 + 2], __fsub_rn(P[3 * i1], P[3 * i0]));
        b = make_float4(__fsub_rn(P[3 * i1 + 1], P[3 * i0 + 1]), __fsub_rn(P[3 * i1 + 2], P[3 * i0 + 2]),
                        __fsub_rn(P[3 * i2], P[3 * i0]), __fsub_rn(P[3 * i2 + 1], P[3 * i0 + 1]));
        c.x = __fsub_rn(P[3 * i2 + 2], P[3 * i0 + 2]);
    }
    c.y = __uint_as_float(tag); c.z = __uint_as_float(local);
    slots[3 * (size_t) k] = a; slots[3 * (size_t) k + 1] = b; slots[3 * (size_t) k + 2] = c;
    slot_tag[k] = tag; slot_prim[k] = local;
}

/* ---- NB_BVH_DEVICE_SAH: top-down binned SAH, one tree level per pass ------------------------------------------------------
 * The linear BVH above is built in 2 ms but renders 12-23 % slower than a SAH tree, and the host's SAH build (host/bvh.cpp, the
 * reference's algorithm) takes seconds for a million triangles. This builder is the SAH criterion of the reference
 * (bvh.cpp:100-233: cost 2 T + (nL SA(L) + nR SA(R)) / SA(node) against the leaf cost n, T = I = 1) run breadth-first on the
 * device: every level bins the primitives of every open node at once (16 bins per axis over the node's CENTROID bounds, all
 * three axes — the reference bins the largest axis of the node's box, which degenerates when a few large triangles span it
 * and sends whole subtrees to its O(n log^2 n) sweep), one thread per node picks the cheapest plane, a prefix sum makes the
 * partition stable and therefore deterministic, and the children's centroid bounds are reduced while the primitives move.
 * Leaves: a node becomes a leaf when no plane beats n and it holds at most NB_SAH_MAX_LEAF primitives; larger nodes that
 * cannot be split by cost are cut at the median bin, nodes whose centroids coincide are cut in half by position.
 * Output: the same node / slot arrays as the other builders, through the same emit kernels. */
#define NB_SAH_BINS 16
#define NB_SAH_MAX_LEAF 4u
#define NB_SAH_BIN_WORDS (3 * NB_SAH_BINS * 7)          /* per node: axis x bin x { min xyz, max xyz (order-preserving uints), count } */
#define NB_SAH_DONE 0x80000000u
#define NB_SAH_CHUNK 8                                   /* primitives per thread in the binning kernel */

struct NbSah {                                           /* node arrays (capacity 2n) and the per-level work lists */
    uint32_t *start, *count, *slot, *split, *left, *nl;  /* split: axis << 8 | bin, NB_SAH_HALF, or NB_SAH_LEAF */
    float3 *bmin, *bmax;
    uint32_t *cb;                                        /* 6 per node: centroid bounds, order-preserving uints */
    uint32_t *bins;                                      /* NB_SAH_BIN_WORDS per slot */
    uint32_t *active, *next_active;                      /* node ids open at this / the next level */
    uint32_t *counters;                                  /* [0] nodes allocated, [1] open nodes of the next level */
};
#define NB_SAH_LEAF 0xffffffffu
#define NB_SAH_HALF 0xfffffffeu

__device__ __forceinline__ int nb_sah_bin(float c, float lo, float hi) {
    if (!(hi > lo)) return 0;
    return min(NB_SAH_BINS - 1, max(0, (int) ((c - lo) * ((float) NB_SAH_BINS / (hi - lo)))));
}

__global__ void nb_k_sah_clear_bins(uint32_t *bins, uint32_t n_slots) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_slots * (uint32_t) NB_SAH_BIN_WORDS) return;
    bins[i] = (i % 7u) < 3u ? 0xffffffffu : 0u;
}

/* every primitive of an open node (slots [slot_lo, slot_hi) of this level) joins one bin per axis. A block's primitives mostly
 * belong to the node of its first one: those are gathered in shared memory and flushed once. */
__global__ void __launch_bounds__(256) nb_k_sah_bin(NbSah t, const NbBuildPrim *prims, const uint32_t *order, const uint32_t *pnode, uint32_t n,
                                                    uint32_t slot_lo, uint32_t slot_hi) {
    __shared__ uint32_t sb[NB_SAH_BIN_WORDS];
    __shared__ uint32_t s_node;
    const uint32_t base = blockIdx.x * (256u * NB_SAH_CHUNK);
    for (uint32_t w = threadIdx.x; w < (uint32_t) NB_SAH_BIN_WORDS; w += 256u) sb[w] = (w % 7u) < 3u ? 0xffffffffu : 0u;
    if (threadIdx.x == 0) s_node = base < n ? pnode[base] : NB_SAH_DONE;
    __syncthreads();
    const uint32_t node0 = s_node;
    for (uint32_t k = 0; k < (uint32_t) NB_SAH_CHUNK; ++k) {
        const uint32_t i = base + k * 256u + threadIdx.x;
        if (i >= n) break;
        const uint32_t node = pnode[i];
        if (node & NB_SAH_DONE) continue;
        const uint32_t slot = t.slot[node];
        if (slot < slot_lo || slot >= slot_hi) continue;
        const NbBuildPrim p = prims[order[i]];
        const uint32_t *cb = t.cb + 6 * (size_t) node;
        const float c[3] = { p.centroid.x, p.centroid.y, p.centroid.z };
        uint32_t *dst = node == node0 ? sb : t.bins + (size_t) (slot - slot_lo) * NB_SAH_BIN_WORDS;
        for (int a = 0; a < 3; ++a) {
            uint32_t *b = dst + (a * NB_SAH_BINS + nb_sah_bin(c[a], nb_o2f(cb[a]), nb_o2f(cb[3 + a]))) * 7;
            atomicMin(b + 0, nb_f2o(p.bmin.x)); atomicMin(b + 1, nb_f2o(p.bmin.y)); atomicMin(b + 2, nb_f2o(p.bmin.z));
            atomicMax(b + 3, nb_f2o(p.bmax.x)); atomicMax(b + 4, nb_f2o(p.bmax.y)); atomicMax(b + 5, nb_f2o(p.bmax.z));
            atomicAdd(b + 6, 1u);
        }
    }
    __syncthreads();
    if (node0 & NB_SAH_DONE) return;
    const uint32_t slot0 = t.slot[node0];
    if (slot0 < slot_lo || slot0 >= slot_hi) return;
    uint32_t *g = t.bins + (size_t) (slot0 - slot_lo) * NB_SAH_BIN_WORDS;
    for (uint32_t w = threadIdx.x; w < (uint32_t) NB_SAH_BIN_WORDS; w += 256u) {
        const uint32_t v = sb[w], kind = w % 7u;
        if (kind < 3u) { if (v != 0xffffffffu) atomicMin(g + w, v); }
        else if (kind < 6u) { if (v != 0u) atomicMax(g + w, v); }
        else if (v) atomicAdd(g + w, v);
    }
}

struct NbBox { float lo[3], hi[3]; };
__device__ __forceinline__ float nb_box_area(const NbBox &b) {                /* BoundingBox3f::getSurfaceArea */
    const float dx = b.hi[0] - b.lo[0], dy = b.hi[1] - b.lo[1], dz = b.hi[2] - b.lo[2];
    return 2.f * (dx * dy + dy * dz + dz * dx);
}
__device__ __forceinline__ void nb_box_join(NbBox &b, const uint32_t *bin) {
    for (int a = 0; a < 3; ++a) { b.lo[a] = fminf(b.lo[a], nb_o2f(bin[a])); b.hi[a] = fmaxf(b.hi[a], nb_o2f(bin[3 + a])); }
}

/* one thread per open node: the cheapest of 3 x 15 planes, the children */
__global__ void nb_k_sah_split(NbSah t, uint32_t n_active, uint32_t slot_lo, uint32_t slot_hi) {
    const uint32_t s = slot_lo + blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= slot_hi || s >= n_active) return;
    const uint32_t node = t.active[s], size = t.count[node];
    const uint32_t *bins = t.bins + (size_t) (s - slot_lo) * NB_SAH_BIN_WORDS;
    NbBox whole;
    { const float3 mn = t.bmin[node], mx = t.bmax[node]; whole.lo[0] = mn.x; whole.lo[1] = mn.y; whole.lo[2] = mn.z; whole.hi[0] = mx.x; whole.hi[1] = mx.y; whole.hi[2] = mx.z; }
    const float area = nb_box_area(whole), factor = area > 0.f ? 1.f / area : 0.f;
    float best = (float) size, best_any = INFINITY;          /* best: must beat the leaf; best_any: the cheapest plane that splits at all */
    uint32_t pick = NB_SAH_LEAF, pick_any = NB_SAH_LEAF, nl_pick = 0, nl_any = 0;
    for (int a = 0; a < 3; ++a) {
        const uint32_t *ab = bins + a * NB_SAH_BINS * 7;
        float right_area[NB_SAH_BINS];
        NbBox r; for (int k = 0; k < 3; ++k) { r.lo[k] = INFINITY; r.hi[k] = -INFINITY; }
        for (int i = NB_SAH_BINS - 1; i >= 1; --i) { if (ab[i * 7 + 6]) nb_box_join(r, ab + i * 7); right_area[i] = nb_box_area(r); }
        NbBox l; for (int k = 0; k < 3; ++k) { l.lo[k] = INFINITY; l.hi[k] = -INFINITY; }
        uint32_t nl = 0;
        for (int i = 0; i + 1 < NB_SAH_BINS; ++i) {          /* plane after bin i */
            if (ab[i * 7 + 6]) nb_box_join(l, ab + i * 7);
            nl += ab[i * 7 + 6];
            if (nl == 0 || nl == size) continue;
            const float cost = 2.f + factor * ((float) nl * nb_box_area(l) + (float) (size - nl) * right_area[i + 1]);
            if (cost < best) { best = cost; pick = ((uint32_t) a << 8) | (uint32_t) i; nl_pick = nl; }
            if (cost < best_any) { best_any = cost; pick_any = ((uint32_t) a << 8) | (uint32_t) i; nl_any = nl; }
        }
    }
    uint32_t nl = nl_pick;
    if (pick == NB_SAH_LEAF && size > NB_SAH_MAX_LEAF) {      /* too large for a leaf: the cheapest plane anyway, or halves by position */
        if (pick_any != NB_SAH_LEAF) { pick = pick_any; nl = nl_any; }
        else { pick = NB_SAH_HALF; nl = size / 2; }
    }
    t.split[node] = pick;
    if (pick == NB_SAH_LEAF) return;
    const uint32_t child = atomicAdd(t.counters, 2u);
    t.left[node] = child; t.nl[node] = nl;
    /* children's boxes: the bins on each side of the plane (exact); the rare positional cut keeps the parent's box */
    for (int side = 0; side < 2; ++side) {
        const uint32_t c = child + side;
        t.start[c] = t.start[node] + (side ? nl : 0u);
        t.count[c] = side ? size - nl : nl;
        NbBox b; for (int k = 0; k < 3; ++k) { b.lo[k] = INFINITY; b.hi[k] = -INFINITY; }
        if (pick != NB_SAH_HALF) {
            const uint32_t a = pick >> 8, plane = pick & 0xffu;
            const uint32_t *ab = bins + a * NB_SAH_BINS * 7;
            for (uint32_t i = side ? plane + 1 : 0u; i <= (side ? (uint32_t) NB_SAH_BINS - 1u : plane); ++i) if (ab[i * 7 + 6]) nb_box_join(b, ab + i * 7);
        } else b = whole;
        t.bmin[c] = make_float3(b.lo[0], b.lo[1], b.lo[2]); t.bmax[c] = make_float3(b.hi[0], b.hi[1], b.hi[2]);
        for (int k = 0; k < 3; ++k) { t.cb[6 * (size_t) c + k] = 0xffffffffu; t.cb[6 * (size_t) c + 3 + k] = 0u; }
        t.split[c] = NB_SAH_LEAF;
        if (t.count[c] > 1u) { const uint32_t at = atomicAdd(t.counters + 1, 1u); t.next_active[at] = c; t.slot[c] = at; }
    }
}

/* which side of its node's plane a primitive falls on (1 = left); primitives of closed nodes and of new leaves stay put */
__global__ void nb_k_sah_flag(NbSah t, const NbBuildPrim *prims, const uint32_t *order, const uint32_t *pnode, uint32_t n, uint32_t *flag) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t node = pnode[i];
    uint32_t f = 0u;
    if (!(node & NB_SAH_DONE)) {
        const uint32_t split = t.split[node];
        if (split == NB_SAH_HALF) f = (i - t.start[node]) < t.nl[node] ? 1u : 0u;
        else if (split != NB_SAH_LEAF) {
            const uint32_t a = split >> 8;
            const float3 c3 = prims[order[i]].centroid;
            const float c = a == 0 ? c3.x : a == 1 ? c3.y : c3.z;
            const uint32_t *cb = t.cb + 6 * (size_t) node;
            f = nb_sah_bin(c, nb_o2f(cb[a]), nb_o2f(cb[3 + a])) <= (int) (split & 0xffu) ? 1u : 0u;
        }
    }
    flag[i] = f;
}

/* stable partition of every split node at once: `before` = exclusive prefix sum of the flags over all positions */
__global__ void nb_k_sah_scatter(NbSah t, const NbBuildPrim *prims, const uint32_t *order, const uint32_t *pnode, uint32_t n, const uint32_t *flag,
                                 const uint32_t *before, uint32_t *order2, uint32_t *pnode2) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t node = pnode[i];
    const uint32_t split = (node & NB_SAH_DONE) ? NB_SAH_LEAF : t.split[node];
    if (split == NB_SAH_LEAF) { order2[i] = order[i]; pnode2[i] = node | NB_SAH_DONE; return; }
    const uint32_t start = t.start[node], nl = t.nl[node], left_before = before[i] - before[start];
    const bool left = flag[i] != 0u;
    const uint32_t pos = left ? start + left_before : start + nl + ((i - start) - left_before);
    const uint32_t child = t.left[node] + (left ? 0u : 1u);
    const uint32_t id = order[i];
    const bool open = t.count[child] > 1u;
    order2[pos] = id;
    pnode2[pos] = open ? child : (child | NB_SAH_DONE);
    if (!open) return;
    const float3 c = prims[id].centroid;                      /* the child's centroid bounds, for its own bins */
    uint32_t lo[3] = { nb_f2o(c.x), nb_f2o(c.y), nb_f2o(c.z) }, hi[3] = { lo[0], lo[1], lo[2] };
    /* the lanes of a warp that feed the same child reduce among themselves (a warp covers a few nodes at most: positions are
     * consecutive) and the first of them does the child's six atomics: the launch list showed this kernel at 78 % of the
     * build with one set of atomics per primitive */
    const uint32_t live = __activemask();
    const uint32_t same = __match_any_sync(live, child);
    for (int a = 0; a < 3; ++a) { lo[a] = __reduce_min_sync(same, lo[a]); hi[a] = __reduce_max_sync(same, hi[a]); }
    if ((threadIdx.x & 31u) != (uint32_t) (__ffs(same) - 1)) return;
    uint32_t *cb = t.cb + 6 * (size_t) child;
    for (int a = 0; a < 3; ++a) { atomicMin(cb + a, lo[a]); atomicMax(cb + 3 + a, hi[a]); }
}

__global__ void nb_k_sah_keys(const NbSah t, uint32_t n_nodes, unsigned long long *keys, uint32_t *ids) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_nodes) return;
    const uint32_t a = t.start[i], b = a + t.count[i] - 1u;
    keys[i] = ((unsigned long long) a << 32) | (unsigned long long) (0xffffffffu - b);
    ids[i] = i;
}

__global__ void nb_k_sah_init(NbSah t, const uint32_t *cbounds, const uint32_t *pbounds, uint32_t n, uint32_t *order, uint32_t *pnode) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) { order[i] = i; pnode[i] = n > 1u ? 0u : NB_SAH_DONE; }
    if (i == 0) {
        t.start[0] = 0u; t.count[0] = n; t.slot[0] = 0u; t.split[0] = NB_SAH_LEAF;
        t.bmin[0] = make_float3(nb_o2f(pbounds[0]), nb_o2f(pbounds[1]), nb_o2f(pbounds[2])); t.bmax[0] = make_float3(nb_o2f(pbounds[3]), nb_o2f(pbounds[4]), nb_o2f(pbounds[5]));
        for (int k = 0; k < 6; ++k) t.cb[k] = cbounds[k];
        t.active[0] = 0u; t.counters[0] = 1u; t.counters[1] = 0u;
    }
}

#define NB_TRY(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { rc = nb_fail(ctx, NB_E_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); goto done; } } while (0)

int nb_build_bvh_device(nb_ctx *ctx, int sah) {
    int rc = nb_render_finish(ctx);              /* queued subframes still walk the old tree */
    if (rc != NB_OK) return rc;
    const uint32_t n = ctx->h_shape_prim_offset.back();
    if (n == 0) return nb_fail(ctx, NB_E_INVALID, "nb_build_bvh: scene has no primitives");
    if (n >= (1u << 27)) return nb_fail(ctx, NB_E_UNSUPPORTED, "nb_build_bvh: too many primitives for the device builder");
    const uint32_t n_all = 2 * n - 1;
    cudaStream_t st = ctx->stream;
    std::vector<void *> tmp;
    auto alloc = [&](size_t bytes) -> void * { void *p = nullptr; if (cudaMalloc(&p, std::max<size_t>(bytes, 16)) != cudaSuccess) return nullptr; tmp.push_back(p); return p; };
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    uint32_t *d_off = (uint32_t *) alloc((ctx->dev.n_shapes + 1) * 4);
    NbBuildPrim *d_prims = (NbBuildPrim *) alloc((size_t) n * sizeof(NbBuildPrim));
    uint32_t *d_cb = (uint32_t *) alloc(8 * 4);
    unsigned long long *d_keys = (unsigned long long *) alloc((size_t) n_all * 8), *d_keys2 = (unsigned long long *) alloc((size_t) n_all * 8);
    uint32_t *d_ids = (uint32_t *) alloc((size_t) n_all * 4), *d_ids2 = (uint32_t *) alloc((size_t) n_all * 4);
    uint32_t *d_order = (uint32_t *) alloc((size_t) n_all * 4);      /* node ids in pre-order */
    uint32_t *d_left = (uint32_t *) alloc((size_t) n * 4), *d_right = (uint32_t *) alloc((size_t) n * 4);
    uint32_t *d_first = (uint32_t *) alloc((size_t) n * 4), *d_last = (uint32_t *) alloc((size_t) n * 4);
    uint32_t *d_parent = (uint32_t *) alloc((size_t) n_all * 4), *d_visits = (uint32_t *) alloc((size_t) n * 4);
    float3 *d_bmin = (float3 *) alloc((size_t) n_all * sizeof(float3)), *d_bmax = (float3 *) alloc((size_t) n_all * sizeof(float3));
    uint32_t *d_leaf_last = (uint32_t *) alloc((size_t) n * 4), *d_nkept = (uint32_t *) alloc(4);
    uint32_t *d_slot_tag = (uint32_t *) alloc((size_t) n * 4), *d_slot_prim = (uint32_t *) alloc((size_t) n * 4);
    void *d_sort_tmp = nullptr; size_t sort_bytes = 0, sort_bytes2 = 0;
    float4 *d_nodes = nullptr, *d_slots = nullptr, *d_wide = nullptr;
    uint32_t n_kept = 0, n_wide = 0;
    NbSah T; memset(&T, 0, sizeof(T));
    uint32_t *d_pnode[2] = { nullptr, nullptr }, *d_sorder[2] = { nullptr, nullptr }, *d_flag = nullptr, *d_before = nullptr, *d_pb = nullptr;
    void *d_scan_tmp = nullptr; size_t scan_bytes = 0;
    uint32_t sah_levels = 0;
    const uint32_t SLOT_CAP = 1u << 18;             /* open nodes binned at once: 352 MB of bins */
    const unsigned B = 256;
    if (!d_off || !d_prims || !d_cb || !d_keys || !d_keys2 || !d_ids || !d_ids2 || !d_order || !d_left || !d_right || !d_first || !d_last ||
        !d_parent || !d_visits || !d_bmin || !d_bmax || !d_leaf_last || !d_nkept || !d_slot_tag || !d_slot_prim) {
        rc = nb_fail(ctx, NB_E_NOMEM, "nb_build_bvh: out of device memory for %u primitives", n);
        goto done;
    }
    cub::DeviceRadixSort::SortPairs(nullptr, sort_bytes, d_keys, d_keys2, d_ids, d_ids2, (int) n, 0, 64, st);
    cub::DeviceRadixSort::SortPairs(nullptr, sort_bytes2, d_keys, d_keys2, d_ids, d_ids2, (int) n_all, 0, 64, st);
    d_sort_tmp = alloc(std::max(sort_bytes, sort_bytes2));
    if (!d_sort_tmp) { rc = nb_fail(ctx, NB_E_NOMEM, "nb_build_bvh: out of device memory (sort)"); goto done; }
    NB_TRY(cudaEventCreate(&e0)); NB_TRY(cudaEventCreate(&e1));
    NB_TRY(cudaMemcpyAsync(d_off, ctx->h_shape_prim_offset.data(), (ctx->dev.n_shapes + 1) * 4, cudaMemcpyHostToDevice, st));
    NB_TRY(cudaEventRecord(e0, st));
    {
        const uint32_t init[8] = { 0xffffffffu, 0xffffffffu, 0xffffffffu, 0u, 0u, 0u, 0u, 0u };
        NB_TRY(cudaMemcpyAsync(d_cb, init, sizeof(init), cudaMemcpyHostToDevice, st));
    }
    sort_bytes = std::max(sort_bytes, sort_bytes2);
    if (sah) {
        const uint32_t cap = n_all + 1;
        T.start = (uint32_t *) alloc((size_t) cap * 4); T.count = (uint32_t *) alloc((size_t) cap * 4); T.slot = (uint32_t *) alloc((size_t) cap * 4);
        T.split = (uint32_t *) alloc((size_t) cap * 4); T.left = (uint32_t *) alloc((size_t) cap * 4); T.nl = (uint32_t *) alloc((size_t) cap * 4);
        T.bmin = (float3 *) alloc((size_t) cap * sizeof(float3)); T.bmax = (float3 *) alloc((size_t) cap * sizeof(float3));
        T.cb = (uint32_t *) alloc((size_t) cap * 6 * 4);
        T.bins = (uint32_t *) alloc((size_t) std::min<uint32_t>(SLOT_CAP, n / 2 + 1) * NB_SAH_BIN_WORDS * 4);
        T.active = (uint32_t *) alloc((size_t) cap * 4); T.next_active = (uint32_t *) alloc((size_t) cap * 4);
        T.counters = (uint32_t *) alloc(16);
        d_pb = (uint32_t *) alloc(8 * 4);
        for (int k = 0; k < 2; ++k) { d_pnode[k] = (uint32_t *) alloc((size_t) n * 4); d_sorder[k] = (uint32_t *) alloc((size_t) n * 4); }
        d_flag = (uint32_t *) alloc((size_t) n * 4); d_before = (uint32_t *) alloc((size_t) n * 4);
        cub::DeviceScan::ExclusiveSum(nullptr, scan_bytes, d_flag, d_before, (int) n, st);
        d_scan_tmp = alloc(scan_bytes);
        if (!T.start || !T.count || !T.slot || !T.split || !T.left || !T.nl || !T.bmin || !T.bmax || !T.cb || !T.bins || !T.active || !T.next_active ||
            !T.counters || !d_pb || !d_pnode[0] || !d_pnode[1] || !d_sorder[0] || !d_sorder[1] || !d_flag || !d_before || !d_scan_tmp) {
            rc = nb_fail(ctx, NB_E_NOMEM, "nb_build_bvh: out of device memory for the SAH build of %u primitives", n);
            goto done;
        }
        {
            const uint32_t init[8] = { 0xffffffffu, 0xffffffffu, 0xffffffffu, 0u, 0u, 0u, 0u, 0u };
            NB_TRY(cudaMemcpyAsync(d_pb, init, sizeof(init), cudaMemcpyHostToDevice, st));
        }
        nb_k_build_bounds<<<(n + B - 1) / B, B, 0, st>>>(ctx->dev, d_off, n, d_prims, d_cb, d_pb);
        nb_k_sah_init<<<(n + B - 1) / B, B, 0, st>>>(T, d_cb, d_pb, n, d_sorder[0], d_pnode[0]);
        uint32_t n_active = n > 1 ? 1u : 0u, n_nodes = 1u;
        int cur = 0;
        for (; n_active && sah_levels < 512u; ++sah_levels) {
            for (uint32_t lo = 0; lo < n_active; lo += SLOT_CAP) {
                const uint32_t hi = std::min(n_active, lo + SLOT_CAP), ns = hi - lo;
                nb_k_sah_clear_bins<<<(ns * NB_SAH_BIN_WORDS + B - 1) / B, B, 0, st>>>(T.bins, ns);
                nb_k_sah_bin<<<(n + 256 * NB_SAH_CHUNK - 1) / (256 * NB_SAH_CHUNK), 256, 0, st>>>(T, d_prims, d_sorder[cur], d_pnode[cur], n, lo, hi);
                nb_k_sah_split<<<(ns + 127) / 128, 128, 0, st>>>(T, n_active, lo, hi);
            }
            nb_k_sah_flag<<<(n + B - 1) / B, B, 0, st>>>(T, d_prims, d_sorder[cur], d_pnode[cur], n, d_flag);
            NB_TRY(cub::DeviceScan::ExclusiveSum(d_scan_tmp, scan_bytes, d_flag, d_before, (int) n, st));
            nb_k_sah_scatter<<<(n + B - 1) / B, B, 0, st>>>(T, d_prims, d_sorder[cur], d_pnode[cur], n, d_flag, d_before, d_sorder[cur ^ 1], d_pnode[cur ^ 1]);
            cur ^= 1;
            uint32_t counters[2] = { 0, 0 };
            NB_TRY(cudaMemcpyAsync(counters, T.counters, 8, cudaMemcpyDeviceToHost, st));
            NB_TRY(cudaStreamSynchronize(st));
            n_nodes = counters[0]; n_active = counters[1];
            if (n_nodes > n_all) { rc = nb_fail(ctx, NB_E_CUDA, "nb_build_bvh: internal: SAH build made %u nodes for %u primitives", n_nodes, n); goto done; }
            std::swap(T.active, T.next_active);
            NB_TRY(cudaMemsetAsync(T.counters + 1, 0, 4, st));
        }
        /* pre-order numbering as in the linear build: sort the nodes by (first primitive, larger range first) */
        nb_k_sah_keys<<<(n_nodes + B - 1) / B, B, 0, st>>>(T, n_nodes, d_keys, d_ids);
        NB_TRY(cudaMemsetAsync(d_leaf_last, 0, (size_t) n * 4, st));
        NB_TRY(cub::DeviceRadixSort::SortPairs(d_sort_tmp, sort_bytes, d_keys, d_keys2, d_ids, d_order, (int) n_nodes, 0, 64, st));
        NB_TRY(cudaMemcpyAsync(d_ids2, d_sorder[cur], (size_t) n * 4, cudaMemcpyDeviceToDevice, st));
        NB_TRY(cudaStreamSynchronize(st));
        NB_TRY(cudaGetLastError());
        n_kept = n_nodes;
    } else {
    nb_k_build_bounds<<<(n + B - 1) / B, B, 0, st>>>(ctx->dev, d_off, n, d_prims, d_cb);
    nb_k_build_morton<<<(n + B - 1) / B, B, 0, st>>>(d_prims, d_cb, n, d_keys, d_ids);
    NB_TRY(cub::DeviceRadixSort::SortPairs(d_sort_tmp, sort_bytes, d_keys, d_keys2, d_ids, d_ids2, (int) n, 0, 64, st));
    /* sorted: d_keys2 (Morton), d_ids2 (global primitive id per sorted position) */
    NB_TRY(cudaMemsetAsync(d_visits, 0, (size_t) n * 4, st));
    NB_TRY(cudaMemsetAsync(d_leaf_last, 0, (size_t) n * 4, st));
    NB_TRY(cudaMemsetAsync(d_nkept, 0, 4, st));
    if (n > 1) nb_k_build_topology<<<(n - 1 + B - 1) / B, B, 0, st>>>(d_keys2, (int) n, d_left, d_right, d_first, d_last, d_parent);
    else { const uint32_t none = 0xffffffffu; NB_TRY(cudaMemcpyAsync(d_parent, &none, 4, cudaMemcpyHostToDevice, st)); }
    nb_k_build_refit<<<(n + B - 1) / B, B, 0, st>>>(d_prims, d_ids2, (int) n, d_left, d_right, d_parent, d_visits, d_bmin, d_bmax);
    nb_k_build_keep<<<(n_all + B - 1) / B, B, 0, st>>>((int) n, d_first, d_last, d_parent, d_keys, d_ids, d_nkept);
    NB_TRY(cub::DeviceRadixSort::SortPairs(d_sort_tmp, sort_bytes, d_keys, d_keys2, d_ids, d_order, (int) n_all, 0, 64, st));
    NB_TRY(cudaMemcpyAsync(&n_kept, d_nkept, 4, cudaMemcpyDeviceToHost, st));
    NB_TRY(cudaStreamSynchronize(st));
    NB_TRY(cudaGetLastError());
    }
    if (n_kept == 0 || n_kept > n_all) { rc = nb_fail(ctx, NB_E_CUDA, "nb_build_bvh: internal: %u nodes kept of %u", n_kept, n_all); goto done; }
    NB_TRY(cudaMalloc((void **) &d_nodes, (size_t) n_kept * 2 * sizeof(float4)));
    NB_TRY(cudaMalloc((void **) &d_slots, (size_t) n * 3 * sizeof(float4)));
    /* pre-order: d_keys2 now holds the node sort keys, d_order the node ids; d_ids2 still the Morton order (or the SAH builder's final order) */
    nb_k_build_emit_nodes<<<(n_kept + B - 1) / B, B, 0, st>>>((int) n, n_kept, d_keys2, d_order, sah ? T.bmin : d_bmin, sah ? T.bmax : d_bmax, d_nodes, d_leaf_last,
                                                             sah ? T.split : nullptr);
    nb_k_build_emit_slots<<<(n + B - 1) / B, B, 0, st>>>(ctx->dev, d_off, d_ids2, d_leaf_last, n, d_slots, d_slot_tag, d_slot_prim);
    NB_TRY(cudaEventRecord(e1, st));
    ctx->h_slot_tag.resize(n); ctx->h_slot_prim.resize(n);
    NB_TRY(cudaMemcpyAsync(ctx->h_slot_tag.data(), d_slot_tag, (size_t) n * 4, cudaMemcpyDeviceToHost, st));
    NB_TRY(cudaMemcpyAsync(ctx->h_slot_prim.data(), d_slot_prim, (size_t) n * 4, cudaMemcpyDeviceToHost, st));
    NB_TRY(cudaStreamSynchronize(st));
    NB_TRY(cudaGetLastError());
    {
        float ms = 0.f;
        cudaEventElapsedTime(&ms, e0, e1);
        ctx->bvh_build_ms = ms;
    }
    /* production tree: collapse the tree just built (host side for now: nodes come back once, 32 B each) */
    {
        std::vector<float4> h_nodes((size_t) n_kept * 2);
        NB_TRY(cudaMemcpyAsync(h_nodes.data(), d_nodes, h_nodes.size() * sizeof(float4), cudaMemcpyDeviceToHost, st));
        NB_TRY(cudaStreamSynchronize(st));
        std::vector<nb_bvh_node> bin(n_kept);
        for (uint32_t i = 0; i < n_kept; ++i) {
            const float4 A = h_nodes[2 * (size_t) i], B = h_nodes[2 * (size_t) i + 1];
            uint32_t w1; memcpy(&w1, &A.y, 4);
            nb_bvh_node &b = bin[i];
            if (w1 & NB_NODE_INNER_BIT) { b.word0 = 0u; b.word1 = w1 & NB_NODE_RIGHT_MASK; }
            else { b.word0 = 1u | ((w1 == NB_NODE_EMPTY_LEAF ? 0u : 1u) << 1); b.word1 = w1; }   /* leaf size: only "empty or not" matters to the collapse */
            b.bmin[0] = A.z; b.bmin[1] = A.w; b.bmin[2] = B.x; b.bmax[0] = B.y; b.bmax[1] = B.z; b.bmax[2] = B.w;
        }
        std::vector<float4> wide;
        uint32_t wide_depth = 0;
        nb_collapse_wide(bin.data(), n_kept, wide, wide_depth);
        NB_TRY(cudaMalloc((void **) &d_wide, std::max<size_t>(wide.size(), 1) * sizeof(float4)));
        NB_TRY(cudaMemcpyAsync(d_wide, wide.data(), wide.size() * sizeof(float4), cudaMemcpyHostToDevice, st));
        NB_TRY(cudaStreamSynchronize(st));
        n_wide = (uint32_t) (wide.size() / 8);
        ctx->wide_depth = wide_depth;
    }
    for (void *p : ctx->bvh_allocs) cudaFree(p);
    ctx->bvh_allocs.clear();
    ctx->bvh_allocs.push_back(d_nodes); ctx->bvh_allocs.push_back(d_slots); ctx->bvh_allocs.push_back(d_wide);
    ctx->dev.wide = d_wide; ctx->dev.n_wide = n_wide;
    d_wide = nullptr;
    ctx->dev.nodes = d_nodes; ctx->dev.slots = d_slots;
    ctx->dev.n_nodes = n_kept; ctx->dev.n_slots = n;
    ctx->dev.flat_in_params = 0;
#ifdef NB_PARAM_SLOTS
    if (n <= NB_FLAT_MAX_SLOTS) {
        NB_TRY(cudaMemcpy(ctx->dev.flat_slots, d_slots, (size_t) n * 3 * sizeof(float4), cudaMemcpyDeviceToHost));
        ctx->dev.flat_in_params = 1;
    }
#endif
    ctx->flat_ok = true;                         /* slots are emitted in the pre-order of the leaves */
    ctx->dev.flat = nb_scene_flat(ctx) ? 1u : 0u;
    nb_update_wide(ctx);
    d_nodes = nullptr; d_slots = nullptr;
    ctx->bvh_ready = true;
done:
    if (d_nodes) cudaFree(d_nodes);
    if (d_slots) cudaFree(d_slots);
    if (d_wide) cudaFree(d_wide);
    for (void *p : tmp) cudaFree(p);
    if (e0) cudaEventDestroy(e0);
    if (e1) cudaEventDestroy(e1);
    return rc;
}
