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
__global__ void nb_k_build_bounds(NbDeviceScene sc, const uint32_t *prim_offset, uint32_t n, NbBuildPrim *prims, uint32_t *cbounds) {
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
                                      const float3 *bmax, float4 *nodes, uint32_t *leaf_last) {
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
    const bool leaf = id >= (uint32_t) (n - 1) || (b - a + 1u) <= NB_LBVH_LEAF;
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

#define NB_TRY(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { rc = nb_fail(ctx, NB_E_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); goto done; } } while (0)

int nb_build_bvh_device(nb_ctx *ctx) {
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
    nb_k_build_bounds<<<(n + B - 1) / B, B, 0, st>>>(ctx->dev, d_off, n, d_prims, d_cb);
    nb_k_build_morton<<<(n + B - 1) / B, B, 0, st>>>(d_prims, d_cb, n, d_keys, d_ids);
    sort_bytes = std::max(sort_bytes, sort_bytes2);
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
    if (n_kept == 0 || n_kept > n_all) { rc = nb_fail(ctx, NB_E_CUDA, "nb_build_bvh: internal: %u nodes kept of %u", n_kept, n_all); goto done; }
    NB_TRY(cudaMalloc((void **) &d_nodes, (size_t) n_kept * 2 * sizeof(float4)));
    NB_TRY(cudaMalloc((void **) &d_slots, (size_t) n * 3 * sizeof(float4)));
    /* pre-order: d_keys2 now holds the node sort keys, d_order the node ids; d_ids2 still the Morton order */
    nb_k_build_emit_nodes<<<(n_kept + B - 1) / B, B, 0, st>>>((int) n, n_kept, d_keys2, d_order, d_bmin, d_bmax, d_nodes, d_leaf_last);
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
