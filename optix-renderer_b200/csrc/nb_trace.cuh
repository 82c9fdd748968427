/* nb_trace.cuh — ray/BVH traversal, bit-exact restatement of the reference's query.
 *
 * Follows BVH::rayIntersect (src/utils/bvh.cpp:402-460), TBoundingBox::rayIntersect
 * (include/nori/bbox.h:336-363), Mesh::rayIntersect (src/shapes/mesh.cpp:101-139) and
 * Sphere::rayIntersect (src/shapes/sphere.cpp:67-94) operation by operation:
 *   - fp32, no FMA contraction: every product/sum that decides a hit is written with the
 *     round-to-nearest intrinsics (__fmul_rn/__fadd_rn/__fsub_rn), which nvcc never fuses, so the
 *     header gives the same bits in any translation unit, whatever -fmad / -prec-div / -prec-sqrt
 *     say (x86-64 gcc emits no FMA for the reference); divisions and the square root likewise,
 *   - Eigen 3.3.8 association: dot3 = a0*b0 + (a1*b1 + a2*b2) (Eigen/src/Core/Redux.h:91-105),
 *     cross as Eigen/src/Geometry/OrthoMethods.h:43-47,
 *   - std::max/std::min/std::swap written as the comparisons they are (NaN behaviour included),
 *   - the sphere roots are evaluated in DOUBLE, because `sqrt(discr)` in sphere.cpp:80-81
 *     resolves to ::sqrt(double) (verified in the oracle's object code: cvtss2sd/sqrtsd/divsd).
 * Visiting order is the reference's (left child first, leaves in m_indices order), expressed
 * with precomputed skip links instead of a stack (see nb_internal.h).
 */
#pragma once
#include "nb_internal.h"

#define NB_EPSILON 1e-4f

struct NbHit { float t, u, v; uint32_t slot; };

__device__ __forceinline__ float nb_dot3(float ax, float ay, float az, float bx, float by, float bz) {
    return __fadd_rn(__fmul_rn(ax, bx), __fadd_rn(__fmul_rn(ay, by), __fmul_rn(az, bz)));
}
/* a*b - c*d, each product rounded (Eigen cross component) */
__device__ __forceinline__ float nb_det2(float a, float b, float c, float d) {
    return __fsub_rn(__fmul_rn(a, b), __fmul_rn(c, d));
}

/* bbox.h:336-363 */
__device__ __forceinline__ bool nb_box_hit(const float4 A, const float4 B, const float3 o, const float3 d,
                                           const float3 r, float mint, float maxt) {
    float nearT = -INFINITY, farT = INFINITY;
    const float mn[3] = { A.z, A.w, B.x }, mx[3] = { B.y, B.z, B.w };
    const float oo[3] = { o.x, o.y, o.z }, dd[3] = { d.x, d.y, d.z }, rr[3] = { r.x, r.y, r.z };
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        if (dd[i] == 0.f) {
            if (oo[i] < mn[i] || oo[i] > mx[i]) return false;
        } else {
            float t1 = __fmul_rn(__fsub_rn(mn[i], oo[i]), rr[i]);
            float t2 = __fmul_rn(__fsub_rn(mx[i], oo[i]), rr[i]);
            if (t1 > t2) { float s = t1; t1 = t2; t2 = s; }
            nearT = (t1 < nearT) ? nearT : t1;      /* std::max(t1, nearT) */
            farT = (farT < t2) ? farT : t2;         /* std::min(t2, farT)  */
            if (!(nearT <= farT)) return false;
        }
    }
    return mint <= farT && nearT <= maxt;
}

/* Experiment (nb_mega.cu NB_RENDER_FAST_HITS, trace option "trace_fast"), not used by the shipped kernels: the same
 * Moeller-Trumbore test with contraction allowed (fma) and a fast reciprocal — 45 instead of ~105 instructions, but it
 * opens cracks along shared edges that the reference's arithmetic does not have. */
__device__ __forceinline__ bool nb_tri_hit_fast(const float4 s0, const float4 s1, const float4 s2, const float3 o,
                                           const float3 d, float mint, float maxt, float &u, float &v, float &t) {
    const float p0x = s0.x, p0y = s0.y, p0z = s0.z;
    const float e1x = s0.w, e1y = s1.x, e1z = s1.y;
    const float e2x = s1.z, e2y = s1.w, e2z = s2.x;
    const float pvx = d.y * e2z - d.z * e2y, pvy = d.z * e2x - d.x * e2z, pvz = d.x * e2y - d.y * e2x;
    const float det = e1x * pvx + e1y * pvy + e1z * pvz;
    if (det > -1e-8f && det < 1e-8f) return false;
    const float inv = __fdividef(1.0f, det);
    const float tvx = o.x - p0x, tvy = o.y - p0y, tvz = o.z - p0z;
    u = (tvx * pvx + tvy * pvy + tvz * pvz) * inv;
    if (u < 0.0f || u > 1.0f) return false;
    const float qvx = tvy * e1z - tvz * e1y, qvy = tvz * e1x - tvx * e1z, qvz = tvx * e1y - tvy * e1x;
    v = (d.x * qvx + d.y * qvy + d.z * qvz) * inv;
    if (v < 0.0f || u + v > 1.0f) return false;
    t = (e2x * qvx + e2y * qvy + e2z * qvz) * inv;
    return t >= mint && t <= maxt;
}
/* mesh.cpp:101-139 with e1/e2 precomputed: the reference's arithmetic, bit for bit */
__device__ __forceinline__ bool nb_tri_hit_exact(const float4 s0, const float4 s1, const float4 s2, const float3 o,
                                           const float3 d, float mint, float maxt, float &u, float &v, float &t) {
    const float p0x = s0.x, p0y = s0.y, p0z = s0.z;
    const float e1x = s0.w, e1y = s1.x, e1z = s1.y;
    const float e2x = s1.z, e2y = s1.w, e2z = s2.x;
    const float pvx = nb_det2(d.y, e2z, d.z, e2y), pvy = nb_det2(d.z, e2x, d.x, e2z), pvz = nb_det2(d.x, e2y, d.y, e2x);
    const float det = nb_dot3(e1x, e1y, e1z, pvx, pvy, pvz);
    if (det > -1e-8f && det < 1e-8f) return false;
    const float inv = __fdiv_rn(1.0f, det);
    const float tvx = __fsub_rn(o.x, p0x), tvy = __fsub_rn(o.y, p0y), tvz = __fsub_rn(o.z, p0z);
    u = __fmul_rn(nb_dot3(tvx, tvy, tvz, pvx, pvy, pvz), inv);
    if (u < 0.0f || u > 1.0f) return false;
    const float qvx = nb_det2(tvy, e1z, tvz, e1y), qvy = nb_det2(tvz, e1x, tvx, e1z), qvz = nb_det2(tvx, e1y, tvy, e1x);
    v = __fmul_rn(nb_dot3(d.x, d.y, d.z, qvx, qvy, qvz), inv);
    if (v < 0.0f || __fadd_rn(u, v) > 1.0f) return false;
    t = __fmul_rn(nb_dot3(e2x, e2y, e2z, qvx, qvy, qvz), inv);
    return t >= mint && t <= maxt;
}

__device__ __forceinline__ bool nb_tri_hit(const float4 s0, const float4 s1, const float4 s2, const float3 o,
                                           const float3 d, float mint, float maxt, float &u, float &v, float &t, bool fast = false) {
    return fast ? nb_tri_hit_fast(s0, s1, s2, o, d, mint, maxt, u, v, t) : nb_tri_hit_exact(s0, s1, s2, o, d, mint, maxt, u, v, t);
}
/* sphere.cpp:67-94 */
__device__ __forceinline__ bool nb_sphere_hit(const float4 s0, const float3 o, const float3 d, float mint,
                                              float maxt, float &t) {
    const float Lx = __fsub_rn(o.x, s0.x), Ly = __fsub_rn(o.y, s0.y), Lz = __fsub_rn(o.z, s0.z);
    const float a = nb_dot3(d.x, d.y, d.z, d.x, d.y, d.z);
    const float b = __fmul_rn(2.f, nb_dot3(d.x, d.y, d.z, Lx, Ly, Lz));
    const float c = __fsub_rn(nb_dot3(Lx, Ly, Lz, Lx, Ly, Lz), __fmul_rn(s0.w, s0.w));
    const float discr = __fsub_rn(__fmul_rn(b, b), __fmul_rn(__fmul_rn(4.f, a), c));
    if (discr < 0.f) return false;
    const double sd = __dsqrt_rn((double) discr), nb = (double) (-b), da = (double) a;
    const float tmin = (float) __ddiv_rn(__ddiv_rn(__dsub_rn(nb, sd), 2.0), da);
    const float tmax = (float) __ddiv_rn(__ddiv_rn(__dadd_rn(nb, sd), 2.0), da);
    if (mint <= tmin && maxt >= tmin) { t = tmin; return true; }
    if (mint <= tmax && maxt >= tmax) { t = tmax; return true; }
    return false;
}

/* bvh.cpp:402-460. ANY = shadow-ray query (returns at the first accepted primitive). */
template <bool ANY>
__device__ __forceinline__ bool nb_traverse(const float4 *__restrict__ nodes, const float4 *__restrict__ slots,
                                            uint32_t n_nodes, float3 o, float3 d, float mint, float maxt,
                                            NbHit &hit) {
    hit.t = INFINITY; hit.u = 0.f; hit.v = 0.f; hit.slot = 0xffffffffu;
    if (mint == NB_EPSILON) {                       /* adaptive ray epsilon, bvh.cpp:409-410 */
        float m = fabsf(o.x);
        m = fmaxf(m, fabsf(o.y));
        m = fmaxf(m, fabsf(o.z));
        const float s = __fmul_rn(mint, m);
        mint = (mint < s) ? s : mint;
    }
    if (n_nodes == 0 || maxt < mint) return false;
    const float3 r = make_float3(__fdiv_rn(1.0f, d.x), __fdiv_rn(1.0f, d.y), __fdiv_rn(1.0f, d.z));   /* Ray3f::update, ray.h */
    bool found = false;
    uint32_t idx = 0;
    /* "while-while": every lane first walks inner nodes until it stands on a leaf (cheap, uniform
     * box tests), then the lanes that found one intersect their leaves together. With the leaf
     * loop nested inside the node step instead, a warp executed the ~100-instruction primitive
     * test for 2 of 32 lanes at a time (ncu, table scene). Per ray the sequence of visits is
     * unchanged. */
    while (true) {
        uint32_t leaf = NB_NODE_END, after = NB_NODE_END;
        while (idx != NB_NODE_END) {
            const float4 A = __ldg(nodes + 2 * idx), B = __ldg(nodes + 2 * idx + 1);
            const uint32_t skip = __float_as_uint(A.x), start = __float_as_uint(A.y);
            if (!nb_box_hit(A, B, o, d, r, mint, maxt)) { idx = skip; continue; }
            if (start & NB_NODE_INNER_BIT) { idx = idx + 1; continue; }
            idx = skip;
            if (start == NB_NODE_EMPTY_LEAF) continue;
            leaf = start; after = skip;
            break;
        }
        if (leaf == NB_NODE_END) break;
        uint32_t s = leaf;
        while (true) {
            const float4 s0 = __ldg(slots + 3 * s), s1 = __ldg(slots + 3 * s + 1), s2 = __ldg(slots + 3 * s + 2);
            const uint32_t tag = __float_as_uint(s2.y);
            float u = 0.f, v = 0.f, t;
            const bool ok = (tag & NB_SLOT_SPHERE) ? nb_sphere_hit(s0, o, d, mint, maxt, t)
                                                   : nb_tri_hit(s0, s1, s2, o, d, mint, maxt, u, v, t);
            if (ok) {
                if (ANY) return true;
                found = true;
                maxt = t;
                hit.t = t; hit.u = u; hit.v = v; hit.slot = s;
            }
            if (tag & NB_SLOT_LAST) break;
            ++s;
        }
        idx = after;
    }
    return found;
}


/* Which child of an inner node the ordered walk enters first: the one whose box centre comes first
 * along the ray on the node's axis (nb_inner_word). */
__device__ __forceinline__ bool nb_right_first(uint32_t word, const float3 d) {
    const uint32_t axis = (word >> 29) & 3u;
    const float da = axis == 0u ? d.x : (axis == 1u ? d.y : d.z);
    return (da < 0.f) != ((word >> 28) & 1u);
}

/* The same query with the children of every inner node visited near-to-far (an explicit stack of far
 * children, as deep as the reference's). The set of primitives a ray can hit does not depend on the
 * order; what the reference's fixed left-first order decides is which of two primitives hit at EXACTLY
 * the same t is reported — it keeps the one visited later, i.e. the higher leaf-order slot, because
 * Mesh::rayIntersect accepts t <= maxt (mesh.cpp:134) — so a tie is resolved here by slot number.
 * Boxes are still tested against the current maxt (bbox.h:362), which only prunes primitives that
 * could not be accepted. Held to the same oracle ray sets as nb_traverse (tests/test_trace_gpu.py). */
template <bool ANY>
__device__ __forceinline__ bool nb_traverse_ordered(const float4 *__restrict__ nodes, const float4 *__restrict__ slots,
                                                    uint32_t n_nodes, float3 o, float3 d, float mint, float maxt,
                                                    NbHit &hit) {
    hit.t = INFINITY; hit.u = 0.f; hit.v = 0.f; hit.slot = 0xffffffffu;
    if (mint == NB_EPSILON) {
        float m = fabsf(o.x);
        m = fmaxf(m, fabsf(o.y));
        m = fmaxf(m, fabsf(o.z));
        const float s = __fmul_rn(mint, m);
        mint = (mint < s) ? s : mint;
    }
    if (n_nodes == 0 || maxt < mint) return false;
    const float3 r = make_float3(__fdiv_rn(1.0f, d.x), __fdiv_rn(1.0f, d.y), __fdiv_rn(1.0f, d.z));
    uint32_t stk[NB_STACK_DEPTH];
    uint32_t sp = 0, idx = 0;
    bool found = false;
    while (true) {
        uint32_t leaf = NB_NODE_END;
        while (idx != NB_NODE_END) {
            const float4 A = __ldg(nodes + 2 * idx), B = __ldg(nodes + 2 * idx + 1);
            const uint32_t start = __float_as_uint(A.y);
            if (nb_box_hit(A, B, o, d, r, mint, maxt)) {
                if (start & NB_NODE_INNER_BIT) {
                    const uint32_t right = start & NB_NODE_RIGHT_MASK;
                    const bool rf = nb_right_first(start, d);
                    stk[sp++] = rf ? idx + 1 : right;
                    idx = rf ? right : idx + 1;
                    continue;
                }
                if (start != NB_NODE_EMPTY_LEAF) leaf = start;
            }
            idx = sp ? stk[--sp] : NB_NODE_END;
            if (leaf != NB_NODE_END) break;
        }
        if (leaf == NB_NODE_END) break;
        uint32_t s = leaf;
        while (true) {
            const float4 s0 = __ldg(slots + 3 * s), s1 = __ldg(slots + 3 * s + 1), s2 = __ldg(slots + 3 * s + 2);
            const uint32_t tag = __float_as_uint(s2.y);
            float u = 0.f, v = 0.f, t;
            const bool ok = (tag & NB_SLOT_SPHERE) ? nb_sphere_hit(s0, o, d, mint, maxt, t)
                                                   : nb_tri_hit(s0, s1, s2, o, d, mint, maxt, u, v, t);
            if (ok) {
                if (ANY) return true;
                if (!(found && t == hit.t && s < hit.slot)) {
                    found = true;
                    maxt = t;
                    hit.t = t; hit.u = u; hit.v = v; hit.slot = s;
                }
            }
            if (tag & NB_SLOT_LAST) break;
            ++s;
        }
    }
    return found;
}

/* ---- production traversal: 4-wide tree, near-first (layout: nb_internal.h, NB_WIDE_*) ---------------------------
 * Only the ORDER of visits and the box arithmetic differ from the reference; every primitive test is the bit-exact
 * nb_tri_hit / nb_sphere_hit above and ties in t are resolved by leaf-order slot as in nb_traverse_ordered, so a ray
 * reports the reference's (primitive, t, u, v) unless a box test decides differently within its rounding (measured:
 * tests/test_wide_gpu.py). The slab test is the usual fma form on precomputed o/d with a 4-ulp safety margin instead
 * of the reference's exact comparisons; a zero direction component becomes a huge finite reciprocal. */
struct NbWideRay { float3 r, o; };
__device__ __forceinline__ NbWideRay nb_wide_ray(float3 o, float3 d) {
    NbWideRay w;
    w.r = make_float3(1.0f / d.x, 1.0f / d.y, 1.0f / d.z);
    if (!(fabsf(w.r.x) < 1e30f)) w.r.x = copysignf(1e30f, d.x);
    if (!(fabsf(w.r.y) < 1e30f)) w.r.y = copysignf(1e30f, d.y);
    if (!(fabsf(w.r.z) < 1e30f)) w.r.z = copysignf(1e30f, d.z);
    w.o = o;
    return w;
}
/* (plane - o) * (1 / d), the subtraction first: for a ray that starts on or next to a flat box (a wall, the quad light)
 * the difference is exact, where fma(plane, r, -(o * r)) cancels two large products and is off by a few ulps of |o * r|,
 * i.e. by the hit distance itself for a grazing ray (measured: 6 rays in 10^6 on the Cornell box lost their hit on the
 * quad light with that form). Making the fma form safe needs that error as per-axis slack, which costs the same six
 * instructions per box as the subtractions; one global slack is cheaper but lets every small box of a fine mesh through
 * (B200, C3: 690 -> 522 M samples/s). */
__device__ __forceinline__ bool nb_box_fast(const float4 A, const float4 B, const NbWideRay &w, float mint, float maxt, float &tn) {
    float t1 = (A.z - w.o.x) * w.r.x, t2 = (B.y - w.o.x) * w.r.x;
    float tnear = fminf(t1, t2), tfar = fmaxf(t1, t2);
    t1 = (A.w - w.o.y) * w.r.y; t2 = (B.z - w.o.y) * w.r.y;
    tnear = fmaxf(tnear, fminf(t1, t2)); tfar = fminf(tfar, fmaxf(t1, t2));
    t1 = (B.x - w.o.z) * w.r.z; t2 = (B.w - w.o.z) * w.r.z;
    tnear = fmaxf(tnear, fminf(t1, t2)); tfar = fminf(tfar, fmaxf(t1, t2));
    tfar *= 1.0000005f;                              /* conservative: a primitive on a box face must not be lost to rounding */
    tn = tnear * 0.9999995f;
    return tn <= tfar && mint <= tfar && tn <= maxt;
}
/* the four children of one wide node: tests them, sorts the hit ones by entry distance (nearest first) */
__device__ __forceinline__ int nb_wide_children(const float4 *__restrict__ node, const NbWideRay &w, float mint, float maxt,
                                                float (&t)[4], uint32_t (&ref)[4]) {
    int n = 0;
#pragma unroll
    for (int c = 0; c < 4; ++c) {
        const float4 A = __ldg(node + 2 * c), B = __ldg(node + 2 * c + 1);
        const uint32_t rf = __float_as_uint(A.x);
        float tn;
        const bool ok = rf != NB_WIDE_EMPTY && nb_box_fast(A, B, w, mint, maxt, tn);
        t[c] = ok ? tn : INFINITY; ref[c] = rf;
        n += ok ? 1 : 0;
    }
#define NB_CSWAP(a, b) { const bool sw = t[b] < t[a]; const float ta = sw ? t[b] : t[a], tb = sw ? t[a] : t[b]; \
                         const uint32_t ra = sw ? ref[b] : ref[a], rb = sw ? ref[a] : ref[b]; t[a] = ta; t[b] = tb; ref[a] = ra; ref[b] = rb; }
    NB_CSWAP(0, 1) NB_CSWAP(2, 3) NB_CSWAP(0, 2) NB_CSWAP(1, 3) NB_CSWAP(1, 2)
#undef NB_CSWAP
    return n;
}

template <bool ANY>
__device__ __forceinline__ bool nb_traverse_wide(const float4 *__restrict__ wide, const float4 *__restrict__ slots, uint32_t n_wide,
                                                 float3 o, float3 d, float mint, float maxt, NbHit &hit, bool fast = false) {
    hit.t = INFINITY; hit.u = 0.f; hit.v = 0.f; hit.slot = 0xffffffffu;
    if (mint == NB_EPSILON) {                       /* adaptive ray epsilon, bvh.cpp:409-410 */
        float m = fabsf(o.x);
        m = fmaxf(m, fabsf(o.y));
        m = fmaxf(m, fabsf(o.z));
        const float s = __fmul_rn(mint, m);
        mint = (mint < s) ? s : mint;
    }
    if (n_wide == 0 || maxt < mint) return false;
    const NbWideRay w = nb_wide_ray(o, d);
    uint2 stk[NB_WIDE_STACK];
    uint32_t sp = 0, cur = 0;
    bool found = false;
    while (true) {
        if (!(cur & NB_WIDE_LEAF)) {
            float t[4]; uint32_t ref[4];
            const int n = nb_wide_children(wide + 8 * (size_t) cur, w, mint, maxt, t, ref);
            if (n > 3) stk[sp++] = make_uint2(ref[3], __float_as_uint(t[3]));
            if (n > 2) stk[sp++] = make_uint2(ref[2], __float_as_uint(t[2]));
            if (n > 1) stk[sp++] = make_uint2(ref[1], __float_as_uint(t[1]));
            if (n > 0) { cur = ref[0]; continue; }
        } else {
            uint32_t s = cur & ~NB_WIDE_LEAF;
            while (true) {
                const float4 s0 = __ldg(slots + 3 * s), s1 = __ldg(slots + 3 * s + 1), s2 = __ldg(slots + 3 * s + 2);
                const uint32_t tag = __float_as_uint(s2.y);
                float u = 0.f, v = 0.f, t;
                const bool ok = (tag & NB_SLOT_SPHERE) ? nb_sphere_hit(s0, o, d, mint, maxt, t)
                                                       : nb_tri_hit(s0, s1, s2, o, d, mint, maxt, u, v, t, fast);
                if (ok) {
                    if (ANY) return true;
                    if (!(found && t == hit.t && s < hit.slot)) {      /* exact-t tie: the later slot wins, as in the reference's order */
                        found = true;
                        maxt = t;
                        hit.t = t; hit.u = u; hit.v = v; hit.slot = s;
                    }
                }
                if (tag & NB_SLOT_LAST) break;
                ++s;
            }
        }
        bool got = false;
        while (sp) {
            const uint2 e = stk[--sp];
            if (__uint_as_float(e.y) <= maxt) { cur = e.x; got = true; break; }
        }
        if (!got) break;
    }
    return found;
}

/* Flat scan for scenes of a few primitives (the Cornell box has 14): test every slot in leaf
 * order, no tree. The reference's left-first walk also visits leaves in exactly this order and
 * only skips primitives whose node box the ray misses within [mint, current maxt] — primitives
 * that could not have been accepted anyway — so the sequence of accepted hits, and with it the
 * reported (t, u, v, primitive), is the same; every lane of a warp executes the same slot at the
 * same time, which the per-lane tree walk cannot offer. nb_trace_rays checks it against the same
 * 1M-ray oracle set as the tree walk (tests/test_trace_gpu.py). */
template <bool ANY>
__device__ __forceinline__ bool nb_scan(const float4 *__restrict__ slots, uint32_t n_slots, float3 o, float3 d,
                                        float mint, float maxt, NbHit &hit, bool fast = false) {
    hit.t = INFINITY; hit.u = 0.f; hit.v = 0.f; hit.slot = 0xffffffffu;
    if (mint == NB_EPSILON) {
        float m = fabsf(o.x);
        m = fmaxf(m, fabsf(o.y));
        m = fmaxf(m, fabsf(o.z));
        const float s = __fmul_rn(mint, m);
        mint = (mint < s) ? s : mint;
    }
    if (maxt < mint) return false;
    bool found = false;
    for (uint32_t s = 0; s < n_slots; ++s) {
        #ifdef NB_SMEM_SLOTS
        const float4 s0 = slots[3 * s], s1 = slots[3 * s + 1], s2 = slots[3 * s + 2];       /* may point into shared memory */
#else
        const float4 s0 = __ldg(slots + 3 * s), s1 = __ldg(slots + 3 * s + 1), s2 = __ldg(slots + 3 * s + 2);
#endif
        float u = 0.f, v = 0.f, t;
        const bool ok = (__float_as_uint(s2.y) & NB_SLOT_SPHERE) ? nb_sphere_hit(s0, o, d, mint, maxt, t)
                                                                : nb_tri_hit(s0, s1, s2, o, d, mint, maxt, u, v, t, fast);
        if (ok) {
            if (ANY) return true;
            found = true;
            maxt = t;
            hit.t = t; hit.u = u; hit.v = v; hit.slot = s;
        }
    }
    return found;
}

/* Experiment (NB_FUSED_SCAN, off: measured slower): two rays per lane in ONE pass over the slots — the closest hit of
 * a main ray (if `main_on`) and the any-hit answer of a shadow ray (if `shadow_on`), i.e. the flat render kernels' NEE ray of
 * bounce k and the path's ray of bounce k + 1. Each ray sees the tests of nb_scan<false> / nb_scan<true> in the same order with
 * the same arithmetic; what is shared is the slot fetch and the loop. Returns the main ray's "found"; `occluded` answers the
 * shadow ray. */
__device__ __forceinline__ bool nb_scan2(const float4 *__restrict__ slots, uint32_t n_slots, bool main_on, float3 o, float3 d, float mint, float maxt,
                                         NbHit &hit, bool shadow_on, float3 so, float3 sd, float smint, float smaxt, bool &occluded) {
    hit.t = INFINITY; hit.u = 0.f; hit.v = 0.f; hit.slot = 0xffffffffu;
    occluded = false;
    if (mint == NB_EPSILON) {
        const float m = fmaxf(fmaxf(fabsf(o.x), fabsf(o.y)), fabsf(o.z));
        const float s = __fmul_rn(mint, m);
        mint = (mint < s) ? s : mint;
    }
    if (smint == NB_EPSILON) {
        const float m = fmaxf(fmaxf(fabsf(so.x), fabsf(so.y)), fabsf(so.z));
        const float s = __fmul_rn(smint, m);
        smint = (smint < s) ? s : smint;
    }
    main_on = main_on && !(maxt < mint);
    shadow_on = shadow_on && !(smaxt < smint);
    bool found = false;
    for (uint32_t s = 0; s < n_slots; ++s) {
        const float4 s0 = __ldg(slots + 3 * s), s1 = __ldg(slots + 3 * s + 1), s2 = __ldg(slots + 3 * s + 2);
        const bool sphere = (__float_as_uint(s2.y) & NB_SLOT_SPHERE) != 0u;
        if (main_on) {
            float u = 0.f, v = 0.f, t;
            const bool ok = sphere ? nb_sphere_hit(s0, o, d, mint, maxt, t) : nb_tri_hit(s0, s1, s2, o, d, mint, maxt, u, v, t);
            if (ok) { found = true; maxt = t; hit.t = t; hit.u = u; hit.v = v; hit.slot = s; }
        }
        if (shadow_on) {
            float u = 0.f, v = 0.f, t;
            const bool ok = sphere ? nb_sphere_hit(s0, so, sd, smint, smaxt, t) : nb_tri_hit(s0, s1, s2, so, sd, smint, smaxt, u, v, t);
            if (ok) { occluded = true; shadow_on = false; }
        }
    }
    return found;
}

/* nb_scan over slots held in kernel-parameter space (NbDeviceScene::flat_slots): same tests, same order, same bits;
 * the warp-uniform slot index makes every fetch a constant-bank broadcast. */
template <bool ANY>
__device__ __forceinline__ bool nb_scan_params(const float4 *slots, uint32_t n_slots, float3 o, float3 d,
                                               float mint, float maxt, NbHit &hit) {
    hit.t = INFINITY; hit.u = 0.f; hit.v = 0.f; hit.slot = 0xffffffffu;
    if (mint == NB_EPSILON) {
        float m = fabsf(o.x);
        m = fmaxf(m, fabsf(o.y));
        m = fmaxf(m, fabsf(o.z));
        const float s = __fmul_rn(mint, m);
        mint = (mint < s) ? s : mint;
    }
    if (maxt < mint) return false;
    bool found = false;
    for (uint32_t s = 0; s < n_slots; ++s) {
        const float4 s0 = slots[3 * s], s1 = slots[3 * s + 1], s2 = slots[3 * s + 2];
        float u = 0.f, v = 0.f, t;
        const bool ok = (__float_as_uint(s2.y) & NB_SLOT_SPHERE) ? nb_sphere_hit(s0, o, d, mint, maxt, t)
                                                                : nb_tri_hit(s0, s1, s2, o, d, mint, maxt, u, v, t);
        if (ok) {
            if (ANY) return true;
            found = true;
            maxt = t;
            hit.t = t; hit.u = u; hit.v = v; hit.slot = s;
        }
    }
    return found;
}

/* The query the render kernels use: flat scan when the scene qualifies, else the tree in the order `mode` names. */
enum { NB_TRAV_REFERENCE = 0, NB_TRAV_ORDERED = 1, NB_TRAV_WIDE = 2 };
template <bool ANY, bool FLAT>
__device__ __forceinline__ bool nb_query(const float4 *__restrict__ nodes, const float4 *__restrict__ slots, uint32_t n_nodes,
                                         uint32_t n_slots, float3 o, float3 d, float mint, float maxt, NbHit &hit, int mode = NB_TRAV_REFERENCE) {
    if (FLAT) return nb_scan<ANY>(slots, n_slots, o, d, mint, maxt, hit);
    if (mode == NB_TRAV_ORDERED) return nb_traverse_ordered<ANY>(nodes, slots, n_nodes, o, d, mint, maxt, hit);
    return nb_traverse<ANY>(nodes, slots, n_nodes, o, d, mint, maxt, hit);
}
