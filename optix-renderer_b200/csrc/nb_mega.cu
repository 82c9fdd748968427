/* nb_mega.cu — K_tile: persistent per-warp path tracing with the whole path state in registers.
 *
 * Second execution strategy next to the wavefront pipeline of nb_render.cu, for scenes whose BVH
 * and shading data stay cache-resident (the Cornell box of the headline metric is < 2 KB): no
 * path pool in HBM, no queues, no per-iteration launches. Every warp repeatedly takes a work item
 * (one 16x8-pixel tile x a chunk of sample passes) from a global counter; a lane whose path ends
 * splats it (ImageBlock::put) and starts the item's next camera sample in the same loop
 * iteration, so all 32 lanes stay busy and stay in the same phase of the loop
 *      regenerate -> closest-hit -> shade/NEE/RR -> any-hit (shadow) -> splat.
 * Sample numbering, pcg32 streams and draw order per path are identical to the wavefront
 * pipeline's, so both strategies produce the same per-sample radiance.
 *
 * Integrators restated here: PathMISIntegrator::Li (src/integrators/path_mis.cpp:16-150),
 * PathMATSIntegrator::Li (path_mats.cpp:16-78), DirectIntegrator::Li (direct.cpp:16-67),
 * DirectEMSIntegrator::Li (direct_ems.cpp:15-74), DirectMATSIntegrator::Li (direct_mats.cpp:16-83),
 * DirectMISIntegrator::Li (direct_mis.cpp:15-139). */
#include "nb_internal.h"
#include "nb_wavefront.h"
#include "nb_shading.cuh"
#include "nb_film.cuh"
#include "nb_trace.cuh"
#include "nb_volume.cuh"

#ifndef NB_TILE_THREADS
#define NB_TILE_THREADS 128
#endif
#ifndef NB_TILE_MIN_BLOCKS
#define NB_TILE_MIN_BLOCKS 7      /* 72 registers. B200, 6 / 7 / 8 CTAs per SM, M samples/s: cbox path_mis 1961 / 1988 / 2032, path_mats 3058 / 3199 / 3137,
                                   * cbox_val_mis 1018 / 1004 / 978, cbox_homog_aniso 2920 / 3052 / 2915, veach direct_mis 1449 / 1489 / 1445;
                                   * cbox path_mis at 8 / 9 / 10 / 12: 2027 / 2026 / 1904 / 1875 */
#endif
/* CTAs per SM asked of the EXT variants (launch bounds): the kernels that carry the texture / disney / spot /
 * environment / heterogeneous code are 1.5-2x larger and wait on instruction fetch (see nb_k_tile), so they trade
 * registers for resident warps; the kernels without that code are best at 7 (72 registers) */
#ifndef NB_TILE_MIN_BLOCKS_EXT
#define NB_TILE_MIN_BLOCKS_EXT 8
#endif
#ifndef NB_WALK_MIN_BLOCKS_EXT
#define NB_WALK_MIN_BLOCKS_EXT 8
#endif

struct NbTilePlan {
    NbPlan plan;
    uint32_t height, passes;          /* image rows; sample passes of this subframe */
    uint32_t tile_w, tile_h;          /* pixels per tile */
    uint32_t tiles_x, n_tiles;
    uint32_t chunk;                   /* passes per work item */
    uint32_t n_items;                 /* n_tiles * ceil(passes / chunk) */
    /* adaptive rounds (nb_render_adaptive_round): the camera samples of this launch are an explicit LIST of pixel indices
     * written by the device (nb_k_adaptive_decide), its length and first sample index in list_ctl[0], [1] — the host sizes
     * nothing. Items are NB_LIST_ITEM consecutive entries. Null for every other launch; read by the EXT kernel variants only
     * (nb_launch_tile runs those for a list). */
    const uint32_t *list, *list_ctl;
};
#define NB_LIST_ITEM 128u

struct NbPathRegs {                   /* path_mis.cpp:18-24 locals + the ray being traced */
    f3 Li, thr, o, d;
    float mint, maxt, w_mats, w_ems, pdfmat, px, py;
    uint32_t flags;
    NbRng rng;
};

struct NbShadowRay { f3 o, d, c; float mint, maxt, cand; };

template <bool EXT>
__device__ __forceinline__ void nb_hit_info(const NbDeviceScene &sc, const NbHit &hit, f3 o, f3 d, nb_shape &sh, NbIts &its) {
#ifdef NB_SMEM_SLOTS
    const float4 sl = sc.slots[3 * hit.slot + 2];
#else
    const float4 sl = __ldg(sc.slots + 3 * hit.slot + 2);
#endif
    const uint32_t sid = __float_as_uint(sl.y) & NB_SLOT_SHAPE_MASK, prim = __float_as_uint(sl.z);
    sh = sc.shapes[sid];
    its.shape = sid;
    if (sh.type == NB_SHAPE_SPHERE) nb_sphere_hit_info<EXT>(sc, sh, o, d, hit.t, its);
    else nb_mesh_hit_info<EXT>(sc, sh, prim, hit.u, hit.v, its);
}

/* AreaEmitter::eval (arealight.cpp:58-73) for EmitterQueryRecord(ref, p, n) */
template <bool EXT>
__device__ __forceinline__ f3 nb_area_eval(const nb_emitter &em, f3 ref, f3 p, f3 n) {
    if (EXT && em.type == NB_EMITTER_VOLUME) return mk3(em.radiance[0], em.radiance[1], em.radiance[2]);   /* VolumeEmitter::eval */
    const f3 wi = normalize(p - ref);
    if (dot(n, -wi) >= 0.f) return mk3(em.radiance[0], em.radiance[1], em.radiance[2]);
    return mk3(0.f);
}

/* Experiment kept behind a macro, NOT adopted: nb_tri_hit_fast (fma-contracted Moeller-Trumbore, fast reciprocal;
 * nb_trace.cuh) as the render kernels' triangle test. B200: cbox path_mis +11-14 %, val_mis +10 %, table +3-16 %, C3 +4 %.
 * But through the trace entry ("trace_fast") it reports another primitive on 3 of 24 656 table rays (1.2e-4, the bar is
 * 1e-6), and the reference's own known-answer scenes scenes/pa3/tests/test-mesh.xml — a camera ray aimed exactly at the
 * edge two triangles share — lose HALF their energy (0.0451 against 0.0898): with the reference's arithmetic u + v == 1
 * exactly on that edge and one triangle takes the ray, contracted it is > 1 in both and the ray falls through the crack. */
#ifndef NB_RENDER_FAST_HITS
#define NB_RENDER_FAST_HITS false
#endif
/* the lock-step kernels on a tree scene (direct_*, the volume integrators, normals / av / preview): one out-of-line copy
 * of the production traversal per translation unit, so that the kernels which call it from several sites do not grow */
template <bool ANY>
__device__ __noinline__ bool nb_wide_query(const float4 *wide, const float4 *slots, uint32_t n_wide, float ox, float oy, float oz,
                                           float dx, float dy, float dz, float mint, float maxt, float4 &out, bool fast) {
    NbHit h;
    const bool r = nb_traverse_wide<ANY>(wide, slots, n_wide, make_float3(ox, oy, oz), make_float3(dx, dy, dz), mint, maxt, h, fast);
    out = make_float4(h.t, h.u, h.v, __uint_as_float(h.slot));
    return r;
}
template <bool FLAT>
__device__ __forceinline__ bool nb_occluded(const NbDeviceScene &sc, f3 o, f3 d, float mint, float maxt) {
    NbHit dummy;
#ifdef NB_PARAM_SLOTS
    if (FLAT) return nb_scan_params<true>(sc.flat_slots, sc.n_slots, o, d, mint, maxt, dummy);
#endif
    if (!FLAT && sc.use_wide) {
        float4 h;
        return nb_wide_query<true>(sc.wide, sc.slots, sc.n_wide, o.x, o.y, o.z, d.x, d.y, d.z, mint, maxt, h, NB_RENDER_FAST_HITS);
    }
    if (FLAT) return nb_scan<true>(sc.slots, sc.n_slots, o, d, mint, maxt, dummy, NB_RENDER_FAST_HITS);
    return nb_query<true, FLAT>(sc.nodes, sc.slots, sc.n_nodes, sc.n_slots, o, d, mint, maxt, dummy);
}
template <bool FLAT>
__device__ __forceinline__ bool nb_closest(const NbDeviceScene &sc, f3 o, f3 d, float mint, float maxt, NbHit &hit) {
#ifdef NB_PARAM_SLOTS
    if (FLAT) return nb_scan_params<false>(sc.flat_slots, sc.n_slots, o, d, mint, maxt, hit);
#endif
    if (!FLAT && sc.use_wide) {
        float4 h;
        const bool r = nb_wide_query<false>(sc.wide, sc.slots, sc.n_wide, o.x, o.y, o.z, d.x, d.y, d.z, mint, maxt, h, NB_RENDER_FAST_HITS);
        hit.t = h.x; hit.u = h.y; hit.v = h.z; hit.slot = __float_as_uint(h.w);
        return r;
    }
    if (FLAT) return nb_scan<false>(sc.slots, sc.n_slots, o, d, mint, maxt, hit, NB_RENDER_FAST_HITS);
    return nb_query<false, FLAT>(sc.nodes, sc.slots, sc.n_nodes, sc.n_slots, o, d, mint, maxt, hit);
}

/* One bounce of path_mis / path_mats for a path whose closest hit just arrived. Returns false
 * when the path ended at this vertex; `shadow` is set when an NEE ray must be resolved. */
template <bool MIS, bool EXT>
__device__ __forceinline__ bool nb_path_bounce(const NbDeviceScene &sc, uint32_t max_depth, NbPathRegs &ps, const NbHit &hit,
                                               bool &shadow, NbShadowRay &so) {
    nb_shape sh; NbIts its;
    nb_hit_info<EXT>(sc, hit, ps.o, ps.d, sh, its);
    const float n_lights = (float) sc.n_emitters;
    /* emitter hit (path_mis.cpp:51-56); the MIS weight of the BSDF-sampled ray that led here
     * (path_mis.cpp:117-140) is evaluated now that its hit is known */
    if (sh.emitter >= 0) {
        const nb_emitter em = sc.emitters[sh.emitter];
        if (MIS && !(ps.flags & NB_FLAG_PRIMARY) && (ps.flags & NB_FLAG_BSDF_OK)) {
            const f3 wi_e = normalize(its.p - ps.o);
            const float pdfmat_ems = nb_hit_pdf<EXT>(sc, sh, ps.o, its.p, its.sh.n, wi_e) / n_lights;
            if (ps.pdfmat + pdfmat_ems > NB_EPS) ps.w_mats = ps.pdfmat / (ps.pdfmat + pdfmat_ems);
        }
        if (MIS && (ps.flags & NB_FLAG_DISCRETE)) ps.w_mats = 1.f;
        ps.Li += (MIS ? ps.w_mats : 1.f) * ps.thr * nb_area_eval<EXT>(em, ps.o, its.p, its.sh.n);
    } else if (MIS && (ps.flags & NB_FLAG_DISCRETE)) ps.w_mats = 1.f;

    const uint32_t bounce = ps.flags >> NB_BOUNCE_SHIFT;
    /* Russian roulette: path_mis.cpp:58-71 (always, clamped) / path_mats.cpp:49-61 (after 3) */
    float succ = fminf(max3(ps.thr), 0.99f);
    if (MIS) {
        succ = fmaxf(succ, NB_EPS);
        if (ps.rng.next_float() > succ) return false;
        ps.thr = ps.thr / succ;
    } else if (bounce >= 3) {
        if (ps.rng.next_float() > succ) return false;
        ps.thr = ps.thr / succ;
    }
    if ((max_depth && bounce >= max_depth) || sh.bsdf < 0) return false;

    nb_bsdf bsdf = sc.bsdfs[sh.bsdf];
    nb_bsdf_resolve<EXT>(sc, bsdf, its.uv);
    const f3 wi = its.sh.to_local(-ps.d);
    if (MIS) {                                                       /* path_mis.cpp:75-106 */
        const uint32_t ei = nb_cdf_sample(sc.emitter_cdf, sc.n_emitters, ps.rng.next_float());
        float2 s; s.x = ps.rng.next_float(); s.y = ps.rng.next_float();
        NbEmitterSample es;
        const f3 ems_col = nb_emitter_sample<EXT>(sc, ei, its.p, s, es);
        if (!is_zero(ems_col, NB_EPS)) {
            const f3 we = its.sh.to_local(es.wi);
            const f3 C = ps.thr * (ems_col * we.z * nb_bsdf_eval<EXT>(bsdf, wi, we, sc.params) * n_lights);
            if (!is_black(C)) {
                const float pdf_b = nb_bsdf_pdf<EXT>(bsdf, wi, we), pdf_e = es.pdf / n_lights;
                shadow = true;
                so.o = es.ray_o; so.d = es.ray_d; so.mint = es.ray_mint; so.maxt = es.ray_maxt; so.c = C;
                so.cand = (pdf_b + pdf_e > NB_EPS) ? pdf_e / (pdf_b + pdf_e) : -1.f;
            }
        }
    }
    /* BSDF sampling (path_mis.cpp:108-146 / path_mats.cpp:63-74) */
    float2 s; s.x = ps.rng.next_float(); s.y = ps.rng.next_float();
    f3 wo; bool discrete;
    const f3 col = nb_bsdf_sample<EXT>(bsdf, wi, s, wo, discrete, ps.pdfmat, sc.params);
    ps.flags = (bounce + 1) << NB_BOUNCE_SHIFT;
    if (discrete) { ps.flags |= NB_FLAG_DISCRETE; ps.w_ems = 0.f; }
    if (!is_zero(col, NB_EPS)) ps.flags |= NB_FLAG_BSDF_OK;
    ps.thr = ps.thr * col;
    ps.o = its.p; ps.d = its.sh.to_world(wo); ps.mint = NB_EPS; ps.maxt = INFINITY;
    return !is_black(ps.thr);                                       /* a black path can gain nothing more */
}

/* The depth-1 integrators, one whole sample. Returns Li. */
template <int INTEG, bool FLAT, bool EXT>
__device__ __forceinline__ f3 nb_direct_sample(const NbDeviceScene &sc, const NbPlan &plan, NbPathRegs &ps, uint32_t &n_closest, uint32_t &n_any,
                                                f3 &layer_normal) {
    NbHit hit;
    layer_normal = mk3(0.f);
    ++n_closest;
    if (!nb_closest<FLAT>(sc, ps.o, ps.d, ps.mint, ps.maxt, hit)) {
        if (INTEG == NB_INTEGRATOR_AV) return mk3(1.f);                                                     /* av.cpp:21-22 */
        return nb_env_radiance<EXT>(sc, ps.d);                                                              /* direct*.cpp, normals.cpp, PreviewIntegrator.cpp: getEnvMap()->eval */
    }
    nb_shape sh; NbIts its;
    nb_hit_info<EXT>(sc, hit, ps.o, ps.d, sh, its);
    if (INTEG == NB_INTEGRATOR_NORMALS) {                             /* normals.cpp:31-35 */
        const f3 n = its.sh.to_world(mk3(plan.iparam[0], plan.iparam[1], plan.iparam[2]));
        return mk3(fabsf(n.x), fabsf(n.y), fabsf(n.z));
    }
    if (INTEG == NB_INTEGRATOR_AV) {                                  /* av.cpp:24-42; Warp::sampleUniformHemisphere warp.cpp:25-39 */
        f3 v;
        do {
            v.x = 1.f - 2.f * ps.rng.next_float();
            v.y = 1.f - 2.f * ps.rng.next_float();
            v.z = 1.f - 2.f * ps.rng.next_float();
        } while (dot(v, v) > 1.f);
        if (dot(v, its.sh.n) < 0.f) v = -v;
        v = v / length(v);
        NbHit h2;
        ++n_closest;                                                  /* the reference asks for the closest hit here, not a shadow query */
        return nb_closest<FLAT>(sc, its.p, v, NB_EPS, plan.iparam[0], h2) ? mk3(0.f) : mk3(1.f);
    }
    if (INTEG == NB_INTEGRATOR_PREVIEW) {                             /* PreviewIntegrator.cpp:37-76 */
        layer_normal = its.sh.n;
        if (sc.n_emitters == 0) return mk3(fabsf(its.sh.n.x), fabsf(its.sh.n.y), fabsf(its.sh.n.z));
        const f3 wo = its.sh.to_local(normalize(ps.o - its.p));
        const uint32_t ei = nb_cdf_sample(sc.emitter_cdf, sc.n_emitters, ps.rng.next_float());
        float2 s; s.x = ps.rng.next_float(); s.y = ps.rng.next_float();
        NbEmitterSample es;
        const f3 li = nb_emitter_sample<EXT>(sc, ei, its.p, s, es) * (float) sc.n_emitters;
        ++n_any;
        if (sh.bsdf < 0 || nb_occluded<FLAT>(sc, es.ray_o, es.ray_d, es.ray_mint, es.ray_maxt)) return mk3(0.f);
        nb_bsdf bsdf = sc.bsdfs[sh.bsdf];
        nb_bsdf_resolve<EXT>(sc, bsdf, its.uv);
        const float c = fabsf(dot(es.wi, its.sh.n)) / length(es.wi);
        return li * c * nb_bsdf_eval<EXT>(bsdf, its.sh.to_local(es.wi), wo, sc.params);
    }
    f3 result = mk3(0.f);
    if (INTEG != NB_INTEGRATOR_DIRECT && sh.emitter >= 0)
        result += nb_area_eval<EXT>(sc.emitters[sh.emitter], ps.o, its.p, its.sh.n);
    if (sh.bsdf < 0) return result;
    nb_bsdf bsdf = sc.bsdfs[sh.bsdf];
    nb_bsdf_resolve<EXT>(sc, bsdf, its.uv);
    const float n_lights = (float) sc.n_emitters;

    if (INTEG == NB_INTEGRATOR_DIRECT) {                              /* direct.cpp:37-63 */
        const f3 wo = its.sh.to_local(normalize(ps.o - its.p));
        for (uint32_t l = 0; l < sc.n_emitters; ++l) {
            float2 s; s.x = ps.rng.next_float(); s.y = ps.rng.next_float();
            NbEmitterSample es;
            const f3 li = nb_emitter_sample<EXT>(sc, l, its.p, s, es);
            const f3 wi = its.sh.to_local(es.wi);
            ++n_any;
            if (nb_occluded<FLAT>(sc, es.ray_o, es.ray_d, es.ray_mint, es.ray_maxt)) continue;
            const float c = fabsf(dot(es.wi, its.sh.n)) / length(es.wi);
            result += li * c * nb_bsdf_eval<EXT>(bsdf, wi, wo, sc.params);
        }
        return result;
    }
    const f3 wi = its.sh.to_local(-ps.d);
    if (INTEG == NB_INTEGRATOR_DIRECT_EMS) {                          /* direct_ems.cpp:42-70 */
        for (uint32_t l = 0; l < sc.n_emitters; ++l) {
            float2 s; s.x = ps.rng.next_float(); s.y = ps.rng.next_float();
            NbEmitterSample es;
            const f3 li = nb_emitter_sample<EXT>(sc, l, its.p, s, es);
            if (is_zero(li, NB_EPS)) continue;
            ++n_any;
            if (nb_occluded<FLAT>(sc, es.ray_o, es.ray_d, es.ray_mint, es.ray_maxt)) continue;
            const f3 we = its.sh.to_local(es.wi);
            result += li * fabsf(we.z) * nb_bsdf_eval<EXT>(bsdf, wi, we, sc.params);
        }
        return result;
    }
    float w_ems = 0.f, w_mat = 0.f;
    f3 result_ems = mk3(0.f), result_mats = mk3(0.f);
    if (INTEG == NB_INTEGRATOR_DIRECT_MIS) {                          /* direct_mis.cpp:49-90 */
        const uint32_t ei = nb_cdf_sample(sc.emitter_cdf, sc.n_emitters, ps.rng.next_float());
        float2 s; s.x = ps.rng.next_float(); s.y = ps.rng.next_float();
        NbEmitterSample es;
        const f3 li = nb_emitter_sample<EXT>(sc, ei, its.p, s, es);
        if (!is_zero(li, NB_EPS)) {
            ++n_any;
            if (!nb_occluded<FLAT>(sc, es.ray_o, es.ray_d, es.ray_mint, es.ray_maxt)) {
                const f3 we = its.sh.to_local(es.wi);
                const float pdf_ems = es.pdf / n_lights, pdf_mat = nb_bsdf_pdf<EXT>(bsdf, wi, we);
                result_ems = li * we.z * nb_bsdf_eval<EXT>(bsdf, wi, we, sc.params) * n_lights;
                if (pdf_ems + pdf_mat > NB_EPS) w_ems = pdf_ems / (pdf_ems + pdf_mat);
            }
        }
    }
    /* BSDF sampling: direct_mats.cpp:46-80 / direct_mis.cpp:92-135 */
    float2 s; s.x = ps.rng.next_float(); s.y = ps.rng.next_float();
    f3 wo; bool discrete; float pdf_mat;
    const f3 col = nb_bsdf_sample<EXT>(bsdf, wi, s, wo, discrete, pdf_mat, sc.params);
    if (!is_zero(col, NB_EPS)) {
        const f3 d2 = its.sh.to_world(wo);
        NbHit hit2;
        ++n_closest;
        if (nb_closest<FLAT>(sc, its.p, d2, NB_EPS, INFINITY, hit2)) {
            nb_shape sh2; NbIts its2;
            nb_hit_info<EXT>(sc, hit2, its.p, d2, sh2, its2);
            if (sh2.emitter >= 0) {
                result_mats = col * nb_area_eval<EXT>(sc.emitters[sh2.emitter], its.p, its2.p, its2.sh.n);
                if (INTEG == NB_INTEGRATOR_DIRECT_MIS) {
                    const float pdf_ems = nb_hit_pdf<EXT>(sc, sh2, its.p, its2.p, its2.sh.n, normalize(its2.p - its.p)) / n_lights;
                    if (pdf_mat + pdf_ems > NB_EPS) w_mat = pdf_mat / (pdf_mat + pdf_ems);
                }
            }
        } else if (INTEG == NB_INTEGRATOR_DIRECT_MATS)
            result_mats = col * nb_env_radiance<EXT>(sc, d2);                 /* direct_mats.cpp:60-69; in direct_mis.cpp:122-133 the weight stays 0 */
    }
    if (INTEG == NB_INTEGRATOR_DIRECT_MATS) return result + result_mats;
    return result + w_ems * result_ems + w_mat * result_mats;
}

/* PathVolMISIntegrator::sampleEmitter + traceShadowray (path_vol_mis.cpp:26-104). `its`/`sh` are
 * the SURFACE hit of the current ray even when the interaction is inside the medium (the
 * reference passes `its` in both cases); exactly one of bsdf / phase is non-null. */
template <bool FLAT, bool EXT>
__device__ __forceinline__ f3 nb_vol_sample_emitter(const NbDeviceScene &sc, NbPathRegs &ps, int32_t medium, const NbIts &its,
                                                    const nb_shape &sh, const nb_bsdf *bsdf,
                                                    const nb_medium *phase, uint32_t &n_closest) {
    const float n_lights = (float) sc.n_emitters;
    const uint32_t ei = nb_cdf_sample(sc.emitter_cdf, sc.n_emitters, ps.rng.next_float());
    float2 s; s.x = ps.rng.next_float(); s.y = ps.rng.next_float();
    NbEmitterSample es;
    const f3 Le = nb_emitter_sample<EXT>(sc, ei, its.p, s, es) * n_lights;
    const float pdf_em = es.pdf / n_lights;
    if (is_black(Le)) return mk3(0.f);                                   /* result would be 0; the shadow walk draws nothing */
    int32_t sm = dot(ps.d, es.ray_d) > 0.f ? nb_medium_behind(sc, sh, its.geo_n, es.ray_d) : medium;
    /* traceShadowray: closest hits along the ray until a surface with a BSDF blocks it */
    f3 tr = mk3(1.f), so = es.ray_o;
    float smaxt = es.ray_maxt;
    for (int seg = 0; seg < 64; ++seg) {
        NbHit h;
        ++n_closest;
        if (!nb_closest<FLAT>(sc, so, es.ray_d, es.ray_mint, smaxt, h)) break;
        nb_shape sh2; NbIts its2;
        nb_hit_info<EXT>(sc, h, so, es.ray_d, sh2, its2);
        if (sh2.bsdf >= 0) return mk3(0.f);
        tr = tr * nb_transmittance<EXT>(sc, sm, so, its2.p, ps.rng);
        sm = nb_medium_behind(sc, sh2, its2.geo_n, es.ray_d);
        so = its2.p;
        smaxt -= h.t;
    }
    float pdf_em_mat;
    f3 f = mk3(1.f);                                                     /* |wi . shadowRay.d| = 1 for a phase function */
    if (phase) pdf_em_mat = nb_phase_pdf(*phase, its.sh.to_local(es.ray_d));
    else {
        const f3 wi = its.sh.to_local(-ps.d), we = its.sh.to_local(es.wi);
        pdf_em_mat = nb_bsdf_pdf<EXT>(*bsdf, wi, we);
        f = nb_bsdf_eval<EXT>(*bsdf, wi, we, sc.params) * fabsf(dot(es.wi, its.sh.n));
    }
    const float w_em = pdf_em > NB_EPS ? pdf_em / (pdf_em + pdf_em_mat) : 0.f;
    if (w_em > NB_EPS) return tr * w_em * ps.thr * Le * f;
    return mk3(0.f);
}

/* One loop iteration of path_vol_mats.cpp:32-119 / path_vol_mis.cpp:124-229 after the closest
 * hit of the current ray arrived. Returns false when the path ended. */
template <bool MIS, bool FLAT, bool EXT>
__device__ __forceinline__ bool nb_vol_bounce(const NbDeviceScene &sc, uint32_t max_depth, NbPathRegs &ps, int32_t &medium,
                                              const NbHit &hit, uint32_t &n_closest) {
    nb_shape sh; NbIts its;
    nb_hit_info<EXT>(sc, hit, ps.o, ps.d, sh, its);
    const float tm = nb_free_path<EXT>(sc, medium, ps.o, ps.d, ps.rng);
    const bool mi = tm < hit.t;                                          /* isMediumInteraction */
    const f3 p = mi ? ps.o + ps.d * tm : its.p;
    ps.thr = ps.thr * nb_transmittance<EXT>(sc, medium, ps.o, p, ps.rng);
    uint32_t measure = (ps.flags >> NB_MEASURE_SHIFT) & 3u;
    /* captured emission (path_vol_mats.cpp:52-68, path_vol_mis.cpp:146-173): inside a medium that carries a volume
     * light, every medium interaction collects its radiance; the MIS pdf is taken at its.p as the reference does */
    if (EXT && mi && sc.medium_emitter) {
        const int32_t me = sc.medium_emitter[medium];
        if (me >= 0) {
            const nb_emitter em = sc.emitters[me];
            float w = 1.f;
            if (MIS && measure != NB_MEASURE_DISCRETE) {
                const f3 dlt = its.p - ps.o;
                const float pdf_mat_em = em.position[0] * dot(dlt, dlt) / (float) sc.n_emitters;
                w = ps.pdfmat > NB_EPS ? ps.pdfmat / (ps.pdfmat + pdf_mat_em) : 0.f;
            }
            ps.Li += w * ps.thr * mk3(em.radiance[0], em.radiance[1], em.radiance[2]);
        }
    }
    if (!mi && sh.emitter >= 0) {
        const f3 Le = nb_area_eval<EXT>(sc.emitters[sh.emitter], ps.o, its.p, its.sh.n);
        float w = 1.f;
        if (MIS && measure != NB_MEASURE_DISCRETE) {
            const float pdf_mat_em = nb_hit_pdf<EXT>(sc, sh, ps.o, its.p, its.sh.n, normalize(its.p - ps.o)) / (float) sc.n_emitters;
            w = ps.pdfmat > NB_EPS ? ps.pdfmat / (ps.pdfmat + pdf_mat_em) : 0.f;
        }
        ps.Li += w * ps.thr * Le;
    }
    const uint32_t bounce = ps.flags >> NB_BOUNCE_SHIFT;
    if (bounce >= 3 && (mi || sh.bsdf >= 0)) {                           /* Russian roulette */
        const float succ = fminf(max3(ps.thr), 0.99f);
        if (ps.rng.next_float() > succ || succ < NB_EPS) return false;
        ps.thr = ps.thr / succ;
    }
    if (max_depth && bounce >= max_depth) return false;
    f3 wo = ps.d;                                                        /* default: unperturbed */
    /* the two kinds of interaction meet again at ONE emitter-sampling site: lanes at a medium interaction and lanes
     * at a surface walk their shadow rays together, and the kernel holds one copy of that code instead of two */
    nb_medium med; nb_bsdf bsdf;
    f3 col = mk3(1.f);
    bool nee = false;
    if (mi) {
        med = sc.media[medium];
        float2 s; s.x = ps.rng.next_float(); s.y = ps.rng.next_float();
        const f3 pwo = nb_phase_sample(med, s);
        wo = frame_from_normal(ps.d).to_world(pwo);
        if (MIS) {
            ps.pdfmat = nb_phase_pdf(med, pwo);
            measure = NB_MEASURE_SOLID_ANGLE;
            nee = true;
        }
    } else if (sh.bsdf >= 0) {
        bsdf = sc.bsdfs[sh.bsdf];
        nb_bsdf_resolve<EXT>(sc, bsdf, its.uv);
        const f3 wi = its.sh.to_local(-ps.d);
        float2 s; s.x = ps.rng.next_float(); s.y = ps.rng.next_float();
        f3 lwo; bool discrete; float pdf;
        col = nb_bsdf_sample<EXT>(bsdf, wi, s, lwo, discrete, pdf, sc.params);
        wo = its.sh.to_world(lwo);
        if (MIS) {
            /* bRec.measure after sample(): diffuse sets ESolidAngle and mirror EDiscrete only when
             * wi is on the upper side, dielectric always EDiscrete, microfacet never sets it */
            measure = discrete ? NB_MEASURE_DISCRETE
                    : ((bsdf.type == NB_BSDF_DIFFUSE || bsdf.type == NB_BSDF_DISNEY) && wi.z > 0.f) ? NB_MEASURE_SOLID_ANGLE : NB_MEASURE_UNKNOWN;
            ps.pdfmat = pdf;
            nee = measure == NB_MEASURE_SOLID_ANGLE;                     /* skip delta bsdfs */
        }
    }
    if (MIS && nee) ps.Li += nb_vol_sample_emitter<FLAT, EXT>(sc, ps, medium, its, sh, mi ? nullptr : &bsdf, mi ? &med : nullptr, n_closest);
    ps.thr = ps.thr * col;
    if (!mi && dot(ps.d, wo) > 0.f) medium = nb_medium_behind(sc, sh, its.geo_n, wo);   /* transmitted: entering or leaving */
    ps.o = p; ps.d = wo; ps.mint = NB_EPS; ps.maxt = INFINITY;
    ps.flags = ((bounce + 1) << NB_BOUNCE_SHIFT) | (measure << NB_MEASURE_SHIFT);
    return !is_black(ps.thr);
}

/* ---- work items ------------------------------------------------------------------------------
 * A warp's position in the image: the current item (16x8-pixel tile x a chunk of sample passes)
 * and how much of it is handed out. All members are warp-uniform. */
struct NbItemCursor {
    uint32_t x0 = 0, y0 = 0, tw = 1, tpx = 0, pass0 = 0, n = 0, next = 0;
    bool exhausted = false;
};

/* Called by all 32 lanes together. Lanes with `want` get the next camera samples of the warp's
 * item (a new item is claimed from the global counter when the current one is used up); returns
 * false for a lane that wanted one but the image is finished. */
template <bool LIST>
__device__ __forceinline__ bool nb_take_sample(NbItemCursor &c, const NbTilePlan &tp, uint32_t *item_counter, bool want,
                                               uint32_t lane, uint32_t lt_mask, uint32_t &pix, uint32_t &sidx) {
    bool got = false;
    uint32_t need = __ballot_sync(0xffffffffu, want);
    while (need) {
        if (c.next == c.n) {
            if (c.exhausted) break;
            uint32_t item = 0;
            if (lane == 0) item = atomicAdd(item_counter, 1u);
            item = __shfl_sync(0xffffffffu, item, 0);
            if (LIST && tp.list) {
                const uint32_t n = __ldg(tp.list_ctl), start = item * NB_LIST_ITEM;
                if (item >= (n + NB_LIST_ITEM - 1) / NB_LIST_ITEM) { c.exhausted = true; break; }
                c.x0 = start; c.n = min(NB_LIST_ITEM, n - start); c.next = 0;
                continue;
            }
            if (item >= tp.n_items) { c.exhausted = true; break; }
            /* chunk-major order: concurrently running warps work on different tiles */
            const uint32_t tile = item % tp.n_tiles, chunk = item / tp.n_tiles;
            c.x0 = (tile % tp.tiles_x) * tp.tile_w; c.y0 = (tile / tp.tiles_x) * tp.tile_h;
            c.tw = min(tp.tile_w, tp.plan.width - c.x0);
            c.tpx = c.tw * min(tp.tile_h, tp.height - c.y0);
            c.pass0 = chunk * tp.chunk;
            c.n = c.tpx * min(tp.chunk, tp.passes - c.pass0);
            c.next = 0;
        }
        const uint32_t take = min((uint32_t) __popc(need), c.n - c.next);
        const uint32_t rank = __popc(need & lt_mask);
        if (want && !got && rank < take) {
            const uint32_t k = c.next + rank;
            if (LIST && tp.list) {
                pix = __ldg(tp.list + c.x0 + k);
                sidx = __ldg(tp.list_ctl + 1) + c.x0 + k;      /* every entry of every pass has its own random stream */
            } else {
                const uint32_t pl = k / c.tpx, q = k - pl * c.tpx;
                const uint32_t y = q / c.tw, x = q - y * c.tw;
                pix = (c.y0 + y) * tp.plan.width + c.x0 + x;
                sidx = tp.plan.first + (c.pass0 + pl) * tp.plan.stride;
            }
            got = true;
        }
        c.next += take;
        need = __ballot_sync(0xffffffffu, want && !got);
    }
    return got;
}

/* ---- K_tile ---------------------------------------------------------------------------------- */
/* Occupancy of the EXT variants, measured (B200, M samples/s at 6 / 8 CTAs per SM): normal-map direct_mis 3146 / 3444,
 * disney direct_mis 1728 / 1977, environment-map path_mis 11.42 / 11.36 G. path_vol_mis with the EXT code
 * (heterogeneous media, volume lights, textures) is the largest kernel (119 KB) and waits on instruction fetch more
 * than on anything else (ncu: 11 `stall_no_instruction` per issue at 6 CTAs/SM), so resident warps pay even at
 * the price of spills: fluid scene / volume-light scene at 6 CTAs/SM (80 regs) 203 / 810, 8 (64) 220 / 875,
 * 10 (48) 236 / 936, 12 (40) 249 / 975, 16 (32) 257 / 1007. path_vol_mats with EXT loses 3-5 % beyond 6 on the
 * fluid scene, and the kernels without EXT lose 2-4 % at 8. At 10 CTAs/SM direct_mis gains another 11-12 % (normal maps
 * 3880, disney 2198) while path_mis / path_mats / preview lose 2-15 % (flat and walk kernels alike), at 12 all lose. */
#ifndef NB_TILE_MIN_BLOCKS_VOL_MIS_EXT
#define NB_TILE_MIN_BLOCKS_VOL_MIS_EXT 16
#endif
template <int INTEG, bool FLAT, bool EXT>
__global__ void __launch_bounds__(NB_TILE_THREADS, !EXT ? (INTEG == NB_INTEGRATOR_PATH_MIS ? NB_TILE_MIN_BLOCKS + 1           /* see NB_TILE_MIN_BLOCKS: 8 for path_mis, */
                                                          : INTEG == NB_INTEGRATOR_PATH_VOL_MIS ? NB_TILE_MIN_BLOCKS - 1 : NB_TILE_MIN_BLOCKS)   /* 6 for path_vol_mis */
                                                   : INTEG == NB_INTEGRATOR_PATH_VOL_MIS ? NB_TILE_MIN_BLOCKS_VOL_MIS_EXT
                                                   : INTEG == NB_INTEGRATOR_PATH_VOL_MATS ? 6
                                                   : INTEG == NB_INTEGRATOR_DIRECT_MIS ? NB_TILE_MIN_BLOCKS_EXT + 2 : NB_TILE_MIN_BLOCKS_EXT)
nb_k_tile(NbDeviceScene sc, NbFilm film, NbTilePlan tp, uint32_t *item_counter, unsigned long long *stats) {
    constexpr bool PATH = INTEG == NB_INTEGRATOR_PATH_MIS || INTEG == NB_INTEGRATOR_PATH_MATS;
    constexpr bool VOL = INTEG == NB_INTEGRATOR_PATH_VOL_MIS || INTEG == NB_INTEGRATOR_PATH_VOL_MATS;
    constexpr bool MIS = INTEG == NB_INTEGRATOR_PATH_MIS || INTEG == NB_INTEGRATOR_PATH_VOL_MIS;
    const uint32_t lane = threadIdx.x & 31u, lt_mask = (1u << lane) - 1u;
    uint32_t n_closest = 0, n_any = 0, n_samples = 0;
    NbItemCursor cur;
    bool active = false;
    NbPathRegs ps;
    ps.flags = 0;
#ifdef NB_FUSED_SCAN
    bool pend = false, fin = false;         /* an NEE ray waits for the next pass over the slots; its path has ended and waits for it */
    NbShadowRay pso;
#else
    const bool pend = false, fin = false;
#endif
    int32_t medium = sc.ambient_medium;     /* volume integrators: the medium the path is in */
#ifdef NB_SMEM_SLOTS
    /* experiment: the flat scene's slots staged in shared memory once per CTA (broadcast LDS instead of per-lane LDG) */
    __shared__ float4 sm_slots[3 * NB_FLAT_MAX_SLOTS];
    if (FLAT) {
        for (uint32_t i = threadIdx.x; i < 3 * sc.n_slots && i < 3 * NB_FLAT_MAX_SLOTS; i += blockDim.x) sm_slots[i] = sc.slots[i];
        __syncthreads();
        if (sc.n_slots <= NB_FLAT_MAX_SLOTS) sc.slots = sm_slots;
    }
#endif
    for (;;) {
        /* --- regenerate: idle lanes take the next camera samples of the warp's item --- */
        uint32_t pix, sidx;
        if (nb_take_sample<EXT>(cur, tp, item_counter, !active && !fin, lane, lt_mask, pix, sidx)) {
            NbCameraRay r;
            nb_camera_sample(sc.camera, tp.plan, pix, sidx, ps.rng, r);
            ps.o = r.o; ps.d = r.d; ps.mint = r.mint; ps.maxt = r.maxt; ps.px = r.px; ps.py = r.py;
            ps.Li = mk3(0.f); ps.thr = mk3(1.f); ps.w_mats = 1.f; ps.w_ems = 0.f; ps.pdfmat = 0.f;
            ps.flags = NB_FLAG_PRIMARY;
            if (VOL) {                       /* path_vol_mis.cpp:118-121: the camera acts like a delta bsdf */
                ps.pdfmat = 1.f; ps.flags = NB_MEASURE_DISCRETE << NB_MEASURE_SHIFT;
                medium = sc.ambient_medium;
            }
            active = true;
            ++n_samples;
        }
        if (!__any_sync(0xffffffffu, active || pend)) break;

#ifdef NB_FUSED_SCAN
        if (PATH && FLAT && MIS) {
            /* experiment: the NEE ray of the last bounce and the path's next ray share one pass over the slots (nb_scan2) */
            NbHit hit;
            bool occluded = false;
            if (active) ++n_closest;
            if (pend) ++n_any;
            const bool found = nb_scan2(sc.slots, sc.n_slots, active, ps.o, ps.d, ps.mint, ps.maxt, hit, pend, pso.o, pso.d, pso.mint, pso.maxt, occluded);
            if (pend) {
                if (!occluded) { if (pso.cand >= 0.f) ps.w_ems = pso.cand; ps.Li += ps.w_ems * pso.c; }
                pend = false;
            }
            if (fin) { nb_film_put<EXT>(film, ps.px, ps.py, ps.Li, stats); fin = false; }
            if (active) {
                bool ended = false, shadow = false;
                if (!found) { ps.Li += ps.thr * nb_env_radiance<EXT>(sc, ps.d); ended = true; }
                else ended = !nb_path_bounce<MIS, EXT>(sc, tp.plan.max_depth, ps, hit, shadow, pso);
                pend = shadow;
                if (ended) {
                    active = false;
                    if (pend) fin = true; else nb_film_put<EXT>(film, ps.px, ps.py, ps.Li, stats);
                }
            }
        } else
#endif
        if (PATH) {
            bool ended = false, shadow = false;
            NbShadowRay so;
            if (active) {
                NbHit hit;
                ++n_closest;
                if (!nb_closest<FLAT>(sc, ps.o, ps.d, ps.mint, ps.maxt, hit)) {
                    ps.Li += ps.thr * nb_env_radiance<EXT>(sc, ps.d);       /* path_mis.cpp:35-43, path_mats.cpp:26-34 */
                    ended = true;
                } else ended = !nb_path_bounce<MIS, EXT>(sc, tp.plan.max_depth, ps, hit, shadow, so);
            }
            if (MIS && shadow) {
                ++n_any;
                if (!nb_occluded<FLAT>(sc, so.o, so.d, so.mint, so.maxt)) {
                    /* path_mis.cpp:102-105,141: w_ems is updated only when the pdf sum passes the
                     * Epsilon test, otherwise the value of an earlier bounce is used */
                    if (so.cand >= 0.f) ps.w_ems = so.cand;
                    ps.Li += ps.w_ems * so.c;
                }
            }
            if (active && ended) {
                nb_film_put<EXT>(film, ps.px, ps.py, ps.Li, stats);
                active = false;
            }
        } else if (VOL) {
            if (active) {
                NbHit hit;
                ++n_closest;
                bool alive = nb_closest<FLAT>(sc, ps.o, ps.d, ps.mint, ps.maxt, hit);
                if (alive) alive = nb_vol_bounce<MIS, FLAT, EXT>(sc, tp.plan.max_depth, ps, medium, hit, n_closest);
                else ps.Li += ps.thr * nb_env_radiance<EXT>(sc, ps.d);      /* path_vol_mats.cpp:34-39 */
                if (!alive) {
                    nb_film_put<EXT>(film, ps.px, ps.py, ps.Li, stats);
                    active = false;
                }
            }
        } else if (active) {
            f3 layer_normal;
            const f3 Li = nb_direct_sample<INTEG, FLAT, EXT>(sc, tp.plan, ps, n_closest, n_any, layer_normal);
            nb_film_put<EXT>(film, ps.px, ps.py, Li, stats);
            if (INTEG == NB_INTEGRATOR_PREVIEW && film.normal) nb_film_put(film, ps.px, ps.py, layer_normal, stats, film.normal);
            active = false;
        }
    }
    for (int off = 16; off; off >>= 1) {
        n_closest += __shfl_xor_sync(0xffffffffu, n_closest, off);
        n_any += __shfl_xor_sync(0xffffffffu, n_any, off);
        n_samples += __shfl_xor_sync(0xffffffffu, n_samples, off);
    }
    if (lane == 0) {
        if (n_closest) atomicAdd(&stats[NB_STAT_RAYS_CLOSEST], (unsigned long long) n_closest);
        if (n_any) atomicAdd(&stats[NB_STAT_RAYS_ANY], (unsigned long long) n_any);
        if (n_samples) atomicAdd(&stats[NB_STAT_SAMPLES], (unsigned long long) n_samples);
    }
}

/* ---- K_tile_walk: the same integrators for scenes with a real tree ----------------------------
 * With a tree, rays of one warp need wildly different numbers of node visits (C3, 1.31 M
 * triangles: mean 38, tail in the hundreds). K_tile's lock-step loop then runs the traversal
 * with 5 of 32 lanes active (ncu), every bounce waiting for the warp's slowest ray. Here each
 * lane is its own little state machine
 *     REGEN -> NODE <-> LEAF -> TDONE -> (NODE ... shadow ray, then main ray) -> SHADE -> ...
 * and the warp loop offers each phase's code in turn: node steps every iteration (cheap, most
 * lanes want them), the expensive phases (leaf intersection, shading, splat + regeneration) only
 * once enough lanes wait for them or nothing else can run. A lane with a long ray keeps walking
 * while its neighbours shade, start new paths and walk their own rays. Per ray the visiting order
 * and arithmetic are nb_traverse's; per path the draws are nb_path_bounce's. */
#ifndef NB_WALK_MIN_BLOCKS
#define NB_WALK_MIN_BLOCKS 7      /* 72 registers: the path state is parked in shared memory during the walk. B200, M samples/s at 4 / 6 / 7 / 8 CTAs per SM:
                                   * table 332 / 355 / 369 / 356, C3 (1.31 M triangles) - / 444 / 479 / 477 (three interleaved runs each) */
#endif
#ifndef NB_WALK_NODE_STEPS
#define NB_WALK_NODE_STEPS 8      /* node steps per loop iteration */
#endif
#ifndef NB_WALK_T_LEAF
#define NB_WALK_T_LEAF 6          /* lanes that must wait before a phase is run (unless nothing else can) */
#endif
#ifndef NB_WALK_T_SHADE
#define NB_WALK_T_SHADE 16        /* B200 at 7 CTAs/SM, M samples/s at 8 / 12 / 16 / 20: table 365 / 386 / 402 / 404, C3 479 / 492 / 493 / 466 */
#endif
#ifndef NB_WALK_T_REGEN
#define NB_WALK_T_REGEN 8
#endif
enum { NB_PH_IDLE = 0, NB_PH_REGEN = 1, NB_PH_NODE = 2, NB_PH_LEAF = 3, NB_PH_TDONE = 4, NB_PH_SHADE = 5 };

/* Path state parked in shared memory while a lane walks the tree (SoA, one word per lane per row:
 * conflict-free). The walk itself needs ~25 registers of ray state; the path's 28 words are only
 * touched by REGEN / TDONE / SHADE, so keeping them out of the register file during NODE / LEAF
 * buys the occupancy the latency-bound node loop wants (124 -> NB_WALK_MIN_BLOCKS-bound registers). */
enum { NB_W_LI = 0, NB_W_THR = 3, NB_W_O = 6, NB_W_D = 9, NB_W_MINT = 12, NB_W_MAXT = 13, NB_W_WMATS = 14, NB_W_WEMS = 15,
       NB_W_PDFMAT = 16, NB_W_PX = 17, NB_W_PY = 18, NB_W_FLAGS = 19, NB_W_RNG = 20, NB_W_SC = 24, NB_W_CAND = 27, NB_W_WORDS = 28 };

/* WIDE: the production tree (nb_internal.h NB_WIDE_*, nb_traverse_wide): a NODE step tests the four children of one
 * 128-byte node, continues into the nearest and parks the others, with their entry distances, on a per-lane stack —
 * its first NB_WSTK_SHARED entries in shared memory (one 8-byte column per lane, conflict-free), the rare deeper ones
 * in local memory; a popped entry farther than the current hit is dropped without being fetched. */
#ifndef NB_WSTK_SHARED
#define NB_WSTK_SHARED 8       /* B200, table / C3 M samples/s at 2 / 4 / 8 / 12 / 16: 458 / 451 / 465 / 451 / 479* / -; (*earlier build) shared memory competes with L1 */
#endif
#ifndef NB_WALK_NODE_STEPS_WIDE
#define NB_WALK_NODE_STEPS_WIDE 4
#endif
#ifndef NB_WALK_MIN_BLOCKS_WIDE
#define NB_WALK_MIN_BLOCKS_WIDE 7
#endif

template <bool MIS, bool EXT, bool WIDE>
__global__ void __launch_bounds__(NB_TILE_THREADS, EXT ? NB_WALK_MIN_BLOCKS_EXT : (WIDE ? NB_WALK_MIN_BLOCKS_WIDE : NB_WALK_MIN_BLOCKS))
nb_k_tile_walk(NbDeviceScene sc, NbFilm film, NbTilePlan tp, uint32_t *item_counter, unsigned long long *stats) {
    __shared__ float park[NB_W_WORDS][NB_TILE_THREADS];
    __shared__ uint2 wstk[WIDE ? NB_WSTK_SHARED : 1][NB_TILE_THREADS];
    uint2 wovf[WIDE ? NB_WIDE_STACK - NB_WSTK_SHARED : 1];
    uint32_t sp = 0;
    NbWideRay wr;
    wr.r = make_float3(0.f, 0.f, 0.f); wr.o = wr.r;
    const uint32_t tid = threadIdx.x;
    const uint32_t lane = threadIdx.x & 31u, lt_mask = (1u << lane) - 1u;
    uint32_t n_closest = 0, n_any = 0, n_samples = 0;
    NbItemCursor cur;
    uint32_t phase = NB_PH_REGEN;
    bool splat = false;               /* REGEN: a finished sample waits to be put on the film */
    bool ending = false;              /* the path ended at the last vertex; splat after its shadow ray */
    /* the ray being walked */
    float3 to = make_float3(0.f, 0.f, 0.f), td = to, tr = to;
    float tmint = 0.f, tmaxt = 0.f;
    uint32_t idx = NB_NODE_END, leaf = 0, after = NB_NODE_END;
    NbHit hit;
    hit.t = INFINITY; hit.u = hit.v = 0.f; hit.slot = 0xffffffffu;
    bool t_any = false, occluded = false;

    auto ld3p = [&](int row) { return mk3(park[row][tid], park[row + 1][tid], park[row + 2][tid]); };
    auto st3p = [&](int row, f3 v) { park[row][tid] = v.x; park[row + 1][tid] = v.y; park[row + 2][tid] = v.z; };
    auto park_path = [&](const NbPathRegs &ps) {
        st3p(NB_W_LI, ps.Li); st3p(NB_W_THR, ps.thr); st3p(NB_W_O, ps.o); st3p(NB_W_D, ps.d);
        park[NB_W_MINT][tid] = ps.mint; park[NB_W_MAXT][tid] = ps.maxt; park[NB_W_WMATS][tid] = ps.w_mats;
        park[NB_W_WEMS][tid] = ps.w_ems; park[NB_W_PDFMAT][tid] = ps.pdfmat; park[NB_W_PX][tid] = ps.px; park[NB_W_PY][tid] = ps.py;
        park[NB_W_FLAGS][tid] = __uint_as_float(ps.flags);
        park[NB_W_RNG][tid] = __uint_as_float((uint32_t) ps.rng.state); park[NB_W_RNG + 1][tid] = __uint_as_float((uint32_t) (ps.rng.state >> 32));
        park[NB_W_RNG + 2][tid] = __uint_as_float((uint32_t) ps.rng.inc); park[NB_W_RNG + 3][tid] = __uint_as_float((uint32_t) (ps.rng.inc >> 32));
    };
    auto unpark_path = [&](NbPathRegs &ps) {
        ps.Li = ld3p(NB_W_LI); ps.thr = ld3p(NB_W_THR); ps.o = ld3p(NB_W_O); ps.d = ld3p(NB_W_D);
        ps.mint = park[NB_W_MINT][tid]; ps.maxt = park[NB_W_MAXT][tid]; ps.w_mats = park[NB_W_WMATS][tid];
        ps.w_ems = park[NB_W_WEMS][tid]; ps.pdfmat = park[NB_W_PDFMAT][tid]; ps.px = park[NB_W_PX][tid]; ps.py = park[NB_W_PY][tid];
        ps.flags = __float_as_uint(park[NB_W_FLAGS][tid]);
        ps.rng.state = (uint64_t) __float_as_uint(park[NB_W_RNG][tid]) | ((uint64_t) __float_as_uint(park[NB_W_RNG + 1][tid]) << 32);
        ps.rng.inc = (uint64_t) __float_as_uint(park[NB_W_RNG + 2][tid]) | ((uint64_t) __float_as_uint(park[NB_W_RNG + 3][tid]) << 32);
    };

    /* prologue of nb_traverse for the lane's next ray */
    auto begin_walk = [&](f3 o, f3 d, float mint, float maxt, bool any_hit) {
        to = o; td = d; t_any = any_hit; occluded = false;
        hit.t = INFINITY; hit.u = 0.f; hit.v = 0.f; hit.slot = 0xffffffffu;
        if (mint == NB_EPSILON) {
            float m = fabsf(o.x);
            m = fmaxf(m, fabsf(o.y));
            m = fmaxf(m, fabsf(o.z));
            const float e = __fmul_rn(mint, m);
            mint = (mint < e) ? e : mint;
        }
        tmint = mint; tmaxt = maxt;
        if constexpr (WIDE) {
            wr = nb_wide_ray(o, d);
            sp = 0;
            idx = (sc.n_wide == 0 || maxt < mint) ? NB_NODE_END : 0u;
        } else {
            tr = make_float3(__fdiv_rn(1.0f, d.x), __fdiv_rn(1.0f, d.y), __fdiv_rn(1.0f, d.z));
            idx = (sc.n_nodes == 0 || maxt < mint) ? NB_NODE_END : 0u;
        }
        phase = idx == NB_NODE_END ? NB_PH_TDONE : NB_PH_NODE;
        if (any_hit) ++n_any; else ++n_closest;
    };
    auto walk_main = [&]() { begin_walk(ld3p(NB_W_O), ld3p(NB_W_D), park[NB_W_MINT][tid], park[NB_W_MAXT][tid], false); };
    auto wpush = [&](uint32_t ref, float t) {
        const uint2 e = make_uint2(ref, __float_as_uint(t));
        if (sp < NB_WSTK_SHARED) wstk[sp][tid] = e; else wovf[sp - NB_WSTK_SHARED] = e;
        ++sp;
    };
    /* next stack entry the ray can still reach; sets the lane's phase (NODE at an inner node, LEAF at a leaf, TDONE when none is left) */
    auto wnext = [&]() {
        phase = NB_PH_TDONE;
        while (sp) {
            --sp;
            const uint2 e = sp < NB_WSTK_SHARED ? wstk[sp][tid] : wovf[sp - NB_WSTK_SHARED];
            if (__uint_as_float(e.y) <= tmaxt) {
                if (e.x & NB_WIDE_LEAF) { leaf = e.x & ~NB_WIDE_LEAF; phase = NB_PH_LEAF; }
                else { idx = e.x; phase = NB_PH_NODE; }
                break;
            }
        }
    };

    for (;;) {
        const uint32_t m_walking = __ballot_sync(0xffffffffu, phase == NB_PH_NODE || phase == NB_PH_LEAF || phase == NB_PH_TDONE);
        /* ---- REGEN: splat the finished sample, start the next camera path ---- */
        const uint32_t m_regen = __ballot_sync(0xffffffffu, phase == NB_PH_REGEN);
        if (m_regen && (__popc(m_regen) >= NB_WALK_T_REGEN || !m_walking)) {
            if (phase == NB_PH_REGEN && splat) { nb_film_put<EXT>(film, park[NB_W_PX][tid], park[NB_W_PY][tid], ld3p(NB_W_LI), stats); splat = false; }
            uint32_t pix, sidx;
            const bool got = nb_take_sample<EXT>(cur, tp, item_counter, phase == NB_PH_REGEN, lane, lt_mask, pix, sidx);
            if (got) {
                NbCameraRay r;
                NbPathRegs ps;
                nb_camera_sample(sc.camera, tp.plan, pix, sidx, ps.rng, r);
                ps.o = r.o; ps.d = r.d; ps.mint = r.mint; ps.maxt = r.maxt; ps.px = r.px; ps.py = r.py;
                ps.Li = mk3(0.f); ps.thr = mk3(1.f); ps.w_mats = 1.f; ps.w_ems = 0.f; ps.pdfmat = 0.f;
                ps.flags = NB_FLAG_PRIMARY;
                park_path(ps);
                ending = false;
                ++n_samples;
                begin_walk(ps.o, ps.d, ps.mint, ps.maxt, false);
            } else if (phase == NB_PH_REGEN) phase = NB_PH_IDLE;      /* image finished */
        }
        if (!__ballot_sync(0xffffffffu, phase != NB_PH_IDLE)) break;

        /* ---- NODE: inner-node steps until the lane stands on a leaf or its walk ends ---- */
#pragma unroll 1
        for (int step = 0; step < (WIDE ? NB_WALK_NODE_STEPS_WIDE : NB_WALK_NODE_STEPS); ++step) {
            if constexpr (WIDE) {
                if (phase == NB_PH_NODE) {
                    float t4[4]; uint32_t r4[4];
                    const int n = nb_wide_children(sc.wide + 8 * (size_t) idx, wr, tmint, tmaxt, t4, r4);
                    if (n > 3) wpush(r4[3], t4[3]);
                    if (n > 2) wpush(r4[2], t4[2]);
                    if (n > 1) wpush(r4[1], t4[1]);
                    if (n > 0) {
                        if (r4[0] & NB_WIDE_LEAF) { leaf = r4[0] & ~NB_WIDE_LEAF; phase = NB_PH_LEAF; }
                        else idx = r4[0];
                    } else wnext();
                }
            } else
            if (phase == NB_PH_NODE) {
                const float4 A = __ldg(sc.nodes + 2 * idx), B = __ldg(sc.nodes + 2 * idx + 1);
                const uint32_t skip = __float_as_uint(A.x), start = __float_as_uint(A.y);
                /* left-first with skip links. The near-child-first walk (nb_traverse_ordered, bit-exact too) was
                 * tried here with a per-lane stack: B200, table 336 -> 309 M samples/s, C3 425 -> 401 M, C5 (SAH
                 * tree) 192 -> 212 M, C5 (LBVH) 144 -> 144 M — the stack traffic costs what the saved nodes give */
                if (!nb_box_hit(A, B, to, td, tr, tmint, tmaxt)) idx = skip;
                else if (start & NB_NODE_INNER_BIT) idx = idx + 1;
                else {
                    idx = skip;
                    if (start != NB_NODE_EMPTY_LEAF) { leaf = start; after = skip; phase = NB_PH_LEAF; }
                }
                if (phase == NB_PH_NODE && idx == NB_NODE_END) phase = NB_PH_TDONE;
            }
        }
        /* ---- LEAF: lanes standing on a leaf intersect it together ---- */
        const uint32_t m_leaf = __ballot_sync(0xffffffffu, phase == NB_PH_LEAF);
        if (m_leaf && (__popc(m_leaf) >= NB_WALK_T_LEAF || !__ballot_sync(0xffffffffu, phase == NB_PH_NODE))) {
            if (phase == NB_PH_LEAF) {
                uint32_t s = leaf;
                while (true) {
                    const float4 s0 = __ldg(sc.slots + 3 * s), s1 = __ldg(sc.slots + 3 * s + 1), s2 = __ldg(sc.slots + 3 * s + 2);
                    const uint32_t tag = __float_as_uint(s2.y);
                    float u = 0.f, v = 0.f, t;
                    const bool ok = (tag & NB_SLOT_SPHERE) ? nb_sphere_hit(s0, to, td, tmint, tmaxt, t)
                                                           : nb_tri_hit(s0, s1, s2, to, td, tmint, tmaxt, u, v, t, WIDE && NB_RENDER_FAST_HITS);
                    if (ok) {
                        if (t_any) { occluded = true; break; }
                        /* the wide walk meets primitives out of slot order: an exact-t tie keeps the later slot, as the reference's
                         * order does. (One primitive per lane per step, popping the stack in between, was tried to lift the 7 of 32
                         * lanes this loop runs at: B200, table 486 -> 451, C3 690 -> 686 M samples/s — the per-step ballots and pops cost more) */
                        if (!WIDE || !(hit.slot != 0xffffffffu && t == hit.t && s < hit.slot)) {
                            tmaxt = t;
                            hit.t = t; hit.u = u; hit.v = v; hit.slot = s;
                        }
                    }
                    if (tag & NB_SLOT_LAST) break;
                    ++s;
                }
                if constexpr (WIDE) {
                    if (occluded) phase = NB_PH_TDONE; else wnext();
                } else {
                    idx = occluded ? NB_NODE_END : after;
                    phase = idx == NB_NODE_END ? NB_PH_TDONE : NB_PH_NODE;
                }
            }
        }
        /* ---- TDONE: a walk ended. Shadow ray: add its contribution, go on to the main ray
         * (or end the path); main ray: queue for shading ---- */
        if (phase == NB_PH_TDONE) {
            if (t_any) {
                if (!occluded) {                           /* path_mis.cpp:102-105,141 */
                    const float cand = park[NB_W_CAND][tid];
                    float w_ems = park[NB_W_WEMS][tid];
                    if (cand >= 0.f) { w_ems = cand; park[NB_W_WEMS][tid] = cand; }
                    st3p(NB_W_LI, ld3p(NB_W_LI) + w_ems * ld3p(NB_W_SC));
                }
                if (ending) { splat = true; phase = NB_PH_REGEN; }
                else walk_main();
            } else phase = NB_PH_SHADE;
        }
        /* ---- SHADE: one bounce of the integrator for lanes whose closest hit arrived ---- */
        const uint32_t m_shade = __ballot_sync(0xffffffffu, phase == NB_PH_SHADE);
        if (m_shade && (__popc(m_shade) >= NB_WALK_T_SHADE || !__ballot_sync(0xffffffffu, phase == NB_PH_NODE || phase == NB_PH_LEAF))) {
            if (phase == NB_PH_SHADE) {
                bool alive = false, shadow = false;
                NbPathRegs ps;
                NbShadowRay so;
                unpark_path(ps);
                if (hit.slot != 0xffffffffu) alive = nb_path_bounce<MIS, EXT>(sc, tp.plan.max_depth, ps, hit, shadow, so);
                else ps.Li += ps.thr * nb_env_radiance<EXT>(sc, ps.d);
                park_path(ps);
                ending = !alive;
                if (MIS && shadow) {
                    st3p(NB_W_SC, so.c); park[NB_W_CAND][tid] = so.cand;
                    begin_walk(so.o, so.d, so.mint, so.maxt, true);
                } else if (alive) begin_walk(ps.o, ps.d, ps.mint, ps.maxt, false);
                else { splat = true; phase = NB_PH_REGEN; }
            }
        }
    }
    for (int off = 16; off; off >>= 1) {
        n_closest += __shfl_xor_sync(0xffffffffu, n_closest, off);
        n_any += __shfl_xor_sync(0xffffffffu, n_any, off);
        n_samples += __shfl_xor_sync(0xffffffffu, n_samples, off);
    }
    if (lane == 0) {
        if (n_closest) atomicAdd(&stats[NB_STAT_RAYS_CLOSEST], (unsigned long long) n_closest);
        if (n_any) atomicAdd(&stats[NB_STAT_RAYS_ANY], (unsigned long long) n_any);
        if (n_samples) atomicAdd(&stats[NB_STAT_SAMPLES], (unsigned long long) n_samples);
    }
}

template <bool MIS, bool EXT, bool WIDE>
static cudaError_t nb_tile_launch_walk(const NbDeviceScene &sc, const NbFilm &film, const NbTilePlan &tp, uint32_t *item_counter,
                                       unsigned long long *stats, int sm_count, cudaStream_t stream) {
    int per_sm = 0;
    cudaError_t e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, nb_k_tile_walk<MIS, EXT, WIDE>, NB_TILE_THREADS, 0);
    if (e != cudaSuccess) return e;
    if (per_sm < 1) per_sm = 1;
    uint64_t blocks = (uint64_t) sm_count * per_sm;
    const uint64_t blocks_useful = ((uint64_t) tp.n_items + NB_TILE_THREADS / 32 - 1) / (NB_TILE_THREADS / 32);
    if (blocks > blocks_useful) blocks = blocks_useful;
    if (blocks < 1) blocks = 1;
    nb_k_tile_walk<MIS, EXT, WIDE><<<(unsigned) blocks, NB_TILE_THREADS, 0, stream>>>(sc, film, tp, item_counter, stats);
    return cudaGetLastError();
}

template <int INTEG, bool FLAT, bool EXT>
static cudaError_t nb_tile_launch_flat(const NbDeviceScene &sc, const NbFilm &film, const NbTilePlan &tp, uint32_t *item_counter,
                                    unsigned long long *stats, int sm_count, cudaStream_t stream) {
    int per_sm = 0;
    cudaError_t e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, nb_k_tile<INTEG, FLAT, EXT>, NB_TILE_THREADS, 0);
    if (e != cudaSuccess) return e;
    if (per_sm < 1) per_sm = 1;
    const uint64_t warps_wanted = ((uint64_t) tp.n_items + 0) ;
    uint64_t blocks = (uint64_t) sm_count * per_sm;
    const uint64_t blocks_useful = (warps_wanted + NB_TILE_THREADS / 32 - 1) / (NB_TILE_THREADS / 32);
    if (blocks > blocks_useful) blocks = blocks_useful;
    if (blocks < 1) blocks = 1;
    nb_k_tile<INTEG, FLAT, EXT><<<(unsigned) blocks, NB_TILE_THREADS, 0, stream>>>(sc, film, tp, item_counter, stats);
    return cudaGetLastError();
}

template <int INTEG>
static cudaError_t nb_tile_launch_t(const NbDeviceScene &sc, const NbFilm &film, const NbTilePlan &tp, uint32_t *item_counter,
                                    unsigned long long *stats, int sm_count, cudaStream_t stream) {
#ifdef NB_PARAM_SLOTS
    const bool flat = sc.flat && sc.flat_in_params;      /* the flat kernels read the slots from their parameters */
#else
    const bool flat = sc.flat != 0;
#endif
    if (sc.ext)
        return flat ? nb_tile_launch_flat<INTEG, true, true>(sc, film, tp, item_counter, stats, sm_count, stream)
                       : nb_tile_launch_flat<INTEG, false, true>(sc, film, tp, item_counter, stats, sm_count, stream);
    return flat ? nb_tile_launch_flat<INTEG, true, false>(sc, film, tp, item_counter, stats, sm_count, stream)
                   : nb_tile_launch_flat<INTEG, false, false>(sc, film, tp, item_counter, stats, sm_count, stream);
}

/* Host entry used by nb_render_subframe. `passes` sample passes over the image, pass k drawing
 * sample index plan.first + k * plan.stride. Asynchronous on `stream`. */
cudaError_t nb_launch_tile(const NbDeviceScene &sc_in, const NbFilm &film, const NbPlan &plan, uint32_t height, uint32_t passes,
                           uint32_t integrator, uint32_t *item_counter, unsigned long long *stats, int sm_count,
                           cudaStream_t stream, bool walk, const uint32_t *list, const uint32_t *list_ctl) {
    NbDeviceScene sc = sc_in;
    if (list) sc.ext = 1u;             /* the list and the block images are read by the EXT variants only (nb_film.cuh) */
    NbTilePlan tp;
    tp.plan = plan;
    tp.list = list; tp.list_ctl = list_ctl;
    tp.height = height; tp.passes = passes;
    /* Work items are small: a warp that draws an expensive item last (glass-heavy tiles cost several
     * times the mean) finishes alone while the rest of the GPU idles, and at a few hundred k samples/s
     * per warp one 4096-sample item is ~10 ms of a ~100 ms render. Measured on B200 (table, 22 766
     * triangles, 64 spp): 32 passes per item 226 M samples/s, 8: 228 M, 4: 280 M, 2: 319 M, 1: 332 M;
     * 8x8 tiles 359 M. Lanes pick up the next item as soon as the current one is handed out
     * (nb_take_sample), so small items cost one atomic each and nothing else. The lock-step kernel
     * keeps 16x8 tiles x 2 passes: its camera rays run coherently (cbox 1.96 G vs 1.94 G at 1 pass). */
#ifdef NB_PARAM_SLOTS
    const bool is_flat = sc.flat && sc.flat_in_params;
#else
    const bool is_flat = sc.flat != 0;
#endif
    const bool walking = !is_flat && walk && (integrator == NB_INTEGRATOR_PATH_MIS || integrator == NB_INTEGRATOR_PATH_MATS);
#ifndef NB_TILE_CHUNK_MAX
#define NB_TILE_CHUNK_MAX 2
#endif
#ifndef NB_WALK_CHUNK_MAX
#define NB_WALK_CHUNK_MAX 1
#endif
    tp.tile_w = walking ? 8 : 16; tp.tile_h = 8;
    tp.tiles_x = (plan.width + tp.tile_w - 1) / tp.tile_w;
    tp.n_tiles = tp.tiles_x * ((height + tp.tile_h - 1) / tp.tile_h);
    uint32_t chunk = walking ? NB_WALK_CHUNK_MAX : NB_TILE_CHUNK_MAX;
    while ((uint64_t) tp.n_tiles * ((passes + chunk - 1) / chunk) >= 0xffff0000ull) chunk <<= 1;   /* item ids are 32-bit */
    tp.chunk = chunk;
    const uint64_t items = (uint64_t) tp.n_tiles * ((passes + chunk - 1) / chunk);
    if (items >= 0xffff0000ull) return cudaErrorInvalidValue;
    tp.n_items = (uint32_t) items;
    cudaError_t e = cudaMemsetAsync(item_counter, 0, sizeof(uint32_t), stream);
    if (e != cudaSuccess) return e;
    if (walking) {
        const bool mis = integrator == NB_INTEGRATOR_PATH_MIS;
        if (sc.use_wide) {
            if (mis) return sc.ext ? nb_tile_launch_walk<true, true, true>(sc, film, tp, item_counter, stats, sm_count, stream)
                                   : nb_tile_launch_walk<true, false, true>(sc, film, tp, item_counter, stats, sm_count, stream);
            return sc.ext ? nb_tile_launch_walk<false, true, true>(sc, film, tp, item_counter, stats, sm_count, stream)
                          : nb_tile_launch_walk<false, false, true>(sc, film, tp, item_counter, stats, sm_count, stream);
        }
        if (mis) return sc.ext ? nb_tile_launch_walk<true, true, false>(sc, film, tp, item_counter, stats, sm_count, stream)
                               : nb_tile_launch_walk<true, false, false>(sc, film, tp, item_counter, stats, sm_count, stream);
        return sc.ext ? nb_tile_launch_walk<false, true, false>(sc, film, tp, item_counter, stats, sm_count, stream)
                      : nb_tile_launch_walk<false, false, false>(sc, film, tp, item_counter, stats, sm_count, stream);
    }
    switch (integrator) {
        case NB_INTEGRATOR_PATH_MIS: return nb_tile_launch_t<NB_INTEGRATOR_PATH_MIS>(sc, film, tp, item_counter, stats, sm_count, stream);
        case NB_INTEGRATOR_PATH_MATS: return nb_tile_launch_t<NB_INTEGRATOR_PATH_MATS>(sc, film, tp, item_counter, stats, sm_count, stream);
        case NB_INTEGRATOR_DIRECT: return nb_tile_launch_t<NB_INTEGRATOR_DIRECT>(sc, film, tp, item_counter, stats, sm_count, stream);
        case NB_INTEGRATOR_DIRECT_EMS: return nb_tile_launch_t<NB_INTEGRATOR_DIRECT_EMS>(sc, film, tp, item_counter, stats, sm_count, stream);
        case NB_INTEGRATOR_DIRECT_MATS: return nb_tile_launch_t<NB_INTEGRATOR_DIRECT_MATS>(sc, film, tp, item_counter, stats, sm_count, stream);
        case NB_INTEGRATOR_DIRECT_MIS: return nb_tile_launch_t<NB_INTEGRATOR_DIRECT_MIS>(sc, film, tp, item_counter, stats, sm_count, stream);
        case NB_INTEGRATOR_PATH_VOL_MATS: return nb_tile_launch_t<NB_INTEGRATOR_PATH_VOL_MATS>(sc, film, tp, item_counter, stats, sm_count, stream);
        case NB_INTEGRATOR_PATH_VOL_MIS: return nb_tile_launch_t<NB_INTEGRATOR_PATH_VOL_MIS>(sc, film, tp, item_counter, stats, sm_count, stream);
        case NB_INTEGRATOR_NORMALS: return nb_tile_launch_t<NB_INTEGRATOR_NORMALS>(sc, film, tp, item_counter, stats, sm_count, stream);
        case NB_INTEGRATOR_AV: return nb_tile_launch_t<NB_INTEGRATOR_AV>(sc, film, tp, item_counter, stats, sm_count, stream);
        case NB_INTEGRATOR_PREVIEW: return nb_tile_launch_t<NB_INTEGRATOR_PREVIEW>(sc, film, tp, item_counter, stats, sm_count, stream);
        default: return cudaErrorInvalidValue;
    }
}
