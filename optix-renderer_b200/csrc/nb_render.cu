/* nb_render.cu — K_shade (ray generation, per-bounce shading, NEE + MIS, Russian roulette, queue
 * compaction, film splat) and the host side of the wavefront loop.
 *
 * Restates, per path and per bounce, PathMISIntegrator::Li (src/integrators/path_mis.cpp:16-150)
 * and PathMATSIntegrator::Li (path_mats.cpp:16-78) on top of renderBlock (src/utils/render.cpp:
 * 421-459), PerspectiveCamera::sampleRay (src/cameras/perspective.cpp:97-141) and
 * ImageBlock::put (src/utils/block.cpp:93-123). The reference traces the BSDF-sampled ray twice
 * (MIS probe path_mis.cpp:119, then as the next main ray :34); here it is traced once and the
 * probe's MIS weight is computed when the hit arrives, from state carried in the path.
 */
#include "nb_internal.h"
#include "nb_wavefront.h"
#include "nb_shading.cuh"
#include "nb_film.cuh"
#include <cstring>
#include <algorithm>

/* start camera sample n of the plan in pool slot p */
__device__ __forceinline__ void nb_generate_path(const NbDeviceScene &sc, const NbWave &w, const NbPlan &plan,
                                                 uint32_t p, unsigned long long n) {
    const uint32_t pass = (uint32_t) (n / plan.npix), pix = (uint32_t) (n % plan.npix);
    const uint32_t sidx = plan.first + pass * plan.stride;
    NbRng rng;
    NbCameraRay r;
    nb_camera_sample(sc.camera, plan, pix, sidx, rng, r);
    w.ray_o[p] = make_float4(r.o.x, r.o.y, r.o.z, r.mint);
    w.ray_d[p] = make_float4(r.d.x, r.d.y, r.d.z, r.maxt);
    w.st0[p] = make_float4(0.f, 0.f, 0.f, 1.f);
    w.st1[p] = make_float4(1.f, 1.f, 1.f, 0.f);
    w.st2[p] = make_float4(r.px, r.py, 0.f, __uint_as_float(NB_FLAG_PRIMARY));
    w.rng[p] = make_uint4((uint32_t) rng.state, (uint32_t) (rng.state >> 32), pix, sidx);
}

/* ---- K_shade ---------------------------------------------------------------------------------
 * One thread per live path slot. MIS = path_mis.cpp, !MIS = path_mats.cpp. Ends with two
 * warp-ballot compactions: surviving / regenerated paths -> queue_next, NEE rays -> shadow queue
 * (one atomicAdd per warp each). */
struct NbShadowOut { f3 o, d, c; float mint, maxt, cand; };

template <bool MIS>
__device__ __forceinline__ bool nb_shade_path(const NbDeviceScene &sc, const NbWave &w, const NbPlan &plan,
                                              uint32_t p, bool &emit_shadow, NbShadowOut &so) {
    /* returns true when the path lives on into the next iteration; an ended path leaves its
     * final radiance in st0 for K_regen to splat */
    /* every per-path vector is requested up front: one memory round trip instead of a chain of
     * dependent ones (the loads of a path that turns out to be finished are simply unused) */
    float4 s2 = w.st2[p];
    const float4 hit = w.hit[p];
    float4 s0 = w.st0[p];
    float4 s1 = w.st1[p];
    const float4 ro = w.ray_o[p], rd = w.ray_d[p];
    uint4 rs = w.rng[p];
    uint32_t flags = __float_as_uint(s2.w);
    if (flags & (NB_FLAG_EMPTY | NB_FLAG_FINALIZE)) return false;
    const uint32_t slot = __float_as_uint(hit.w);
    f3 Li = mk3(s0.x, s0.y, s0.z);
    const f3 o = mk3(ro.x, ro.y, ro.z), d = mk3(rd.x, rd.y, rd.z);
    f3 thr = mk3(s1.x, s1.y, s1.z);
    if (slot == 0xffffffffu) {              /* miss: Scene::getEnvMap()->eval (path_mis.cpp:35-43, path_mats.cpp:26-34) */
        if (sc.envmap_emitter >= 0) {
            Li += thr * nb_env_radiance(sc, d);
            w.st0[p] = make_float4(Li.x, Li.y, Li.z, s0.w);
        }
        return false;
    }
    float w_mats = s0.w;
    const float4 sl = __ldg(sc.slots + 3 * slot + 2);
    const uint32_t sid = __float_as_uint(sl.y) & NB_SLOT_SHAPE_MASK, prim = __float_as_uint(sl.z);
    const nb_shape sh = sc.shapes[sid];
    NbIts its;
    its.shape = sid;
    if (sh.type == NB_SHAPE_SPHERE) nb_sphere_hit_info(sc, sh, o, d, hit.x, its);
    else nb_mesh_hit_info(sc, sh, prim, hit.y, hit.z, its);
    const float n_lights = (float) sc.n_emitters;

    /* emitter hit (path_mis.cpp:51-56) with the probe's MIS weight (path_mis.cpp:117-140) */
    if (sh.emitter >= 0) {
        const nb_emitter em = sc.emitters[sh.emitter];
        const f3 wi_e = normalize(its.p - o);                       /* EmitterQueryRecord(ref, p, n) */
        if (MIS && !(flags & NB_FLAG_PRIMARY) && (flags & NB_FLAG_BSDF_OK)) {
            const float pdfmat = s1.w;
            const float pdfmat_ems = nb_hit_pdf<true>(sc, sh, o, its.p, its.sh.n, wi_e) / n_lights;
            if (pdfmat + pdfmat_ems > NB_EPS) w_mats = pdfmat / (pdfmat + pdfmat_ems);
        }
        if (MIS && (flags & NB_FLAG_DISCRETE)) w_mats = 1.f;
        if (dot(its.sh.n, -wi_e) >= 0.f || em.type == NB_EMITTER_VOLUME) {   /* AreaEmitter::eval / VolumeEmitter::eval */
            const f3 Le = mk3(em.radiance[0], em.radiance[1], em.radiance[2]);
            Li += (MIS ? w_mats : 1.f) * thr * Le;
        }
    } else if (MIS && (flags & NB_FLAG_DISCRETE)) w_mats = 1.f;

    const uint32_t bounce = flags >> NB_BOUNCE_SHIFT;
    NbRng rng;
    rng.state = ((uint64_t) rs.y << 32) | rs.x;
    rng.inc = ((uint64_t) rs.z << 1) | 1u;

    /* Russian roulette: path_mis.cpp:58-71 (always, clamped) / path_mats.cpp:49-61 (after 3) */
    float succ = fminf(max3(thr), 0.99f);
    bool die = false;
    if (MIS) {
        succ = fmaxf(succ, NB_EPS);
        if (rng.next_float() > succ) die = true; else thr = thr / succ;
    } else if (bounce >= 3) {
        if (rng.next_float() > succ) die = true; else thr = thr / succ;
    }
    if (plan.max_depth && bounce >= plan.max_depth) die = true;
    if (die || sh.bsdf < 0) {
        if (sh.emitter >= 0) w.st0[p] = make_float4(Li.x, Li.y, Li.z, w_mats);
        return false;
    }

    nb_bsdf bsdf = sc.bsdfs[sh.bsdf];
    nb_bsdf_resolve(sc, bsdf, its.uv);
    const f3 wi = its.sh.to_local(-d);
    float w_ems = s2.z;

    if (MIS) {                                                       /* path_mis.cpp:75-106 */
        const uint32_t ei = nb_cdf_sample(sc.emitter_cdf, sc.n_emitters, rng.next_float());
        float2 s; s.x = rng.next_float(); s.y = rng.next_float();
        NbEmitterSample es;
        const f3 ems_col = nb_emitter_sample(sc, ei, its.p, s, es);
        if (!is_zero(ems_col, NB_EPS)) {
            const f3 we = its.sh.to_local(es.wi);
            const f3 f = nb_bsdf_eval<true>(bsdf, wi, we, sc.params);
            const f3 C = thr * (ems_col * we.z * f * n_lights);
            if (!is_black(C)) {
                const float pdf_b = nb_bsdf_pdf<true>(bsdf, wi, we), pdf_e = es.pdf / n_lights;
                emit_shadow = true;
                so.o = es.ray_o; so.d = es.ray_d; so.mint = es.ray_mint; so.maxt = es.ray_maxt;
                so.c = C;
                so.cand = (pdf_b + pdf_e > NB_EPS) ? pdf_e / (pdf_b + pdf_e) : -1.f;
            }
        }
    }

    /* BSDF sampling (path_mis.cpp:108-146 / path_mats.cpp:63-74) */
    float2 s; s.x = rng.next_float(); s.y = rng.next_float();
    f3 wo; bool discrete; float pdfmat;
    const f3 col = nb_bsdf_sample<true>(bsdf, wi, s, wo, discrete, pdfmat, sc.params);
    uint32_t nf = (bounce + 1) << NB_BOUNCE_SHIFT;
    if (discrete) { nf |= NB_FLAG_DISCRETE; w_ems = 0.f; }
    if (!is_zero(col, NB_EPS)) nf |= NB_FLAG_BSDF_OK;
    thr = thr * col;
    const f3 nd = its.sh.to_world(wo);
    float maxt = INFINITY;
    if (is_black(thr)) { nf |= NB_FLAG_FINALIZE; maxt = -1.f; }     /* nothing further can contribute */
    w.ray_o[p] = make_float4(its.p.x, its.p.y, its.p.z, NB_EPS);
    w.ray_d[p] = make_float4(nd.x, nd.y, nd.z, maxt);
    w.st0[p] = make_float4(Li.x, Li.y, Li.z, w_mats);
    w.st1[p] = make_float4(thr.x, thr.y, thr.z, pdfmat);
    w.st2[p] = make_float4(s2.x, s2.y, w_ems, __uint_as_float(nf));
    rs.x = (uint32_t) rng.state; rs.y = (uint32_t) (rng.state >> 32);
    w.rng[p] = rs;
    return true;
}

#ifndef NB_SHADE_MIN_BLOCKS
#define NB_SHADE_MIN_BLOCKS 3
#endif
template <bool MIS>
__global__ void __launch_bounds__(256, NB_SHADE_MIN_BLOCKS)
nb_k_shade(NbDeviceScene sc, NbWave w, NbPlan plan) {
    const uint32_t n = w.counts[NB_CNT_ACTIVE];
    const uint32_t lane = threadIdx.x & 31u, lt_mask = (1u << lane) - 1u;
    for (uint32_t base = blockIdx.x * blockDim.x; base < n; base += gridDim.x * blockDim.x) {
        const uint32_t i = base + threadIdx.x;
        const bool valid = i < n;
        const uint32_t p = valid ? w.queue_cur[i] : 0u;
        bool alive = false, emit_shadow = false;
        NbShadowOut so;
        if (valid) alive = nb_shade_path<MIS>(sc, w, plan, p, emit_shadow, so);
        /* warp-ballot compaction into three queues; lanes 0..2 each issue one aggregated atomic
         * (LIVE, SHADOW, DEAD are consecutive counter words) so the three latencies overlap */
        const uint32_t amask = __ballot_sync(0xffffffffu, alive);
        const uint32_t smask = __ballot_sync(0xffffffffu, emit_shadow);
        const uint32_t dmask = __ballot_sync(0xffffffffu, valid && !alive);
        const uint32_t mine = lane == 0 ? amask : (lane == 1 ? smask : (lane == 2 ? dmask : 0u));
        uint32_t first = 0;
        if (mine) first = atomicAdd(&w.counts[lane], (uint32_t) __popc(mine));
        const uint32_t first_live = __shfl_sync(0xffffffffu, first, NB_CNT_LIVE);
        const uint32_t first_shadow = __shfl_sync(0xffffffffu, first, NB_CNT_SHADOW);
        const uint32_t first_dead = __shfl_sync(0xffffffffu, first, NB_CNT_DEAD);
        if (alive) w.queue_next[first_live + __popc(amask & lt_mask)] = p;
        else if (valid) w.queue_dead[first_dead + __popc(dmask & lt_mask)] = p;
        if (emit_shadow) {
            const uint32_t j = first_shadow + __popc(smask & lt_mask);
            w.sh_o[j] = make_float4(so.o.x, so.o.y, so.o.z, so.mint);
            w.sh_d[j] = make_float4(so.d.x, so.d.y, so.d.z, so.maxt);
            w.sh_c[j] = make_float4(so.c.x, so.c.y, so.c.z, __uint_as_float(p));
            w.sh_w[j] = so.cand;
        }
    }
}

/* ---- K_regen ---------------------------------------------------------------------------------
 * One thread per ended path: ImageBlock::put of its radiance, then the slot starts the next
 * camera sample (renderBlock prologue + sampleRay). Sample numbers and live-queue positions
 * follow from the thread's index in the dead queue — no atomics besides the film's. */
__global__ void __launch_bounds__(256)
nb_k_regen(NbDeviceScene sc, NbWave w, NbFilm film, NbPlan plan) {
    const uint32_t n = w.counts[NB_CNT_DEAD], live = w.counts[NB_CNT_LIVE];
    const unsigned long long sbase = ((unsigned long long) w.counts[NB_CNT_BASE_HI] << 32) | w.counts[NB_CNT_BASE_LO];
    unsigned long long started = 0;
    for (uint32_t j = blockIdx.x * blockDim.x + threadIdx.x; j < n; j += gridDim.x * blockDim.x) {
        const uint32_t p = w.queue_dead[j];
        const float4 s2 = w.st2[p];
        const float4 s0 = w.st0[p];
        if (!(__float_as_uint(s2.w) & NB_FLAG_EMPTY)) {
            nb_film_put(film, s2.x, s2.y, mk3(s0.x, s0.y, s0.z), w.stats);
        }
        const unsigned long long nidx = sbase + j;
        if (nidx < plan.total) {
            nb_generate_path(sc, w, plan, p, nidx);
            w.queue_next[live + j] = p;
            ++started;
        } else {
            w.st2[p].w = __uint_as_float(NB_FLAG_EMPTY);
        }
    }
    for (int off = 16; off; off >>= 1) started += __shfl_xor_sync(0xffffffffu, started, off);
    if ((threadIdx.x & 31) == 0 && started) atomicAdd(&w.stats[NB_STAT_SAMPLES], started);
}

__global__ void nb_k_pool_init(NbWave w, uint32_t n_live) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < w.pool) {
        w.st2[i] = make_float4(0.f, 0.f, 0.f, __uint_as_float(NB_FLAG_EMPTY));
        w.ray_o[i] = make_float4(0.f, 0.f, 0.f, 0.f);
        w.ray_d[i] = make_float4(0.f, 0.f, 1.f, -1.f);
    }
    if (i < n_live) w.queue_dead[i] = i;            /* every slot starts as an ended, empty path */
    if (i < NB_CNT_WORDS) { w.counts[i] = (i == NB_CNT_DEAD) ? n_live : 0u; w.counts_next[i] = 0u; }
}

/* ImageBlock::toBitmap + Color4f::divideByFilterWeight (block.cpp:76-82, color.h:113-118) */
__global__ void nb_k_film_finalize(const float4 *film, uint32_t cols, uint32_t border, uint32_t W, uint32_t H, float *rgb) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= W * H) return;
    const uint32_t x = i % W, y = i / W;
    const float4 c = film[(size_t) (y + border) * cols + (x + border)];
    float3 o = make_float3(0.f, 0.f, 0.f);
    if (fabsf(c.w) > NB_EPS) o = make_float3(c.x / c.w, c.y / c.w, c.z / c.w);
    rgb[3 * (size_t) i] = o.x; rgb[3 * (size_t) i + 1] = o.y; rgb[3 * (size_t) i + 2] = o.z;
}

/* ---- host side --------------------------------------------------------------------------------- */
struct NbRenderState {
    nb_render_params params;
    NbFilm film;
    NbWave wave;                  /* pointers; queue/count ping-pong resolved per launch */
    uint32_t *queue[2] = { nullptr, nullptr };
    uint32_t *counts = nullptr;   /* 2 sets x NB_CNT_WORDS */
    unsigned long long *stats = nullptr;
    uint32_t *tile_counter = nullptr;   /* work-item counter of K_tile (nb_mega.cu) */
    uint32_t *h_counts = nullptr; /* pinned mirror of counts */
    unsigned long long *h_stats = nullptr;
    std::vector<void *> allocs;
    cudaEvent_t ev_begin = nullptr, ev_end = nullptr, ev_batch[2] = { nullptr, nullptr };
    nb_counters totals;
    int grid = 0;
    int pending = 0;              /* K_tile subframes queued since the last nb_render_finish */
    std::vector<cudaEvent_t> prof_events;   /* "profile" option: 4 per iteration (shade | regen | trace) */
};

int nb_render_finish(nb_ctx *ctx) {
    NbRenderState *r = ctx->render;
    if (!r || !r->pending) return NB_OK;
    const int n = r->pending;
    r->pending = 0;
    cudaError_t e = cudaStreamSynchronize(ctx->stream);
    if (e == cudaSuccess) e = cudaGetLastError();
    if (e != cudaSuccess) return nb_fail(ctx, NB_E_CUDA, "nb_render_subframe: %s", cudaGetErrorString(e));
    float ms = 0.f;
    cudaEventElapsedTime(&ms, r->ev_begin, r->ev_end);
    r->totals.device_ms += ms;
    r->totals.iterations += n;
    r->totals.kernel_launches += n;
    r->totals.samples = r->h_stats[NB_STAT_SAMPLES];
    r->totals.rays_closest = r->h_stats[NB_STAT_RAYS_CLOSEST];
    r->totals.rays_any = r->h_stats[NB_STAT_RAYS_ANY];
    r->totals.invalid_samples = r->h_stats[NB_STAT_INVALID];
    return NB_OK;
}

int nb_render_film(nb_ctx *ctx, float **data, float **normal, uint64_t *n_floats) {
    if (!ctx->render || !ctx->render_begun) return nb_fail(ctx, NB_E_STATE, "no render in progress");
    *data = (float *) ctx->render->film.data; *normal = (float *) ctx->render->film.normal;
    *n_floats = (uint64_t) ctx->render->film.rows * ctx->render->film.cols * 4;
    return NB_OK;
}

int nb_render_free(nb_ctx *ctx) {
    NbRenderState *r = ctx->render;
    if (!r) return NB_OK;
    if (r->pending) { cudaStreamSynchronize(ctx->stream); r->pending = 0; }
    for (void *p : r->allocs) cudaFree(p);
    if (r->h_counts) cudaFreeHost(r->h_counts);
    if (r->h_stats) cudaFreeHost(r->h_stats);
    if (r->ev_begin) cudaEventDestroy(r->ev_begin);
    if (r->ev_end) cudaEventDestroy(r->ev_end);
    for (int i = 0; i < 2; ++i) if (r->ev_batch[i]) cudaEventDestroy(r->ev_batch[i]);
    for (cudaEvent_t e : r->prof_events) cudaEventDestroy(e);
    delete r;
    ctx->render = nullptr;
    return NB_OK;
}

template <typename T>
static int nb_dev_alloc(nb_ctx *ctx, NbRenderState *r, T **out, size_t count) {
    void *p = nullptr;
    cudaError_t e = cudaMalloc(&p, count * sizeof(T));
    if (e != cudaSuccess) return nb_fail(ctx, NB_E_NOMEM, "cudaMalloc(%zu bytes): %s", count * sizeof(T), cudaGetErrorString(e));
    r->allocs.push_back(p);
    *out = (T *) p;
    return NB_OK;
}

/* Which execution strategy renders the path integrators: the wavefront pipeline keeps thousands
 * of rays per SM in flight to cover L2/HBM latency of a large BVH; K_tile keeps everything in
 * registers and wins while the scene is cache-resident. */
static bool nb_use_tile_strategy(const nb_ctx *ctx) {
    if (ctx->opt_strategy == NB_STRATEGY_TILE) return true;
    if (ctx->opt_strategy == NB_STRATEGY_WAVEFRONT) return false;
    /* measured on B200: K_tile / K_tile_walk is ahead of the wavefront pipeline on every scene tried —
     * 2.1x on the Cornell box (2 KB of scene), 2.7x on table (2.5 MB), 2.0x on C3 (92 MB: 445 vs 216 M
     * samples/s) and 2.0x on C5 (102 MB, beyond L2 with the film: 196 vs 97 M samples/s) — so it is
     * the default; the wavefront pipeline stays selectable (nb_set_option "strategy") */
    (void) ctx;
    return true;
}

extern "C" {

int nb_render_begin(nb_ctx *ctx, const nb_render_params *params) {
    if (!ctx || !params) return nb_fail(ctx, NB_E_INVALID, "nb_render_begin: NULL argument");
    if (!ctx->bvh_ready) return nb_fail(ctx, NB_E_STATE, "nb_render_begin: scene/BVH not ready");
    switch (params->integrator) {
        case NB_INTEGRATOR_PATH_MIS: case NB_INTEGRATOR_PATH_MATS: case NB_INTEGRATOR_DIRECT:
        case NB_INTEGRATOR_DIRECT_EMS: case NB_INTEGRATOR_DIRECT_MATS: case NB_INTEGRATOR_DIRECT_MIS:
        case NB_INTEGRATOR_PATH_VOL_MATS: case NB_INTEGRATOR_PATH_VOL_MIS:
        case NB_INTEGRATOR_NORMALS: case NB_INTEGRATOR_AV: case NB_INTEGRATOR_PREVIEW: break;
        default: return nb_fail(ctx, NB_E_UNSUPPORTED, "integrator tag %u has no kernel in this backend", params->integrator);
    }
    if ((params->integrator == NB_INTEGRATOR_PATH_MIS || params->integrator == NB_INTEGRATOR_DIRECT_MIS ||
         params->integrator == NB_INTEGRATOR_PATH_VOL_MIS) && ctx->dev.n_emitters == 0)
        return nb_fail(ctx, NB_E_INVALID, "No Emitter in scene!");                 /* path_mis.cpp:76-79 */
    if (!(params->filter_radius > 0.f)) return nb_fail(ctx, NB_E_INVALID, "filter radius must be positive");
    const nb_camera &cam = ctx->dev.camera;
    if (cam.width == 0 || cam.height == 0) return nb_fail(ctx, NB_E_INVALID, "camera has no output size");
    NB_CUDA(ctx, cudaSetDevice(ctx->device));
    { const int rc0 = nb_render_finish(ctx); if (rc0 != NB_OK) return rc0; }
    uint32_t pool = params->pool_paths ? params->pool_paths : (1u << 20);
    pool = (pool + 255u) & ~255u;
    const uint32_t border = (uint32_t) ceilf(params->filter_radius - 0.5f);      /* block.cpp:57 */
    const bool want_normal = (params->layers & NB_LAYER_NORMAL) != 0;
    if (ctx->render && ctx->render->wave.pool == pool && ctx->render->film.rows == cam.height + 2 * border &&
        ctx->render->film.cols == cam.width + 2 * border && (ctx->render->film.normal != nullptr) == want_normal) {
        /* same geometry as the previous render: keep the device buffers, start a fresh film */
        NbRenderState *r = ctx->render;
        r->params = *params;
        r->film.border = border; r->film.radius = params->filter_radius;
        r->film.lookup = NB_FILTER_RESOLUTION / params->filter_radius;
        memcpy(r->film.table, params->filter_table, sizeof(r->film.table));
        ctx->render_begun = true;
        return nb_film_clear(ctx);
    }
    nb_render_free(ctx);
    NbRenderState *r = new NbRenderState();
    ctx->render = r;
    r->params = *params;
    memset(&r->totals, 0, sizeof(r->totals));
    r->film.border = border;
    r->film.rows = cam.height + 2 * r->film.border;
    r->film.cols = cam.width + 2 * r->film.border;
    r->film.radius = params->filter_radius;
    r->film.lookup = NB_FILTER_RESOLUTION / params->filter_radius;               /* block.cpp:65 */
    memcpy(r->film.table, params->filter_table, sizeof(r->film.table));
    int rc;
    if ((rc = nb_dev_alloc(ctx, r, &r->film.data, (size_t) r->film.rows * r->film.cols))) return rc;
    r->film.normal = nullptr;
    if (want_normal) {
        if ((rc = nb_dev_alloc(ctx, r, &r->film.normal, (size_t) r->film.rows * r->film.cols))) return rc;
        NB_CUDA(ctx, cudaMemsetAsync(r->film.normal, 0, (size_t) r->film.rows * r->film.cols * sizeof(float4), ctx->stream));
    }
    NbWave &w = r->wave;
    w.pool = pool;
    if ((rc = nb_dev_alloc(ctx, r, &w.ray_o, pool)) || (rc = nb_dev_alloc(ctx, r, &w.ray_d, pool)) ||
        (rc = nb_dev_alloc(ctx, r, &w.hit, pool)) || (rc = nb_dev_alloc(ctx, r, &w.st0, pool)) ||
        (rc = nb_dev_alloc(ctx, r, &w.st1, pool)) || (rc = nb_dev_alloc(ctx, r, &w.st2, pool)) ||
        (rc = nb_dev_alloc(ctx, r, &w.rng, pool)) || (rc = nb_dev_alloc(ctx, r, &w.sh_o, pool)) ||
        (rc = nb_dev_alloc(ctx, r, &w.sh_d, pool)) || (rc = nb_dev_alloc(ctx, r, &w.sh_c, pool)) ||
        (rc = nb_dev_alloc(ctx, r, &w.sh_w, pool)) || (rc = nb_dev_alloc(ctx, r, &r->queue[0], pool)) ||
        (rc = nb_dev_alloc(ctx, r, &r->queue[1], pool)) || (rc = nb_dev_alloc(ctx, r, &w.queue_dead, pool)) || (rc = nb_dev_alloc(ctx, r, &r->counts, 2 * NB_CNT_WORDS)) ||
        (rc = nb_dev_alloc(ctx, r, &r->stats, NB_STAT_WORDS)) || (rc = nb_dev_alloc(ctx, r, &r->tile_counter, 1)))
        return rc;
    w.stats = r->stats;
    NB_CUDA(ctx, cudaMallocHost((void **) &r->h_counts, 2 * 2 * NB_CNT_WORDS * sizeof(uint32_t)));
    NB_CUDA(ctx, cudaMallocHost((void **) &r->h_stats, NB_STAT_WORDS * sizeof(unsigned long long)));
    NB_CUDA(ctx, cudaEventCreate(&r->ev_begin));
    NB_CUDA(ctx, cudaEventCreate(&r->ev_end));
    NB_CUDA(ctx, cudaEventCreateWithFlags(&r->ev_batch[0], cudaEventDisableTiming));
    NB_CUDA(ctx, cudaEventCreateWithFlags(&r->ev_batch[1], cudaEventDisableTiming));
    NB_CUDA(ctx, cudaMemsetAsync(r->film.data, 0, (size_t) r->film.rows * r->film.cols * sizeof(float4), ctx->stream));
    NB_CUDA(ctx, cudaMemsetAsync(r->stats, 0, NB_STAT_WORDS * sizeof(unsigned long long), ctx->stream));
    NB_CUDA(ctx, cudaMemsetAsync(r->counts, 0, 2 * NB_CNT_WORDS * sizeof(uint32_t), ctx->stream));
    r->grid = ctx->sm_count * 8;
    NB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    ctx->render_begun = true;
    return NB_OK;
}

int nb_sample_slice(uint32_t first, uint32_t count, uint32_t stride, uint32_t offset, uint32_t *first_sample, uint32_t *n_samples) {
    if (stride == 0 || offset >= stride) return NB_E_INVALID;
    const uint64_t s0 = (uint64_t) first + ((offset + stride - first % stride) % stride);
    const uint64_t end = (uint64_t) first + count;
    if (first_sample) *first_sample = (uint32_t) s0;
    if (n_samples) *n_samples = s0 < end ? (uint32_t) ((end - s0 + stride - 1) / stride) : 0u;
    return NB_OK;
}

int nb_film_clear(nb_ctx *ctx) {
    if (!ctx || !ctx->render || !ctx->render_begun) return nb_fail(ctx, NB_E_STATE, "nb_film_clear: no render in progress");
    NbRenderState *r = ctx->render;
    NB_CUDA(ctx, cudaSetDevice(ctx->device));
    { const int rc0 = nb_render_finish(ctx); if (rc0 != NB_OK) return rc0; }
    NB_CUDA(ctx, cudaMemsetAsync(r->film.data, 0, (size_t) r->film.rows * r->film.cols * sizeof(float4), ctx->stream));
    if (r->film.normal) NB_CUDA(ctx, cudaMemsetAsync(r->film.normal, 0, (size_t) r->film.rows * r->film.cols * sizeof(float4), ctx->stream));
    NB_CUDA(ctx, cudaMemsetAsync(r->stats, 0, NB_STAT_WORDS * sizeof(unsigned long long), ctx->stream));
    memset(&r->totals, 0, sizeof(r->totals));
    return NB_OK;
}

/* The wavefront loop. Iteration = K_shade (consumes hits, regenerates, compacts) then K_trace
 * (main rays + the shadow rays K_shade just made). The host never reads a queue length to size
 * a launch; it only polls "anything left?" one batch behind the GPU. */
int nb_render_subframe(nb_ctx *ctx, uint32_t first, uint32_t count, uint32_t stride, uint32_t offset) {
    if (!ctx || !ctx->render || !ctx->render_begun) return nb_fail(ctx, NB_E_STATE, "nb_render_subframe: call nb_render_begin first");
    if (stride == 0 || offset >= stride) return nb_fail(ctx, NB_E_INVALID, "nb_render_subframe: bad sample slice %u/%u", offset, stride);
    NbRenderState *r = ctx->render;
    NB_CUDA(ctx, cudaSetDevice(ctx->device));
    const nb_camera &cam = ctx->dev.camera;
    uint32_t s0 = 0, passes = 0;
    nb_sample_slice(first, count, stride, offset, &s0, &passes);
    NbPlan plan;
    plan.npix = cam.width * cam.height; plan.width = cam.width;
    plan.total = (unsigned long long) passes * plan.npix;
    plan.seed = r->params.seed; plan.first = s0; plan.stride = stride; plan.max_depth = r->params.max_depth;
    for (int i = 0; i < 3; ++i) plan.iparam[i] = r->params.integrator_params[i];
    if (plan.total == 0) return NB_OK;
    const uint32_t integ = r->params.integrator;
    const bool is_path = integ == NB_INTEGRATOR_PATH_MIS || integ == NB_INTEGRATOR_PATH_MATS;
    if (!is_path || nb_use_tile_strategy(ctx)) {
        /* K_tile: one persistent launch, path state in registers (nb_mega.cu). Asynchronous: the launch, its end
         * event and the copy of the device counters are queued on the context's stream; nb_render_finish (called by
         * nb_sync, nb_get_counters, nb_film_download*, nb_film_clear, nb_render_begin) waits and reads them */
        if (!r->pending) cudaEventRecord(r->ev_begin, ctx->stream);
        cudaError_t e = nb_launch_tile(ctx->dev, r->film, plan, cam.height, passes, integ, r->tile_counter, r->stats,
                                       ctx->sm_count, ctx->stream, ctx->opt_walk != 0);
        if (e != cudaSuccess) return nb_fail(ctx, NB_E_CUDA, "nb_render_subframe: K_tile launch: %s", cudaGetErrorString(e));
        cudaEventRecord(r->ev_end, ctx->stream);
        cudaMemcpyAsync(r->h_stats, r->stats, NB_STAT_WORDS * sizeof(unsigned long long), cudaMemcpyDeviceToHost, ctx->stream);
        r->pending += 1;
        return NB_OK;
    }
    { const int rc0 = nb_render_finish(ctx); if (rc0 != NB_OK) return rc0; }
    cudaEventRecord(r->ev_begin, ctx->stream);
    NbWave w = r->wave;
    const uint32_t n_live = (uint32_t) std::min<unsigned long long>(w.pool, plan.total);
    w.total_samples = plan.total;
    w.queue_cur = r->queue[0]; w.queue_next = r->queue[1];
    w.counts = r->counts; w.counts_next = r->counts + NB_CNT_WORDS;
    nb_k_pool_init<<<(w.pool + 255) / 256, 256, 0, ctx->stream>>>(w, n_live);
    const bool mis = r->params.integrator == NB_INTEGRATOR_PATH_MIS;
    const int BATCH = 8;                      /* even: after a batch, "next" inputs sit in set 0 */
    uint64_t iter = 0, launches = 1;
    int pending[2] = { 0, 0 };
    bool done = false;
    const bool prof = ctx->opt_profile != 0;
    size_t prof_used = 0;
    auto prof_mark = [&]() {
        if (!prof) return;
        if (prof_used == r->prof_events.size()) { cudaEvent_t e; cudaEventCreate(&e); r->prof_events.push_back(e); }
        cudaEventRecord(r->prof_events[prof_used++], ctx->stream);
    };
    for (int batch = 0; !done; ++batch) {
        for (int k = 0; k < BATCH; ++k, ++iter) {
            const int a = (int) (iter & 1), b = a ^ 1;
            NbWave ws = w;
            ws.queue_cur = r->queue[a]; ws.queue_next = r->queue[b];
            ws.counts = r->counts + a * NB_CNT_WORDS; ws.counts_next = r->counts + b * NB_CNT_WORDS;
            prof_mark();
            if (iter > 0) {                   /* the first iteration only has empty slots to fill */
                if (mis) nb_k_shade<true><<<r->grid, 256, 0, ctx->stream>>>(ctx->dev, ws, plan);
                else nb_k_shade<false><<<r->grid, 256, 0, ctx->stream>>>(ctx->dev, ws, plan);
                ++launches;
            }
            prof_mark();
            nb_k_regen<<<r->grid, 256, 0, ctx->stream>>>(ctx->dev, ws, r->film, plan);
            prof_mark();
            nb_launch_trace_wave(ctx->dev, ws, r->grid, ctx->stream);
            prof_mark();
            launches += 2;
        }
        const int slot = batch & 1;
        cudaMemcpyAsync(r->h_counts + slot * 2 * NB_CNT_WORDS, r->counts, 2 * NB_CNT_WORDS * sizeof(uint32_t),
                        cudaMemcpyDeviceToHost, ctx->stream);
        cudaEventRecord(r->ev_batch[slot], ctx->stream);
        pending[slot] = 1;
        const int prev = slot ^ 1;
        if (pending[prev]) {
            NB_CUDA(ctx, cudaEventSynchronize(r->ev_batch[prev]));
            pending[prev] = 0;
            if (r->h_counts[prev * 2 * NB_CNT_WORDS + NB_CNT_ACTIVE] == 0) done = true;
        }
        if (!done && batch > (1 << 22)) return nb_fail(ctx, NB_E_CUDA, "nb_render_subframe: wavefront did not drain");
    }
    cudaEventRecord(r->ev_end, ctx->stream);
    cudaMemcpyAsync(r->h_stats, r->stats, NB_STAT_WORDS * sizeof(unsigned long long), cudaMemcpyDeviceToHost, ctx->stream);
    cudaError_t e = cudaStreamSynchronize(ctx->stream);
    if (e == cudaSuccess) e = cudaGetLastError();
    if (e != cudaSuccess) return nb_fail(ctx, NB_E_CUDA, "nb_render_subframe: %s", cudaGetErrorString(e));
    /* the last polled batch may be stale by one: verify the queue really drained */
    NB_CUDA(ctx, cudaMemcpy(r->h_counts, r->counts, 2 * NB_CNT_WORDS * sizeof(uint32_t), cudaMemcpyDeviceToHost));
    if (r->h_counts[NB_CNT_ACTIVE] != 0) return nb_fail(ctx, NB_E_CUDA, "nb_render_subframe: internal: live paths left");
    float ms = 0.f;
    cudaEventElapsedTime(&ms, r->ev_begin, r->ev_end);
    for (size_t k = 0; prof && k + 3 < prof_used; k += 4) {
        float a = 0.f, b = 0.f, c = 0.f;
        cudaEventElapsedTime(&a, r->prof_events[k], r->prof_events[k + 1]);
        cudaEventElapsedTime(&b, r->prof_events[k + 1], r->prof_events[k + 2]);
        cudaEventElapsedTime(&c, r->prof_events[k + 2], r->prof_events[k + 3]);
        r->totals.shade_ms += a; r->totals.regen_ms += b; r->totals.trace_ms += c;
    }
    r->totals.device_ms += ms;
    r->totals.iterations += iter;
    r->totals.kernel_launches += launches;
    r->totals.samples = r->h_stats[NB_STAT_SAMPLES];
    r->totals.rays_closest = r->h_stats[NB_STAT_RAYS_CLOSEST];
    r->totals.rays_any = r->h_stats[NB_STAT_RAYS_ANY];
    r->totals.invalid_samples = r->h_stats[NB_STAT_INVALID];
    return NB_OK;
}

int nb_sync(nb_ctx *ctx) {
    if (!ctx) return NB_E_INVALID;
    NB_CUDA(ctx, cudaSetDevice(ctx->device));
    { const int rc0 = nb_render_finish(ctx); if (rc0 != NB_OK) return rc0; }
    NB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return NB_OK;
}

int nb_film_shape(nb_ctx *ctx, uint32_t *rows, uint32_t *cols, uint32_t *border) {
    if (!ctx || !ctx->render) return nb_fail(ctx, NB_E_STATE, "nb_film_shape: no render in progress");
    if (rows) *rows = ctx->render->film.rows;
    if (cols) *cols = ctx->render->film.cols;
    if (border) *border = ctx->render->film.border;
    return NB_OK;
}

int nb_film_download(nb_ctx *ctx, float *out, int normalize) { return nb_film_download_layer(ctx, NB_LAYER_COMPOSITE, out, normalize); }

int nb_film_download_layer(nb_ctx *ctx, int layer, float *out, int normalize) {
    if (!ctx || !ctx->render || !out) return nb_fail(ctx, NB_E_STATE, "nb_film_download: no render in progress or NULL buffer");
    NbRenderState *r = ctx->render;
    if (layer != NB_LAYER_COMPOSITE && layer != NB_LAYER_NORMAL) return nb_fail(ctx, NB_E_INVALID, "nb_film_download_layer: layer %d unknown", layer);
    float4 *const film_data = layer == NB_LAYER_NORMAL ? r->film.normal : r->film.data;
    if (!film_data) return nb_fail(ctx, NB_E_STATE, "nb_film_download_layer: layer %d was not requested in nb_render_params.layers", layer);
    NB_CUDA(ctx, cudaSetDevice(ctx->device));
    { const int rc0 = nb_render_finish(ctx); if (rc0 != NB_OK) return rc0; }
    if (!normalize) {
        NB_CUDA(ctx, cudaMemcpyAsync(out, film_data, (size_t) r->film.rows * r->film.cols * sizeof(float4), cudaMemcpyDeviceToHost, ctx->stream));
        NB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        return NB_OK;
    }
    const nb_camera &cam = ctx->dev.camera;
    const size_t n = (size_t) cam.width * cam.height;
    float *d_rgb = nullptr;
    NB_CUDA(ctx, cudaMalloc((void **) &d_rgb, n * 3 * sizeof(float)));
    nb_k_film_finalize<<<(unsigned) ((n + 255) / 256), 256, 0, ctx->stream>>>(film_data, r->film.cols, r->film.border, cam.width, cam.height, d_rgb);
    cudaMemcpyAsync(out, d_rgb, n * 3 * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream);
    cudaError_t e = cudaStreamSynchronize(ctx->stream);
    cudaFree(d_rgb);
    if (e != cudaSuccess) return nb_fail(ctx, NB_E_CUDA, "nb_film_download: %s", cudaGetErrorString(e));
    return NB_OK;
}

int nb_film_device_ptr(nb_ctx *ctx, void **ptr, uint64_t *n_floats) {
    if (!ctx || !ctx->render) return nb_fail(ctx, NB_E_STATE, "nb_film_device_ptr: no render in progress");
    if (ptr) *ptr = ctx->render->film.data;
    if (n_floats) *n_floats = (uint64_t) ctx->render->film.rows * ctx->render->film.cols * 4;
    return NB_OK;
}

int nb_get_counters(nb_ctx *ctx, nb_counters *out) {
    if (!ctx || !ctx->render || !out) return nb_fail(ctx, NB_E_STATE, "nb_get_counters: no render in progress");
    if (ctx->render->pending) {
        NB_CUDA(ctx, cudaSetDevice(ctx->device));
        const int rc0 = nb_render_finish(ctx);
        if (rc0 != NB_OK) return rc0;
    }
    *out = ctx->render->totals;
    return NB_OK;
}

int nb_set_option(nb_ctx *ctx, const char *name, int64_t value) {
    if (!ctx || !name) return NB_E_INVALID;
    if (!strcmp(name, "profile")) ctx->opt_profile = (int) value;
    else if (!strcmp(name, "sort")) ctx->opt_sort = (int) value;
    else if (!strcmp(name, "walk")) ctx->opt_walk = (int) value;
    else if (!strcmp(name, "wide")) { ctx->opt_wide = value != 0; nb_update_wide(ctx); }
    else if (!strcmp(name, "trace_wide")) ctx->opt_trace_wide = value != 0;
    else if (!strcmp(name, "trace_fast")) ctx->opt_trace_fast = value != 0;
    else if (!strcmp(name, "ordered")) { ctx->opt_ordered = value != 0; ctx->dev.ordered = value != 0 ? 1u : 0u; }
    else if (!strcmp(name, "flat_scan")) {
        ctx->opt_flat = (int) value;
        ctx->dev.flat = nb_scene_flat(ctx) ? 1u : 0u;
    }
    else if (!strcmp(name, "strategy")) {
        if (value < NB_STRATEGY_AUTO || value > NB_STRATEGY_TILE) return nb_fail(ctx, NB_E_INVALID, "nb_set_option: strategy %lld out of range", (long long) value);
        ctx->opt_strategy = (int) value;
    }
    else return nb_fail(ctx, NB_E_INVALID, "nb_set_option: unknown option '%s'", name);
    return NB_OK;
}

} /* extern "C" */
