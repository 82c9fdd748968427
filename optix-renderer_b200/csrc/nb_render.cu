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
    /* adaptive rounds (nb_render_adaptive_round), allocated on first use */
    float *adp_old_var = nullptr, *adp_old_norm = nullptr;     /* m_oldVariance (per pixel), m_oldNorm (per block) of every block's sampler */
    uint32_t *adp_total = nullptr;                             /* per block: passes rendered since nb_render_begin */
    uint32_t *adp_count = nullptr;                             /* per block: `count` of render.cpp:326, NB_ADP_DONE once its loop ended */
    uint32_t *adp_list = nullptr, *adp_ctl = nullptr;          /* sample list; [0] length [1] first sample index [2] blocks in this pass */
    unsigned long long *adp_acc = nullptr, *adp_h_acc = nullptr;   /* round totals: samples, block passes, passes, length of the last pass */
    uint32_t adp_block = 0, adp_blocks = 0, adp_serial = 0;
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
    if (r->adp_h_acc) cudaFreeHost(r->adp_h_acc);
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
        if (r->adp_total) NB_CUDA(ctx, cudaMemsetAsync(r->adp_total, 0, r->adp_blocks * sizeof(uint32_t), ctx->stream));   /* nb_adaptive_block_passes counts per render */
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
    r->film.pass = nullptr; r->film.pass_block = 0;
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

/* ---- adaptive rounds: AdaptiveSampler (src/samplers/adaptive.cpp) as a device feedback loop --------------------------------
 * The reference renders an adaptive image block by block (NORI_BLOCK_SIZE_ADAPTIVE = 4, block.h:31): for every block, in
 * every round of the outer sample loop, `while (sampler->computeVariance(block)) { renderBlock(..); m_block.put(block); }`
 * (render.cpp:323-333) — each pass draws width x height camera samples, the first ones one per pixel, the later ones placed by
 * a discrete pdf made from the 3x3 luminance variance of the block's PREVIOUS pass (common.cpp:339-399), until the change of
 * that variance map stops shrinking. Here all blocks advance together: one warp per block takes the decision the block's
 * sampler would take and appends the block's samples to a list, the persistent render kernel consumes the list
 * (NbTilePlan::list), its puts also fill the blocks' own images (NbFilm::pass) for the next decision. The host only reads back
 * "how many blocks went on". Kept from the reference, because they shape what it computes: the rounds computeVariance sees
 * (0, 0, 1, 2, .. — render.cpp:327, :334 post-increment), the reset at round 0, the cap at round 1000, the [1, 1.254] rescale,
 * the Frobenius normalisation of the stored map, the stop rule |sum(old - new)| > previous, and the transposed placement
 * (getSampleIndices returns {row, col}, renderBlock reads it as {x, y}: adaptive.cpp:166-168, render.cpp:438-440).
 * A block whose variance map is constant (black, a flat emitter, a dim corner with no lucky sample) is not special either:
 * common.cpp:381-386 makes the map zero, Eigen's normalize() leaves a zero matrix alone, DiscretePDF::normalize() leaves a zero
 * cdf alone, and DiscretePDF::sample() then answers "the last entry" for every draw — so the reference renders one pass with
 * all of the block's samples on its last pixel, and the NEXT decision stops the block (the sum of any non-zero map exceeds the
 * stored 0) unless that pass was black too, in which case it goes round again, up to the cap: a black block costs the
 * reference 1000 passes per round. All of that is reproduced (tests/test_adaptive_gpu.py compares the law of the passes per
 * block with the reference's, black blocks included); `round_cap` lets a caller who does not want to pay for it lower the cap.
 * Partial blocks at the image edge place samples inside the block (the reference's transposed indices leave it). */
#define NB_ADP_DONE 0xffffffffu
#define NB_ADP_MAX_BLOCK 8u
__global__ void __launch_bounds__(256) nb_k_adaptive_decide(NbFilm film, uint32_t W, uint32_t H, uint32_t bs, uint32_t nbx, uint32_t n_blocks,
                                                            uint32_t initial_uniform, uint32_t round_cap, float *old_var, float *old_norm, uint32_t *count, uint32_t *total,
                                                            uint32_t *list, uint32_t *ctl, unsigned long long seed, uint32_t serial) {
    __shared__ float sm[8][2][NB_ADP_MAX_BLOCK * NB_ADP_MAX_BLOCK];
    const uint32_t block = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31u;
    if (block >= n_blocks) return;
    const uint32_t cnt = count[block];
    if (cnt == NB_ADP_DONE) return;
    const uint32_t round = cnt == 0 ? 0u : cnt - 1u;           /* m_sampleRound at this call of computeVariance */
    const uint32_t x0 = (block % nbx) * bs, y0 = (block / nbx) * bs, w = min(bs, W - x0), h = min(bs, H - y0), n = w * h;
    float *lum = sm[threadIdx.x >> 5][0], *cdf = sm[threadIdx.x >> 5][1];
    auto pixel = [&](uint32_t p) { return (size_t) (y0 + p / w + film.border) * film.cols + (x0 + p % w + film.border); };
    auto own = [&](uint32_t p) { return (size_t) (y0 + p / w) * W + (x0 + p % w); };          /* old_var is W x H */
    auto sum = [&](float v) { for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o); return v; };
    bool go = true, uniform = true, zero_pdf = false;
    float onorm = old_norm[block];
    if (round == 0) {                                          /* adaptive.cpp:72-78 */
        for (uint32_t p = lane; p < n; p += 32) old_var[own(p)] = 0.f;
        onorm = 10000.f;
    } else if (round >= round_cap) go = false;                 /* adaptive.cpp:79-82: 1000 */
    if (go && round >= initial_uniform) {
        uniform = false;
        for (uint32_t p = lane; p < n; p += 32) {              /* |luminance| of the block's own image, Color4f::divideByFilterWeight */
            const float4 c = film.pass[pixel(p)];
            lum[p] = c.w != 0.f ? fabsf((c.x / c.w) * 0.212671f + (c.y / c.w) * 0.715160f + (c.z / c.w) * 0.072169f) : 0.f;
        }
        __syncwarp();
        float v[2] = { 0.f, 0.f }, lo = INFINITY, hi = -INFINITY;
        for (uint32_t p = lane, q = 0; p < n; p += 32, ++q) {  /* computeVarianceFromImage, common.cpp:339-378: neighbours inside the block */
            const int i = (int) (p / w), j = (int) (p % w);
            float mean = 0.f, cntf = 0.f;
            for (int k = -1; k <= 1; ++k)
                for (int l = -1; l <= 1; ++l)
                    if (i + k >= 0 && i + k < (int) h && j + l >= 0 && j + l < (int) w) { mean += lum[(i + k) * w + j + l]; cntf += 1.f; }
            mean /= cntf;
            float col = 0.f;
            for (int k = -1; k <= 1; ++k)
                for (int l = -1; l <= 1; ++l)
                    if (i + k >= 0 && i + k < (int) h && j + l >= 0 && j + l < (int) w) {
                        const float d = lum[(i + k) * w + j + l] - mean;
                        col += 1.f / cntf * (d * d);
                    }
            v[q] = col; lo = fminf(lo, col); hi = fmaxf(hi, col);
        }
        for (int o = 16; o; o >>= 1) { lo = fminf(lo, __shfl_xor_sync(0xffffffffu, lo, o)); hi = fmaxf(hi, __shfl_xor_sync(0xffffffffu, hi, o)); }
        const bool flat = hi - lo < NB_EPS;                    /* common.cpp:381-386: the map becomes zero */
        float diff = 0.f, sq = 0.f, tot = 0.f;
        for (uint32_t p = lane, q = 0; p < n; p += 32, ++q) {
            v[q] = flat ? 0.f : 1.f + (v[q] - lo) / (hi - lo) * 0.254f;
            diff += old_var[own(p)] - v[q]; sq += v[q] * v[q]; tot += v[q];
        }
        const float var_diff = fabsf(sum(diff));               /* adaptive.cpp:98 */
        sq = sum(sq); tot = sum(tot);
        if (var_diff > onorm) go = false;                      /* adaptive.cpp:127-128 */
        else {
            onorm = var_diff;
            const float inv = sq > 0.f ? 1.f / sqrtf(sq) : 1.f; /* variance.normalize(): the unit-norm map is what is kept; a zero map stays */
            zero_pdf = !(tot > 0.f);                           /* DiscretePDF::normalize() leaves a zero cdf alone */
            for (uint32_t p = lane, q = 0; p < n; p += 32, ++q) { old_var[own(p)] = v[q] * inv; lum[p] = zero_pdf ? 0.f : v[q] / tot; }
            __syncwarp();
            if (lane == 0) { float c = 0.f; for (uint32_t p = 0; p + 1 < n; ++p) { c += lum[p]; cdf[p + 1] = c; } }
            __syncwarp();                                      /* cdf[p] = sum of the first p probabilities, 0 < p < n (the bounds 0 and 1 are implied) */
        }
    }
    if (!go) { if (lane == 0) count[block] = NB_ADP_DONE; return; }
    uint32_t at = 0;
    if (lane == 0) {
        old_norm[block] = onorm;
        count[block] = cnt + 1u;                               /* render.cpp:334 */
        total[block] += 1u;
        at = atomicAdd(ctl, n);
        atomicAdd(ctl + 2, 1u);
    }
    at = __shfl_sync(0xffffffffu, at, 0);
    for (uint32_t p = lane; p < n; p += 32) {
        uint32_t x = p % w, y = p / w;                         /* uniform pass: every pixel once (adaptive.cpp:150-160) */
        if (!uniform) {
            NbRng rng; rng.seed(seed ^ 0xada9715eULL, ((uint64_t) block << 32) | ((uint64_t) serial << 6) | p);
            const float u = rng.next_float();                  /* DiscretePDF::sample (dpdf.h): last bound not above u */
            uint32_t e = 0;
            for (uint32_t k = 1; k < n; ++k) e += cdf[k] < u ? 1u : 0u;
            if (zero_pdf) e = u > 0.f ? n - 1u : 0u;           /* every bound of an unnormalised zero cdf lies below u */
            const uint32_t row = e / w, col = e % w;           /* {row, col} read as {x, y} */
            x = min(row, w - 1u); y = min(col, h - 1u);
        }
        list[at + p] = (y0 + y) * W + x0 + x;
    }
    for (uint32_t p = lane; p < n; p += 32) film.pass[pixel(p)] = make_float4(0.f, 0.f, 0.f, 0.f);   /* renderBlock: block.clear() */
}

/* after a pass: its length joins the running totals and the first sample index of the next pass; the list starts empty again */
__global__ void nb_k_adaptive_next(uint32_t *ctl, unsigned long long *acc) {
    const uint32_t n = ctl[0];
    acc[0] += n; acc[1] += ctl[2]; acc[2] += n ? 1u : 0u; acc[3] = n;
    ctl[1] += n; ctl[0] = 0u; ctl[2] = 0u;
}

/* One round of the outer sample loop of render.cpp:283-360 for a scene whose sampler is `adaptive`: every block's
 * while-loop, all blocks in step. Synchronous. */
int nb_render_adaptive_round(nb_ctx *ctx, uint32_t block_size, uint32_t initial_uniform, uint32_t round_cap, nb_adaptive_info *out) {
    if (!ctx || !ctx->render || !ctx->render_begun) return nb_fail(ctx, NB_E_STATE, "nb_render_adaptive_round: call nb_render_begin first");
    if (block_size < 2 || block_size > NB_ADP_MAX_BLOCK) return nb_fail(ctx, NB_E_INVALID, "nb_render_adaptive_round: block size %u outside 2..%u", block_size, NB_ADP_MAX_BLOCK);
    if (initial_uniform < 1) initial_uniform = 1;              /* adaptive.cpp:21: clamp(initialUniform, 1, sampleCount) */
    if (round_cap == 0 || round_cap > 1000) round_cap = 1000;  /* adaptive.cpp:79 */
    NbRenderState *r = ctx->render;
    NB_CUDA(ctx, cudaSetDevice(ctx->device));
    { const int rc0 = nb_render_finish(ctx); if (rc0 != NB_OK) return rc0; }
    const nb_camera &cam = ctx->dev.camera;
    const uint32_t W = cam.width, H = cam.height, npix = W * H;
    const uint32_t nbx = (W + block_size - 1) / block_size, nby = (H + block_size - 1) / block_size, n_blocks = nbx * nby;
    int rc;
    if (!r->adp_list) {
        if ((rc = nb_dev_alloc(ctx, r, &r->film.pass, (size_t) r->film.rows * r->film.cols)) || (rc = nb_dev_alloc(ctx, r, &r->adp_old_var, npix)) ||
            (rc = nb_dev_alloc(ctx, r, &r->adp_list, npix)) || (rc = nb_dev_alloc(ctx, r, &r->adp_ctl, 4)) || (rc = nb_dev_alloc(ctx, r, &r->adp_acc, 4)))
            return rc;
        NB_CUDA(ctx, cudaMallocHost((void **) &r->adp_h_acc, 4 * sizeof(unsigned long long)));
        NB_CUDA(ctx, cudaMemsetAsync(r->film.pass, 0, (size_t) r->film.rows * r->film.cols * sizeof(float4), ctx->stream));
        NB_CUDA(ctx, cudaMemsetAsync(r->adp_ctl, 0, 4 * sizeof(uint32_t), ctx->stream));      /* the sample index runs on from round to round */
    }
    if (r->adp_block != block_size) {                          /* per-block state for this block size */
        if ((rc = nb_dev_alloc(ctx, r, &r->adp_old_norm, n_blocks)) || (rc = nb_dev_alloc(ctx, r, &r->adp_count, n_blocks)) ||
            (rc = nb_dev_alloc(ctx, r, &r->adp_total, n_blocks))) return rc;
        NB_CUDA(ctx, cudaMemsetAsync(r->adp_total, 0, n_blocks * sizeof(uint32_t), ctx->stream));
        r->adp_block = block_size; r->adp_blocks = n_blocks;
    }
    r->film.pass_block = block_size;
    NB_CUDA(ctx, cudaMemsetAsync(r->adp_count, 0, n_blocks * sizeof(uint32_t), ctx->stream));      /* render.cpp:326: count = 0 */
    NB_CUDA(ctx, cudaMemsetAsync(r->adp_acc, 0, 4 * sizeof(unsigned long long), ctx->stream));
    NbPlan plan;
    plan.npix = npix; plan.width = W; plan.total = 0;
    plan.seed = r->params.seed; plan.first = 0; plan.stride = 1; plan.max_depth = r->params.max_depth;
    for (int i = 0; i < 3; ++i) plan.iparam[i] = r->params.integrator_params[i];
    cudaEventRecord(r->ev_begin, ctx->stream);
    /* passes are queued in batches and the host looks at the totals once per batch: a pass past the end of the round finds
     * every block done and costs two empty launches */
    const uint32_t BATCH = 8;
    uint32_t queued = 0;
    for (;;) {
        for (uint32_t k = 0; k < BATCH; ++k, ++queued) {
            nb_k_adaptive_decide<<<(n_blocks + 7) / 8, 256, 0, ctx->stream>>>(r->film, W, H, block_size, nbx, n_blocks, initial_uniform, round_cap, r->adp_old_var,
                                                                            r->adp_old_norm, r->adp_count, r->adp_total, r->adp_list, r->adp_ctl, r->params.seed, r->adp_serial++);
            const cudaError_t e = nb_launch_tile(ctx->dev, r->film, plan, H, 1, r->params.integrator, r->tile_counter, r->stats, ctx->sm_count, ctx->stream,
                                                 ctx->opt_walk != 0, r->adp_list, r->adp_ctl);
            if (e != cudaSuccess) return nb_fail(ctx, NB_E_CUDA, "nb_render_adaptive_round: K_tile launch: %s", cudaGetErrorString(e));
            nb_k_adaptive_next<<<1, 1, 0, ctx->stream>>>(r->adp_ctl, r->adp_acc);
        }
        NB_CUDA(ctx, cudaMemcpyAsync(r->adp_h_acc, r->adp_acc, 4 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, ctx->stream));
        cudaError_t e = cudaStreamSynchronize(ctx->stream);
        if (e == cudaSuccess) e = cudaGetLastError();
        if (e != cudaSuccess) return nb_fail(ctx, NB_E_CUDA, "nb_render_adaptive_round: %s", cudaGetErrorString(e));
        if (r->adp_h_acc[3] == 0) break;                       /* the last pass of the batch drew nothing: every block's loop has ended */
        if (queued > round_cap + 2 + BATCH) return nb_fail(ctx, NB_E_CUDA, "nb_render_adaptive_round: internal: blocks still running after the round cap");
    }
    cudaEventRecord(r->ev_end, ctx->stream);
    NB_CUDA(ctx, cudaMemcpyAsync(r->h_stats, r->stats, NB_STAT_WORDS * sizeof(unsigned long long), cudaMemcpyDeviceToHost, ctx->stream));
    NB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    nb_adaptive_info info; memset(&info, 0, sizeof(info));
    info.samples = r->adp_h_acc[0]; info.block_passes = r->adp_h_acc[1]; info.passes = (uint32_t) r->adp_h_acc[2]; info.blocks = n_blocks;
    float ms = 0.f;
    cudaEventElapsedTime(&ms, r->ev_begin, r->ev_end);
    r->totals.device_ms += ms;
    r->totals.iterations += info.passes;
    r->totals.kernel_launches += 3ull * queued;
    r->totals.samples = r->h_stats[NB_STAT_SAMPLES];
    r->totals.rays_closest = r->h_stats[NB_STAT_RAYS_CLOSEST];
    r->totals.rays_any = r->h_stats[NB_STAT_RAYS_ANY];
    r->totals.invalid_samples = r->h_stats[NB_STAT_INVALID];
    r->film.pass_block = 0;                                    /* plain subframes after this round do not touch the block images */
    if (out) *out = info;
    return NB_OK;
}

int nb_adaptive_block_passes(nb_ctx *ctx, uint32_t *out, uint32_t n) {
    if (!ctx || !ctx->render || !ctx->render->adp_total) return nb_fail(ctx, NB_E_STATE, "nb_adaptive_block_passes: no adaptive round has run");
    const nb_camera &cam = ctx->dev.camera;
    const uint32_t bs = ctx->render->adp_block, n_blocks = ((cam.width + bs - 1) / bs) * ((cam.height + bs - 1) / bs);
    if (n != n_blocks) return nb_fail(ctx, NB_E_INVALID, "nb_adaptive_block_passes: the image has %u blocks of %u x %u, not %u", n_blocks, bs, bs, n);
    NB_CUDA(ctx, cudaSetDevice(ctx->device));
    NB_CUDA(ctx, cudaMemcpy(out, ctx->render->adp_total, n * sizeof(uint32_t), cudaMemcpyDeviceToHost));
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
