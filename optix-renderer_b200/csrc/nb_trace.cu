/* nb_trace.cu — traversal kernels. Built with -fmad=false (see nb_trace.cuh). */
#include "nb_trace.cuh"
#include "nb_wavefront.h"

/* Parity / benchmark entry: one thread per ray, rays as two float4 {o, mint} {d, maxt}. */
template <bool FLAT>
__global__ void __launch_bounds__(256)
nb_k_trace_hits(const float4 *__restrict__ nodes, const float4 *__restrict__ slots, uint32_t n_nodes, uint32_t n_slots,
                const float4 *__restrict__ rays, uint64_t n, int any_hit, bool ordered, bool fast,
                float4 *__restrict__ out, uint32_t *__restrict__ out_any) {
    for (uint64_t i = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t) gridDim.x * blockDim.x) {
        const float4 ro = __ldg(rays + 2 * i), rd = __ldg(rays + 2 * i + 1);
        const float3 o = make_float3(ro.x, ro.y, ro.z), d = make_float3(rd.x, rd.y, rd.z);
        NbHit hit;
        if (FLAT) nb_scan<false>(slots, n_slots, o, d, ro.w, rd.w, hit, fast);
        else nb_query<false, FLAT>(nodes, slots, n_nodes, n_slots, o, d, ro.w, rd.w, hit, ordered);
        out[i] = make_float4(hit.t, hit.u, hit.v, __uint_as_float(hit.slot));
        if (any_hit) {
            NbHit dummy;
            out_any[i] = (FLAT ? nb_scan<true>(slots, n_slots, o, d, ro.w, rd.w, dummy, fast)
                               : nb_query<true, FLAT>(nodes, slots, n_nodes, n_slots, o, d, ro.w, rd.w, dummy, ordered)) ? 1u : 0u;
        }
    }
}

/* the same entry through the production tree (nb_traverse_wide) */
__global__ void __launch_bounds__(256)
nb_k_trace_hits_wide(const float4 *__restrict__ wide, const float4 *__restrict__ slots, uint32_t n_wide,
                     const float4 *__restrict__ rays, uint64_t n, int any_hit, bool fast, float4 *__restrict__ out, uint32_t *__restrict__ out_any) {
    for (uint64_t i = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t) gridDim.x * blockDim.x) {
        const float4 ro = __ldg(rays + 2 * i), rd = __ldg(rays + 2 * i + 1);
        const float3 o = make_float3(ro.x, ro.y, ro.z), d = make_float3(rd.x, rd.y, rd.z);
        NbHit hit;
        nb_traverse_wide<false>(wide, slots, n_wide, o, d, ro.w, rd.w, hit, fast);
        out[i] = make_float4(hit.t, hit.u, hit.v, __uint_as_float(hit.slot));
        if (any_hit) {
            NbHit dummy;
            out_any[i] = nb_traverse_wide<true>(wide, slots, n_wide, o, d, ro.w, rd.w, dummy, fast) ? 1u : 0u;
        }
    }
}

void nb_launch_trace_hits(const NbDeviceScene &scene, const float4 *rays, uint64_t n, int any_hit,
                          float4 *out, uint32_t *out_any, int sm_count, cudaStream_t stream) {
    if (n == 0) return;
    uint64_t blocks = (n + 255) / 256;
    const uint64_t cap = (uint64_t) sm_count * 64;      /* grid-stride beyond 64 CTAs per SM */
    if (blocks > cap) blocks = cap;
    if (scene.flat)
        nb_k_trace_hits<true><<<(unsigned) blocks, 256, 0, stream>>>(scene.nodes, scene.slots, scene.n_nodes, scene.n_slots, rays, n,
                                                                     any_hit, false, scene.fast_hits != 0, out, out_any);
    else if (scene.use_wide && scene.wide && !scene.ordered)
        nb_k_trace_hits_wide<<<(unsigned) blocks, 256, 0, stream>>>(scene.wide, scene.slots, scene.n_wide, rays, n, any_hit, scene.fast_hits != 0, out, out_any);
    else
        nb_k_trace_hits<false><<<(unsigned) blocks, 256, 0, stream>>>(scene.nodes, scene.slots, scene.n_nodes, scene.n_slots, rays, n,
                                                                      any_hit, scene.ordered != 0, false, out, out_any);
}

/* ---- wavefront stage: K_trace -------------------------------------------------------------
 * One launch serves both ray populations of an iteration:
 *   [0, n_main)            closest-hit queries for the live paths (BVH::rayIntersect, shadowRay=false)
 *   [n_main, n_main+n_sh)  any-hit queries for the NEE shadow rays made by the previous K_shade
 *                          (path_mis.cpp:87-89); an unoccluded one adds its pre-weighted
 *                          contribution to the path's radiance and commits the candidate w_ems.
 * Grid-stride over counts read from device memory, so the host never waits on them. */
template <bool FLAT>
__global__ void __launch_bounds__(256)
nb_k_trace_wave(const float4 *__restrict__ nodes, const float4 *__restrict__ slots, uint32_t n_nodes, uint32_t n_slots, NbWave w) {
    /* live paths = survivors of K_shade + the paths K_regen could start within the sample budget */
    const unsigned long long base = ((unsigned long long) w.counts[NB_CNT_BASE_HI] << 32) | w.counts[NB_CNT_BASE_LO];
    const unsigned long long left = w.total_samples - base;
    const uint32_t n_dead = w.counts[NB_CNT_DEAD];
    const uint32_t n_regen = left < n_dead ? (uint32_t) left : n_dead;
    const uint32_t n_main = w.counts[NB_CNT_LIVE] + n_regen, n_sh = w.counts[NB_CNT_SHADOW];
    const uint32_t total = n_main + n_sh;
    /* the other counter set is idle during this launch: publish the next iteration's inputs there */
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        const unsigned long long nb = base + n_regen;
        w.counts_next[NB_CNT_LIVE] = 0; w.counts_next[NB_CNT_SHADOW] = 0; w.counts_next[NB_CNT_DEAD] = 0;
        w.counts_next[NB_CNT_ACTIVE] = n_main;
        w.counts_next[NB_CNT_BASE_LO] = (uint32_t) nb; w.counts_next[NB_CNT_BASE_HI] = (uint32_t) (nb >> 32);
    }
    unsigned long long closest = 0, any = 0;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
        if (i < n_main) {
            const uint32_t p = w.queue_next[i];
            const float4 ro = w.ray_o[p], rd = w.ray_d[p];
            NbHit hit;
            nb_query<false, FLAT>(nodes, slots, n_nodes, n_slots, make_float3(ro.x, ro.y, ro.z), make_float3(rd.x, rd.y, rd.z),
                                  ro.w, rd.w, hit);
            w.hit[p] = make_float4(hit.t, hit.u, hit.v, __uint_as_float(hit.slot));
            if (!(rd.w < ro.w)) ++closest;      /* ended paths park their ray with maxt < mint */
        } else {
            const uint32_t j = i - n_main;
            const float4 ro = w.sh_o[j], rd = w.sh_d[j], c = w.sh_c[j];
            NbHit dummy;
            const bool occluded = nb_query<true, FLAT>(nodes, slots, n_nodes, n_slots, make_float3(ro.x, ro.y, ro.z),
                                                       make_float3(rd.x, rd.y, rd.z), ro.w, rd.w, dummy);
            ++any;
            if (!occluded) {
                const uint32_t p = __float_as_uint(c.w) & 0x7fffffffu;
                /* c.rgb = t * Li_EMS; weight: candidate w_ems when the reference would update it
                 * (path_mis.cpp:102-105), else the value carried from earlier bounces. */
                float4 aux = w.st2[p];
                float wems = aux.z;
                const float cand = w.sh_w[j];
                if (cand >= 0.f) wems = cand;
                aux.z = wems;
                w.st2[p] = aux;
                float4 li = w.st0[p];
                li.x += wems * c.x; li.y += wems * c.y; li.z += wems * c.z;
                w.st0[p] = li;
            }
        }
    }
    /* ray counters: one atomic per warp */
    for (int off = 16; off; off >>= 1) {
        closest += __shfl_xor_sync(0xffffffffu, closest, off);
        any += __shfl_xor_sync(0xffffffffu, any, off);
    }
    if ((threadIdx.x & 31) == 0) {
        if (closest) atomicAdd(&w.stats[NB_STAT_RAYS_CLOSEST], closest);
        if (any) atomicAdd(&w.stats[NB_STAT_RAYS_ANY], any);
    }
}

void nb_launch_trace_wave(const NbDeviceScene &scene, const NbWave &w, int grid, cudaStream_t stream) {
    if (scene.flat) nb_k_trace_wave<true><<<grid, 256, 0, stream>>>(scene.nodes, scene.slots, scene.n_nodes, scene.n_slots, w);
    else nb_k_trace_wave<false><<<grid, 256, 0, stream>>>(scene.nodes, scene.slots, scene.n_nodes, scene.n_slots, w);
}
