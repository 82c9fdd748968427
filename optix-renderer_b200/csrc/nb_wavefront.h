/* nb_wavefront.h — path pool, queues and counters shared by K_trace (nb_trace.cu) and
 * K_shade (nb_render.cu).
 *
 * Path state is SoA, every lane one 16-byte vector, indexed by pool slot:
 *   ray_o {o.xyz, mint}  ray_d {d.xyz, maxt}            current main ray (Ray3f)
 *   hit   {t, u, v, bits(prim slot)}                     K_trace -> K_shade
 *   st0   {Li.rgb, w_mats}                               path_mis.cpp:18,23
 *   st1   {throughput.rgb, pdfmat}                       path_mis.cpp:19; bsdf->pdf of the last sample
 *   st2   {pixelSample.x, pixelSample.y, w_ems, bits(flags | bounce << 8)}
 *   rng   {state lo, state hi, pixel index, sample index}   pcg32 (inc derives from the pixel)
 * Iteration = K_shade (consume hits, shade, NEE, RR; compact survivors / NEE rays / ended paths
 * into three queues) -> K_regen (splat ended paths into the film, start new camera paths in
 * their slots) -> K_trace (closest hit for every live path + any-hit for the NEE rays).
 * Shadow-ray queue (written by K_shade, resolved by the next K_trace):
 *   sh_o, sh_d as above; sh_c {t*Li_EMS rgb, bits(path slot)}; sh_w candidate w_ems (<0: keep)
 */
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

/* One counter set per iteration parity. K_shade fills LIVE/SHADOW/DEAD (warp-aggregated atomics);
 * K_regen and K_trace read them; K_trace publishes ACTIVE and SAMPLE_BASE for the next iteration
 * into the other set and clears that set's LIVE/SHADOW/DEAD. */
enum { NB_CNT_LIVE = 0, NB_CNT_SHADOW = 1, NB_CNT_DEAD = 2, NB_CNT_ACTIVE = 3, NB_CNT_BASE_LO = 4, NB_CNT_BASE_HI = 5,
       NB_CNT_WORDS = 8 };
enum { NB_STAT_RAYS_CLOSEST = 0, NB_STAT_RAYS_ANY = 1, NB_STAT_SAMPLES = 2, NB_STAT_INVALID = 3,
       NB_STAT_WORDS = 8 };

#define NB_FLAG_EMPTY      1u    /* slot holds no path: nothing to splat */
#define NB_FLAG_FINALIZE   2u    /* path ended (throughput became 0): splat and regenerate */
#define NB_FLAG_DISCRETE   4u    /* last BSDF sample had measure EDiscrete */
#define NB_FLAG_BSDF_OK    8u    /* last BSDF sample weight passed !isZero(Epsilon) */
#define NB_FLAG_PRIMARY   16u    /* current main ray is the camera ray */
#define NB_BOUNCE_SHIFT 8

struct NbWave {
    float4 *ray_o, *ray_d, *hit, *st0, *st1, *st2;
    uint4  *rng;
    float4 *sh_o, *sh_d, *sh_c;
    float  *sh_w;
    const uint32_t *queue_cur;   /* live path slots K_shade consumes this iteration */
    uint32_t *queue_next;        /* survivors (K_shade) then regenerated paths (K_regen); K_trace reads it */
    uint32_t *queue_dead;        /* slots whose path ended: K_shade -> K_regen */
    uint32_t *counts;            /* this iteration's counter set */
    uint32_t *counts_next;       /* the other set (published by K_trace) */
    unsigned long long total_samples;   /* camera samples of this subframe */
    unsigned long long *stats;   /* NB_STAT_* */
    uint32_t pool;
};

struct NbDeviceScene;
void nb_launch_trace_wave(const NbDeviceScene &scene, const NbWave &w, int grid, cudaStream_t stream);
