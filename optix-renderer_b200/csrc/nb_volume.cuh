/* nb_volume.cuh — participating media for the volume path tracers.
 *
 * Restates Medium::sampleFreePath / getTransmittance for the vacuum and homogeneous media
 * (src/media/vacuum.cpp:29-38, src/media/homogmedium.cpp:48-73), the three phase functions
 * (src/bsdf/isophase.cpp:20-28, anisophase.cpp:24-33, schlickphase.cpp:24-33 with their warps,
 * src/utils/warp.cpp:168-241) and one loop iteration of PathVolMATSIntegrator::Li
 * (src/integrators/path_vol_mats.cpp:20-122) / PathVolMISIntegrator::Li with its sampleEmitter and
 * traceShadowray helpers (src/integrators/path_vol_mis.cpp:26-243).
 *
 * The reference's estimator is reproduced as written, including what a textbook would call
 * mistakes: throughput is multiplied by the transmittance but never divided by the free-path pdf,
 * a medium interaction scatters with weight 1 (no albedo), emitter sampling at a medium
 * interaction uses the SURFACE point behind it as reference point and the surface frame for the
 * phase pdf, and the shadow ray's transmittance only counts segments that end on an interface.
 * The t-test scenes and images this path is checked against were produced by exactly that code. */
#pragma once
#include "nb_shading.cuh"

enum { NB_MEASURE_SOLID_ANGLE = 0, NB_MEASURE_DISCRETE = 1, NB_MEASURE_UNKNOWN = 2 };   /* EMeasure, common.h */
#define NB_MEASURE_SHIFT 4          /* kept in NbPathRegs::flags bits 4..5 by the volume integrators */

/* Warp::squareToHenyeyGreenstein / Pdf (warp.cpp:168-205) */
__device__ __forceinline__ f3 nb_dir_from_cos(float cosTheta, float u) {
    const float sinTheta = sqrtf(1.f - cosTheta * cosTheta);
    float sn, cs; __sincosf(2.f * NB_PI * u, &sn, &cs);
    return mk3(sinTheta * cs, sinTheta * sn, cosTheta);
}
__device__ __forceinline__ f3 square_to_henyey_greenstein(float2 s, float g) {
    float cosTheta;
    if (fabsf(g) < NB_EPS) cosTheta = 1.f - 2.f * s.x;
    else {
        const float factor = (1.f - g * g) / (1.f - g + 2.f * g * s.x);
        cosTheta = (1.f + g * g - factor * factor) / (2.f * g);
    }
    return nb_dir_from_cos(cosTheta, s.y);
}
__device__ __forceinline__ float henyey_greenstein_pdf(f3 m, float g) {
    const float g2 = g * g, b = 1.f + g2 - 2.f * g * m.z;
    return NB_INV_FOURPI * (1.f - g2) / (b * sqrtf(b));                 /* pow(b, 1.5) */
}
/* Warp::squareToSchlick / Pdf (warp.cpp:207-241) */
__device__ __forceinline__ f3 square_to_schlick(float2 s, float k) {
    float cosTheta;
    if (fabsf(k) < NB_EPS) cosTheta = 1.f - 2.f * s.x;
    else cosTheta = 1.f / k * (1.f - (1.f - k) / (4.f * NB_PI * k / (1.f + k) * (s.x - 1.f) + 1.f));
    return nb_dir_from_cos(cosTheta, s.y);
}
__device__ __forceinline__ float schlick_pdf(f3 m, float k) {
    const float factor = 1.f - k * m.z;
    return NB_INV_FOURPI * (1.f - k * k) / (factor * factor);
}

/* PhaseFunction::sample / pdf around wi = (0, 0, 1) */
__device__ __forceinline__ f3 nb_phase_sample(const nb_medium &m, float2 s) {
    if (m.phase == NB_PHASE_HENYEY_GREENSTEIN) return square_to_henyey_greenstein(s, m.phase_param);
    if (m.phase == NB_PHASE_SCHLICK) return square_to_schlick(s, m.phase_param);
    return square_to_uniform_sphere(s);
}
__device__ __forceinline__ float nb_phase_pdf(const nb_medium &m, f3 wo) {
    if (m.phase == NB_PHASE_HENYEY_GREENSTEIN) return henyey_greenstein_pdf(wo, m.phase_param);
    if (m.phase == NB_PHASE_SCHLICK) return schlick_pdf(wo, m.phase_param);
    return NB_INV_FOURPI;
}

/* Medium::sampleFreePath. medium < 0 is the vacuum (no draws). */
__device__ __forceinline__ float nb_free_path(const NbDeviceScene &sc, int32_t medium, NbRng &rng) {
    if (medium < 0) return INFINITY;
    const nb_medium &m = sc.media[medium];
    if (m.type != NB_MEDIUM_HOMOGENEOUS) return INFINITY;
    const int ch = (int) (3.f * rng.next_float());                       /* homogmedium.cpp:62-66 */
    const float mu = m.mu_t[ch];
    if (mu < NB_EPS) return INFINITY;                                    /* the second draw is skipped */
    return -__logf(rng.next_float()) / mu;
}

/* Medium::getTransmittance(from, to) */
__device__ __forceinline__ f3 nb_transmittance(const NbDeviceScene &sc, int32_t medium, f3 from, f3 to) {
    if (medium < 0) return mk3(1.f);
    const nb_medium &m = sc.media[medium];
    if (m.type != NB_MEDIUM_HOMOGENEOUS) return mk3(1.f);
    const float dist = length(from - to);                                /* homogmedium.cpp:68-72 */
    return mk3(__expf(-m.mu_t[0] * dist), __expf(-m.mu_t[1] * dist), __expf(-m.mu_t[2] * dist));
}

/* the medium behind a surface crossed in direction w (path_vol_mats.cpp:110-113) */
__device__ __forceinline__ int32_t nb_medium_behind(const NbDeviceScene &sc, const nb_shape &sh, f3 geo_n, f3 w) {
    return (dot(w, geo_n) < 0.f && sh.medium >= 0) ? sh.medium : sc.ambient_medium;
}
