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

/* ---- heterogeneous medium (src/media/heterogmedium.cpp:105-174) ---------------------------------
 * The reference samples a NanoVDB grid with its 27-tap triquadratic sampler (NvdbVolume.h:39-41,
 * nanovdb/util/SampleFromVoxels.h: TriquadraticSampler) inside the grid's bounding box, delta
 * tracking for the free path and ratio tracking for the transmittance against the majorant
 * max(0.001, densityScale * gridMax) * max(sigma_t). Here the grid is a dense array with a
 * border of NB_GRID_BORDER background voxels on every side (include/norib200.h), so the stencil
 * of any point inside the box reads in bounds; world_to_grid maps a world point to array
 * coordinates, the ray direction is mapped without renormalising, so t stays a world distance.
 * What this restates is the reference WITH the corrections SURVEY.md 8c lists (as written the
 * reference returns transmittance 0, mixes index- and world-space t, and holds a dangling accessor):
 * transmittance starts at 1, a ray that misses the box has free path infinity / transmittance 1,
 * a free path that leaves the box without a collision is infinity, t is a world distance. */
__device__ __forceinline__ f3 nb_grid_point(const nb_medium &m, f3 p) {
    const float *M = m.world_to_grid;
    return mk3(M[0] * p.x + M[1] * p.y + M[2] * p.z + M[3], M[4] * p.x + M[5] * p.y + M[6] * p.z + M[7],
               M[8] * p.x + M[9] * p.y + M[10] * p.z + M[11]);
}
__device__ __forceinline__ f3 nb_grid_dir(const nb_medium &m, f3 d) {
    const float *M = m.world_to_grid;
    return mk3(M[0] * d.x + M[1] * d.y + M[2] * d.z, M[4] * d.x + M[5] * d.y + M[6] * d.z, M[8] * d.x + M[9] * d.y + M[10] * d.z);
}

/* nanovdb::Ray::intersects(bbox, t0, t1) (nanovdb/util/Ray.h:343-363) against the box of active
 * voxels, in grid coordinates; [t0, t1] comes back clipped to [mint, maxt] */
__device__ __forceinline__ bool nb_grid_clip(const nb_medium &m, f3 og, f3 dg, float mint, float maxt, float &t0, float &t1) {
    t0 = mint; t1 = maxt;
    const float o[3] = { og.x, og.y, og.z }, d[3] = { dg.x, dg.y, dg.z };
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        const float lo = (float) NB_GRID_BORDER, hi = (float) (m.grid_dims[a] - NB_GRID_BORDER);
        if (d[a] == 0.f) { if (o[a] < lo || o[a] > hi) return false; continue; }
        const float inv = 1.f / d[a];
        float ta = (lo - o[a]) * inv, tb = (hi - o[a]) * inv;
        if (ta > tb) { const float x = ta; ta = tb; tb = x; }
        t0 = fmaxf(t0, ta); t1 = fminf(t1, tb);
    }
    return t0 <= t1;
}

/* TriquadraticSampler::sample: one axis of the quadratic B-spline-like kernel */
__device__ __forceinline__ float nb_quad_kernel(float v0, float v1, float v2, float w) {
    return w * (w * (0.5f * (v0 + v2) - v1) + 0.5f * (v2 - v0)) + v1;
}

/* NvdbVolume::getDensity (src/textures/NvdbVolume.cpp:39-48): SampleFromVoxels<.., 2> at a point
 * in grid coordinates; the stencil is ijk-1 .. ijk+1 around ijk = floor(p), z innermost */
__device__ __forceinline__ float nb_grid_density(const float *density, const nb_medium &m, f3 p) {
    const float fx = floorf(p.x), fy = floorf(p.y), fz = floorf(p.z);
    const float u = p.x - fx, v = p.y - fy, w = p.z - fz;
    const int nx = (int) m.grid_dims[0], ny = (int) m.grid_dims[1], nz = (int) m.grid_dims[2];
    const int i = (int) fx, j = (int) fy, k = (int) fz;
    const float *base = density + m.first_voxel;
    const bool bricks = m.type == NB_MEDIUM_HETEROGENEOUS_BRICKS;
    /* a stencil that would leave the array lies wholly in the border or beyond it: the sampler reads background there (the
     * tracking loops never ask — they stay inside the box — but nb_medium_probe may) */
    if (i < 1 || i > nx - 2 || j < 1 || j > ny - 2 || k < 1 || k > nz - 2)
        return __ldg(bricks ? base + (size_t) ((nx + 7) >> 3) * ((ny + 7) >> 3) * ((nz + 7) >> 3) : base);   /* brick 0 / the array's corner voxel */
    size_t sx, sy;                                     /* strides of one step in x / y; z is contiguous in both layouts */
    if (bricks) {
        /* sparse grid (include/norib200.h): table entry of the 8^3 block -> its 10^3 brick, whose apron holds the stencil */
        const int by = (ny + 7) >> 3, bz = (nz + 7) >> 3;
        const size_t entries = (size_t) ((nx + 7) >> 3) * by * bz;
        const uint32_t brick = __float_as_uint(__ldg(base + ((size_t) (i >> 3) * by + (j >> 3)) * bz + (k >> 3)));
        base += entries + (size_t) brick * NB_BRICK_VOXELS + ((i & 7) * NB_BRICK_EDGE + (j & 7)) * NB_BRICK_EDGE + (k & 7);
        sx = NB_BRICK_EDGE * NB_BRICK_EDGE; sy = NB_BRICK_EDGE;
    } else {
        base += ((size_t) (i - 1) * ny + (j - 1)) * nz + (k - 1);
        sx = (size_t) ny * nz; sy = (size_t) nz;
    }
    float vx[3];
#pragma unroll
    for (int dx = 0; dx < 3; ++dx) {
        float vy[3];
#pragma unroll
        for (int dy = 0; dy < 3; ++dy) {
            const float *row = base + dx * sx + dy * sy;
            vy[dy] = nb_quad_kernel(__ldg(row), __ldg(row + 1), __ldg(row + 2), w);
        }
        vx[dx] = nb_quad_kernel(vy[0], vy[1], vy[2], v);
    }
    return nb_quad_kernel(vx[0], vx[1], vx[2], u);
}

struct NbMajorant { float inv_mu, accept; };       /* step scale 1 / (densityMax * max sigma_t); density -> collision probability */
__device__ __forceinline__ NbMajorant nb_majorant(const nb_medium &m) {
    const float dmax = fmaxf(0.001f, m.density_scale * m.max_density);                    /* heterogmedium.cpp:81 */
    NbMajorant r;
    r.inv_mu = 1.f / (dmax * fmaxf(m.mu_t[0], fmaxf(m.mu_t[1], m.mu_t[2])));
    r.accept = m.density_scale / dmax;                                                    /* = 1 / gridMax unless clamped */
    return r;
}

/* HeterogeneousMedium::sampleFreePath (delta tracking from ray.mint = Epsilon; `ratio` = false, returns t or
 * infinity) and HeterogeneousMedium::getTransmittance (ratio tracking over [Epsilon, maxt]; `ratio` = true, returns the
 * transmittance) share one out-of-line body: the 27-tap sampler inlined at every call site of the volume
 * integrators made nb_k_tile<path_vol_mis> 160 KB of code, and ncu showed 20 `stall_no_instruction` cycles per issued
 * instruction (22 % issue-active). Plain-pointer arguments and the pcg32 state passed and returned by value keep the
 * caller's records in registers. */
struct NbTrackOut { float value; uint64_t state; };
static __device__ __noinline__ NbTrackOut nb_track_heterog(const float *density, const nb_medium *mp, float ox, float oy, float oz,
                                                           float dx, float dy, float dz, float maxt, uint64_t state, uint64_t inc, bool ratio) {
    const nb_medium &m = *mp;
    NbRng rng; rng.state = state; rng.inc = inc;
    NbTrackOut out;
    out.value = ratio ? 1.f : INFINITY;
    const f3 og = nb_grid_point(m, mk3(ox, oy, oz)), dg = nb_grid_dir(m, mk3(dx, dy, dz));
    float t, t1;
    if (nb_grid_clip(m, og, dg, NB_EPS, maxt, t, t1)) {
        const NbMajorant mj = nb_majorant(m);
        float tr = 1.f;
        for (int guard = 0; guard < (1 << 20); ++guard) {
            t -= __logf(rng.next_float()) * mj.inv_mu;
            if (!(t < t1)) break;
            const float p = nb_grid_density(density, m, og + dg * t) * mj.accept;
            if (ratio) {
                tr *= 1.f - p;
                if (tr <= 0.f) { tr = 0.f; break; }
            } else if (p >= rng.next_float()) { out.value = t; break; }
        }
        if (ratio) out.value = tr;
    }
    out.state = rng.state;
    return out;
}

__device__ __forceinline__ float nb_free_path_heterog(const NbDeviceScene &sc, int32_t medium, f3 o, f3 d, NbRng &rng) {
    const NbTrackOut r = nb_track_heterog(sc.density, sc.media + medium, o.x, o.y, o.z, d.x, d.y, d.z, INFINITY, rng.state, rng.inc, false);
    rng.state = r.state;
    return r.value;
}

/* HeterogeneousMedium::getTransmittance: ratio tracking between two points */
__device__ __forceinline__ f3 nb_transmittance_heterog(const NbDeviceScene &sc, int32_t medium, f3 from, f3 to, NbRng &rng) {
    const f3 seg = to - from;
    const float dist = length(seg);
    if (!(dist > 0.f)) return mk3(1.f);
    const f3 d = seg / dist;
    const NbTrackOut r = nb_track_heterog(sc.density, sc.media + medium, from.x, from.y, from.z, d.x, d.y, d.z, dist, rng.state, rng.inc, true);
    rng.state = r.state;
    return mk3(r.value);
}

/* Medium::sampleFreePath along the ray (o, d) from mint = Epsilon. medium < 0 is the vacuum (no draws). */
template <bool EXT = true>
__device__ __forceinline__ float nb_free_path(const NbDeviceScene &sc, int32_t medium, f3 o, f3 d, NbRng &rng) {
    if (medium < 0) return INFINITY;
    const nb_medium &m = sc.media[medium];
    if (EXT && m.type >= NB_MEDIUM_HETEROGENEOUS) return nb_free_path_heterog(sc, medium, o, d, rng);
    if (m.type != NB_MEDIUM_HOMOGENEOUS) return INFINITY;
    const int ch = (int) (3.f * rng.next_float());                       /* homogmedium.cpp:62-66 */
    const float mu = m.mu_t[ch];
    if (mu < NB_EPS) return INFINITY;                                    /* the second draw is skipped */
    return -__logf(rng.next_float()) / mu;
}

/* Medium::getTransmittance(from, to) */
template <bool EXT = true>
__device__ __forceinline__ f3 nb_transmittance(const NbDeviceScene &sc, int32_t medium, f3 from, f3 to, NbRng &rng) {
    if (medium < 0) return mk3(1.f);
    const nb_medium &m = sc.media[medium];
    if (EXT && m.type >= NB_MEDIUM_HETEROGENEOUS) return nb_transmittance_heterog(sc, medium, from, to, rng);
    if (m.type != NB_MEDIUM_HOMOGENEOUS) return mk3(1.f);
    const float dist = length(from - to);                                /* homogmedium.cpp:68-72 */
    return mk3(__expf(-m.mu_t[0] * dist), __expf(-m.mu_t[1] * dist), __expf(-m.mu_t[2] * dist));
}

/* the medium behind a surface crossed in direction w (path_vol_mats.cpp:110-113) */
__device__ __forceinline__ int32_t nb_medium_behind(const NbDeviceScene &sc, const nb_shape &sh, f3 geo_n, f3 w) {
    return (dot(w, geo_n) < 0.f && sh.medium >= 0) ? sh.medium : sc.ambient_medium;
}
