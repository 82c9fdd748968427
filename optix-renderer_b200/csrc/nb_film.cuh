/* nb_film.cuh — film accumulation (ImageBlock::put, src/utils/block.cpp:93-123), the sample plan
 * and camera-ray generation (renderBlock prologue src/utils/render.cpp:441-447 +
 * PerspectiveCamera::sampleRay src/cameras/perspective.cpp:97-141), shared by the wavefront
 * kernels (nb_render.cu) and the tile kernels (nb_mega.cu). */
#pragma once
#include "nb_internal.h"
#include "nb_wavefront.h"
#include "nb_math.cuh"

struct NbFilm {
    float4 *data;
    float4 *normal;               /* NB_LAYER_NORMAL accumulators, same geometry; null unless requested */
    float4 *pass;                 /* adaptive rounds only (nb_render_adaptive_round; pass_block != 0): the block-local images of the
                                   * current pass — renderBlock's `block` (render.cpp:421-459) for every block at once, same
                                   * geometry as data; a sample only reaches the pixels of the block_size^2 block it was drawn in */
    uint32_t pass_block;
    uint32_t rows, cols, border;
    float radius, lookup;
    float table[NB_FILTER_RESOLUTION + 1];
};

struct NbPlan {
    unsigned long long total;     /* camera samples of this subframe */
    unsigned long long seed;
    uint32_t npix, width;
    uint32_t first, stride;       /* sample index of pass k = first + k * stride */
    uint32_t max_depth;
    float iparam[3];              /* nb_render_params.integrator_params */
};

/* ImageBlock::put(pos, value), block.cpp:93-123, on the full-image block (offset 0). */
/* PASS: also fill the sample's own block image (adaptive rounds). Only the EXT kernel variants are built with it — adaptive
 * rounds run those — because the headline kernels are issue-bound: the untaken branch and its live registers cost the cbox
 * path_mis kernel 3 % (121.7 -> 125.7 ms, measured). */
template <bool PASS = false>
__device__ __forceinline__ void nb_film_put(const NbFilm &film, float px, float py, f3 v, unsigned long long *stats, float4 *layer = nullptr) {
    const bool valid = v.x >= 0.f && v.y >= 0.f && v.z >= 0.f && isfinite(v.x) && isfinite(v.y) && isfinite(v.z);
    if (!valid) { if (!layer) atomicAdd(&stats[NB_STAT_INVALID], 1ULL); return; }      /* Color3f::isValid, common.cpp:254 */
    float4 *const data = layer ? layer : film.data;
    const float posx = px - 0.5f + (float) film.border, posy = py - 0.5f + (float) film.border;
    int x0 = (int) ceilf(posx - film.radius), x1 = (int) floorf(posx + film.radius);
    int y0 = (int) ceilf(posy - film.radius), y1 = (int) floorf(posy + film.radius);
    x0 = max(x0, 0); y0 = max(y0, 0);
    x1 = min(x1, (int) film.cols - 1); y1 = min(y1, (int) film.rows - 1);
    for (int y = y0; y <= y1; ++y) {
        const float wy = film.table[(int) (fabsf((float) y - posy) * film.lookup)];
        float4 *row = data + (size_t) y * film.cols;
        for (int x = x0; x <= x1; ++x) {
            const float wgt = film.table[(int) (fabsf((float) x - posx) * film.lookup)] * wy;
            atomicAdd(row + x, make_float4(v.x * wgt, v.y * wgt, v.z * wgt, wgt));   /* RED.128 on sm_90+ */
        }
    }
    if (PASS && film.pass_block && !layer) {
        /* the same put into the sample's own block image: what AdaptiveSampler::computeVariance reads (adaptive.cpp:70-133) */
        const int bs = (int) film.pass_block, b = (int) film.border;
        const int bx = ((int) px / bs) * bs + b, by = ((int) py / bs) * bs + b;
        x0 = max(x0, bx); y0 = max(y0, by);
        x1 = min(x1, bx + bs - 1); y1 = min(y1, by + bs - 1);
        for (int y = y0; y <= y1; ++y) {
            const float wy = film.table[(int) (fabsf((float) y - posy) * film.lookup)];
            float4 *row = film.pass + (size_t) y * film.cols;
            for (int x = x0; x <= x1; ++x) {
                const float wgt = film.table[(int) (fabsf((float) x - posx) * film.lookup)] * wy;
                atomicAdd(row + x, make_float4(v.x * wgt, v.y * wgt, v.z * wgt, wgt));
            }
        }
    }
}

__device__ __forceinline__ f3 nb_xform_point(const float *m, f3 p) {           /* transform.h:83-86 */
    const float x = m[0] * p.x + m[1] * p.y + m[2] * p.z + m[3];
    const float y = m[4] * p.x + m[5] * p.y + m[6] * p.z + m[7];
    const float z = m[8] * p.x + m[9] * p.y + m[10] * p.z + m[11];
    const float w = m[12] * p.x + m[13] * p.y + m[14] * p.z + m[15];
    return mk3(x / w, y / w, z / w);
}
__device__ __forceinline__ f3 nb_xform_vector(const float *m, f3 v) {          /* transform.h:73-75 */
    return mk3(m[0] * v.x + m[1] * v.y + m[2] * v.z, m[4] * v.x + m[5] * v.y + m[6] * v.z,
               m[8] * v.x + m[9] * v.y + m[10] * v.z);
}

/* renderBlock's per-sample prologue (render.cpp:441-447) + PerspectiveCamera::sampleRay.
 * The path's pcg32 stream is seeded from (sample index, pixel index); the first four draws are
 * the pixel jitter and the (always drawn, render.cpp:443) aperture sample. */
struct NbCameraRay { f3 o, d; float mint, maxt, px, py; };
__device__ __forceinline__ void nb_camera_sample(const nb_camera &cam, const NbPlan &plan, uint32_t pix, uint32_t sidx,
                                                 NbRng &rng, NbCameraRay &r) {
    const uint32_t x = pix % plan.width, y = pix / plan.width;
    rng.seed(plan.seed * 0x9E3779B97F4A7C15ULL + sidx, pix);
    r.px = (float) x + rng.next_float(); r.py = (float) y + rng.next_float();
    const float2 ap = make_float2(rng.next_float(), rng.next_float());
    const f3 nearP = nb_xform_point(cam.sample_to_camera, mk3(r.px * cam.inv_output_size[0], r.py * cam.inv_output_size[1], 0.f));
    const f3 d0 = normalize(nearP);
    f3 o = mk3(0.f), d = d0;
    if (cam.lens_radius > NB_EPS) {
        /* thin lens; the reference draws the lens point from a hidden static sampler
         * (perspective.cpp:118-122) — here it comes from the path's own stream */
        const float2 dk = square_to_uniform_disk(ap);
        const float ft = cam.focal_distance / d0.z;
        const f3 focus = d0 * ft;
        o = mk3(cam.lens_radius * dk.x, cam.lens_radius * dk.y, 0.f);
        d = normalize(focus - o);
    }
    r.o = nb_xform_point(cam.camera_to_world, o);
    r.d = nb_xform_vector(cam.camera_to_world, d);
    const float invZ = 1.0f / d0.z;
    r.mint = cam.near_clip * invZ; r.maxt = cam.far_clip * invZ;
}
