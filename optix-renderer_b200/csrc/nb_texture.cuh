/* nb_texture.cuh — Texture<Color3f>::eval for the three colour textures and the environment-map
 * emitter built on it. Restates ConstantTexture (src/textures/consttexture.cpp:37), Checkerboard
 * (src/textures/checkerboard.cpp:34-52), PNGTexture::eval (src/textures/PNGTexture.cpp:125-165,
 * sRGB images; normal maps have no kernel) and EnvMap::{eval,pdf,sample}
 * (src/emitters/environmentmap.cpp:73-140) as written, index arithmetic included: nearest-texel
 * lookup with `h = height - uint(v * scaleV * height)` wrapped modulo width*height, the
 * luminance CDF walked with u = i/height, v = j/width, the sampled direction
 * sphericalDirection(j*pi, i*2pi), the point at v/Epsilon and the pdf
 * luminance * normalization * 4pi * width * height. */
#pragma once
#include "nb_internal.h"
#include "nb_math.cuh"

__device__ __forceinline__ float2 nb_spherical_coordinates(f3 v) {           /* common.cpp:283-291: (theta, phi) */
    float2 r; r.x = acosf(v.z); r.y = atan2f(v.y, v.x);
    if (r.y < 0.f) r.y += 2.f * NB_PI;
    return r;
}
__device__ __forceinline__ f3 nb_spherical_direction(float theta, float phi) {  /* common.cpp:273-281 */
    float st, ct, sp, cp; __sincosf(theta, &st, &ct); __sincosf(phi, &sp, &cp);   /* arguments in [0, 2 pi] */
    return mk3(st * cp, st * sp, ct);
}
__device__ __forceinline__ float nb_luminance(f3 c) { return c.x * 0.212671f + c.y * 0.715160f + c.z * 0.072169f; }   /* color.h getLuminance */

/* Cold code, kept out of line so that scenes without textures or an environment map pay one
 * predicated call and no registers: the out-of-line functions take plain pointers (a reference to
 * the kernel's NbDeviceScene parameter would force a copy of it into local memory). */
static __device__ __noinline__ float4 nb_texture_eval_raw(const nb_texture *textures, const float *texels, uint32_t index, float u, float v) {
    const nb_texture &t = textures[index];
    if (t.type == NB_TEXTURE_CONSTANT) return make_float4(t.value[0], t.value[1], t.value[2], 0.f);
    if (t.type == NB_TEXTURE_CHECKERBOARD) {
        const float ox = u / t.scale[0] - t.offset[0], oy = v / t.scale[1] - t.offset[1];
        const int x = (int) ox + (ox < 0.f), y = (int) oy + (oy < 0.f);
        return ((x + y) % 2 == 0) ? make_float4(t.value[0], t.value[1], t.value[2], 0.f) : make_float4(t.value2[0], t.value2[1], t.value2[2], 0.f);
    }
    if (t.flags & NB_TEXTURE_SPHERICAL) {
        const f3 wi = nb_spherical_direction(v * NB_PI, u * 2.f * NB_PI);
        const float *R = t.rotation;
        const f3 w = mk3(R[0] * wi.x + R[1] * wi.y + R[2] * wi.z, R[3] * wi.x + R[4] * wi.y + R[5] * wi.z, R[6] * wi.x + R[7] * wi.y + R[8] * wi.z);
        const float2 c = nb_spherical_coordinates(w);
        u = c.y / (2.f * NB_PI); v = c.x / NB_PI;
    } else { u += t.offset[0]; v += t.offset[1]; }
    /* float -> unsigned as the x86-64 reference converts it (cvttss2si to 64 bits, low word kept) */
    const uint32_t w = (uint32_t) (long long) (u * t.scale[0] * (float) t.width);
    const uint32_t h = t.height - (uint32_t) (long long) (v * t.scale[1] * (float) t.height);
    const uint32_t idx = (h * t.width + w) % (t.width * t.height);
    return __ldg(reinterpret_cast<const float4 *>(texels) + t.first_texel + idx);
}
__device__ __forceinline__ f3 nb_texture_eval(const NbDeviceScene &sc, uint32_t index, float2 uv) {
    const float4 c = nb_texture_eval_raw(sc.textures, sc.texels, index, uv.x, uv.y);
    return mk3(c.x, c.y, c.z);
}

/* EXT: the kernels are compiled twice — EXT = false for scenes without textures, environment map or
 * heterogeneous media (nb_scene_ext, nb_internal.h), where none of that code exists, and EXT = true. */

/* Diffuse::m_albedo->eval(bRec.uv): bsdf records that carry a texture get their colour here */
template <bool EXT = true>
__device__ __forceinline__ void nb_bsdf_resolve(const NbDeviceScene &sc, nb_bsdf &b, float2 uv) {
    if (!EXT) return;
    const uint32_t tex = b.type >> 8;
    if (tex) {
        const f3 c = nb_texture_eval(sc, tex - 1u, uv);
        b.albedo[0] = c.x; b.albedo[1] = c.y; b.albedo[2] = c.z;
        b.type &= 0xffu;
    }
}

/* EnvMap::eval (environmentmap.cpp:111-124) for lRec.wi = d, and EnvMap::pdf (:98-110) in .w */
static __device__ __noinline__ float4 nb_env_eval_pdf(const nb_texture *textures, const float *texels, const nb_emitter *emp, float dx, float dy, float dz) {
    const nb_emitter em = *emp;
    const float2 c = nb_spherical_coordinates(mk3(dx, dy, dz));
    const float4 t = nb_texture_eval_raw(textures, texels, (uint32_t) em.shape, c.y / (2.f * NB_PI), c.x / NB_PI);
    const f3 e = mk3(t.x, t.y, t.z) * mk3(em.radiance[0], em.radiance[1], em.radiance[2]);
    const nb_texture &tx = textures[em.shape];
    float pdf = NB_INV_FOURPI;
    if (!(tx.width == 1 && tx.height == 1)) pdf = nb_luminance(e) * em.position[1] / NB_INV_FOURPI * (float) tx.height * (float) tx.width;
    return make_float4(e.x, e.y, e.z, pdf);
}
__device__ __forceinline__ f3 nb_env_eval(const NbDeviceScene &sc, uint32_t emitter, f3 d) {
    const float4 r = nb_env_eval_pdf(sc.textures, sc.texels, sc.emitters + emitter, d.x, d.y, d.z);
    return mk3(r.x, r.y, r.z);
}
/* Scene::getEnvMap()->eval on a ray that left the scene; black without an environment map */
template <bool EXT = true>
__device__ __forceinline__ f3 nb_env_radiance(const NbDeviceScene &sc, f3 d) {
    if (!EXT || sc.envmap_emitter < 0) return mk3(0.f);
    return nb_env_eval(sc, (uint32_t) sc.envmap_emitter, d);
}
