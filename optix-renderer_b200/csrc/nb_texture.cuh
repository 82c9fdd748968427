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
    float st, ct, sp, cp; sincosf(theta, &st, &ct); sincosf(phi, &sp, &cp);
    return mk3(st * cp, st * sp, ct);
}
__device__ __forceinline__ float nb_luminance(f3 c) { return c.x * 0.212671f + c.y * 0.715160f + c.z * 0.072169f; }   /* color.h getLuminance */

__device__ __forceinline__ f3 nb_texture_eval(const NbDeviceScene &sc, uint32_t index, float2 uv) {
    const nb_texture &t = sc.textures[index];
    if (t.type == NB_TEXTURE_CONSTANT) return mk3(t.value[0], t.value[1], t.value[2]);
    if (t.type == NB_TEXTURE_CHECKERBOARD) {
        const float ox = uv.x / t.scale[0] - t.offset[0], oy = uv.y / t.scale[1] - t.offset[1];
        const int x = (int) ox + (ox < 0.f), y = (int) oy + (oy < 0.f);
        return ((x + y) % 2 == 0) ? mk3(t.value[0], t.value[1], t.value[2]) : mk3(t.value2[0], t.value2[1], t.value2[2]);
    }
    if (t.flags & NB_TEXTURE_SPHERICAL) {
        const f3 wi = nb_spherical_direction(uv.y * NB_PI, uv.x * 2.f * NB_PI);
        const float *R = t.rotation;
        const f3 w = mk3(R[0] * wi.x + R[1] * wi.y + R[2] * wi.z, R[3] * wi.x + R[4] * wi.y + R[5] * wi.z, R[6] * wi.x + R[7] * wi.y + R[8] * wi.z);
        const float2 c = nb_spherical_coordinates(w);
        uv.x = c.y / (2.f * NB_PI); uv.y = c.x / NB_PI;
    } else { uv.x += t.offset[0]; uv.y += t.offset[1]; }
    /* float -> unsigned as the x86-64 reference converts it (cvttss2si to 64 bits, low word kept) */
    const uint32_t w = (uint32_t) (long long) (uv.x * t.scale[0] * (float) t.width);
    const uint32_t h = t.height - (uint32_t) (long long) (uv.y * t.scale[1] * (float) t.height);
    const uint32_t idx = (h * t.width + w) % (t.width * t.height);
    const float4 c = __ldg(reinterpret_cast<const float4 *>(sc.texels) + t.first_texel + idx);
    return mk3(c.x, c.y, c.z);
}

/* Diffuse::m_albedo->eval(bRec.uv): bsdf records that carry a texture get their colour here */
__device__ __forceinline__ void nb_bsdf_resolve(const NbDeviceScene &sc, nb_bsdf &b, float2 uv) {
    const uint32_t tex = b.type >> 8;
    if (tex) {
        const f3 c = nb_texture_eval(sc, tex - 1u, uv);
        b.albedo[0] = c.x; b.albedo[1] = c.y; b.albedo[2] = c.z;
        b.type &= 0xffu;
    }
}

/* EnvMap::eval (environmentmap.cpp:111-124) for lRec.wi = d */
__device__ __forceinline__ f3 nb_env_eval(const NbDeviceScene &sc, const nb_emitter &em, f3 d) {
    const float2 c = nb_spherical_coordinates(d);
    float2 uv; uv.x = c.y / (2.f * NB_PI); uv.y = c.x / NB_PI;
    return nb_texture_eval(sc, (uint32_t) em.shape, uv) * mk3(em.radiance[0], em.radiance[1], em.radiance[2]);
}
/* Scene::getEnvMap()->eval on a ray that left the scene; black without an environment map */
__device__ __forceinline__ f3 nb_env_radiance(const NbDeviceScene &sc, f3 d) {
    if (sc.envmap_emitter < 0) return mk3(0.f);
    return nb_env_eval(sc, sc.emitters[sc.envmap_emitter], d);
}
/* EnvMap::pdf (environmentmap.cpp:98-110) */
__device__ __forceinline__ float nb_env_pdf(const NbDeviceScene &sc, const nb_emitter &em, f3 wi) {
    const nb_texture &t = sc.textures[em.shape];
    if (t.width == 1 && t.height == 1) return NB_INV_FOURPI;
    return nb_luminance(nb_env_eval(sc, em, wi)) * em.position[1] / NB_INV_FOURPI * (float) t.height * (float) t.width;
}
