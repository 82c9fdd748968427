/* nb_shading.cuh — device restatement of the plugin math K_shade needs: hit reconstruction,
 * the four BSDFs, the emitters. One function per reference method, cited beside it. */
#pragma once
#include "nb_internal.h"
#include "nb_math.cuh"
#include "nb_texture.cuh"

struct NbIts {          /* Intersection, include/nori/shape.h:44-76 */
    f3 p;
    float2 uv;
    NbFrame sh;
    f3 geo_n;
    uint32_t shape;
};

__device__ __forceinline__ f3 ld3(const float *a, size_t i) { return mk3(__ldg(a + 3 * i), __ldg(a + 3 * i + 1), __ldg(a + 3 * i + 2)); }

/* Texture<Normal3f>::eval of a normal map (PNGTexture.cpp:125-163 with sRGB = false) */
static __device__ __noinline__ float4 nb_normal_map_eval(const nb_texture *textures, const float *texels, uint32_t index, float u, float v) {
    const float4 c = nb_texture_eval_raw(textures, texels, index, u, v);
    const float k = textures[index].intensity;
    const f3 n = normalize(mk3(c.x * k, c.y * k, c.z * k + (1.f - k)));
    return make_float4(n.x, n.y, n.z, 0.f);
}

/* Mesh::setHitInformation (src/shapes/mesh.cpp:141-196); EXT: with the normal map of :168-180 */
template <bool EXT = true>
__device__ __forceinline__ void nb_mesh_hit_info(const NbDeviceScene &sc, const nb_shape &sh, uint32_t prim,
                                                 float u, float v, NbIts &its) {
    const float b0 = 1.f - (u + v), b1 = u, b2 = v;
    const uint32_t *f = sc.faces + ((size_t) sh.face_offset + prim) * 3;
    const size_t i0 = sh.vertex_offset + __ldg(f), i1 = sh.vertex_offset + __ldg(f + 1), i2 = sh.vertex_offset + __ldg(f + 2);
    const f3 p0 = ld3(sc.positions, i0), p1 = ld3(sc.positions, i1), p2 = ld3(sc.positions, i2);
    its.p = b0 * p0 + b1 * p1 + b2 * p2;
    its.uv = make_float2(u, v);
    if (sh.flags & NB_SHAPE_HAS_UVS) {
        its.uv.x = b0 * __ldg(sc.uvs + 2 * i0) + b1 * __ldg(sc.uvs + 2 * i1) + b2 * __ldg(sc.uvs + 2 * i2);
        its.uv.y = b0 * __ldg(sc.uvs + 2 * i0 + 1) + b1 * __ldg(sc.uvs + 2 * i1 + 1) + b2 * __ldg(sc.uvs + 2 * i2 + 1);
    }
    its.geo_n = normalize(cross(p1 - p0, p2 - p0));
    if (sh.flags & NB_SHAPE_HAS_NORMALS) {
        const f3 n = normalize(b0 * ld3(sc.normals, i0) + b1 * ld3(sc.normals, i1) + b2 * ld3(sc.normals, i2));
        if ((sh.flags & NB_SHAPE_HAS_UVS) && sc.tangents) {
            /* Frame(aTangent, aBitangent, normal), mesh.cpp:168-186 (not re-orthogonalised) */
            its.sh.s = normalize(b0 * ld3(sc.tangents, i0) + b1 * ld3(sc.tangents, i1) + b2 * ld3(sc.tangents, i2));
            its.sh.t = normalize(b0 * ld3(sc.bitangents, i0) + b1 * ld3(sc.bitangents, i1) + b2 * ld3(sc.bitangents, i2));
            its.sh.n = n;
            if (EXT && sh.normal_map) {                           /* normal = (TBN * map(uv)).normalized() */
                const float4 m = nb_normal_map_eval(sc.textures, sc.texels, sh.normal_map - 1u, its.uv.x, its.uv.y);
                its.sh.n = normalize(its.sh.s * m.x + its.sh.t * m.y + n * m.z);
            }
        } else {
            its.sh = frame_from_normal(n);
        }
    } else {
        its.sh = frame_from_normal(its.geo_n);
    }
}

/* Sphere::setHitInformation (src/shapes/sphere.cpp:96-124); EXT: with the normal map of :115-123 */
template <bool EXT = true>
__device__ __forceinline__ void nb_sphere_hit_info(const NbDeviceScene &sc, const nb_shape &sh, f3 o, f3 d, float t, NbIts &its) {
    its.p = o + d * t;
    const f3 c = mk3(sh.center[0], sh.center[1], sh.center[2]);
    const f3 n = normalize(its.p - c);
    /* sphericalCoordinates(-n), swapped, mapped to [0,1] (common.cpp:283-291) */
    float phi = atan2f(-n.y, -n.x);
    if (phi < 0.f) phi += 2.f * NB_PI;
    its.uv = make_float2(phi / (2.f * NB_PI), acosf(-n.z) / NB_PI);
    its.geo_n = n;
    const f3 tt = normalize(cross(mk3(0.f, 0.f, 1.f), n));
    its.sh.s = tt; its.sh.t = cross(n, tt); its.sh.n = n;
    if (EXT && sh.normal_map) {
        const float4 m = nb_normal_map_eval(sc.textures, sc.texels, sh.normal_map - 1u, its.uv.x, its.uv.y);
        const f3 n2 = normalize(its.sh.to_world(mk3(m.x, m.y, m.z)));
        const f3 t2 = normalize(cross(mk3(0.f, 0.f, 1.f), n2));
        its.sh.s = t2; its.sh.t = cross(n2, t2); its.sh.n = n2;
    }
}

/* fresnel(), src/utils/common.cpp:308-337 */
__device__ __forceinline__ float nb_fresnel(float cosThetaI, float extIOR, float intIOR) {
    float etaI = extIOR, etaT = intIOR;
    if (extIOR == intIOR) return 0.0f;
    if (cosThetaI < 0.0f) { const float s = etaI; etaI = etaT; etaT = s; cosThetaI = -cosThetaI; }
    const float eta = etaI / etaT, sinThetaTSqr = eta * eta * (1.f - cosThetaI * cosThetaI);
    if (sinThetaTSqr > 1.0f) return 1.0f;
    const float cosThetaT = sqrtf(1.0f - sinThetaTSqr);
    const float Rs = (etaI * cosThetaI - etaT * cosThetaT) / (etaI * cosThetaI + etaT * cosThetaT);
    const float Rp = (etaT * cosThetaI - etaI * cosThetaT) / (etaT * cosThetaI + etaI * cosThetaT);
    return (Rs * Rs + Rp * Rp) / 2.0f;
}

/* ---- microfacet (src/bsdf/microfacet.cpp:60-148) ------------------------------------------- */
__device__ __forceinline__ float nb_tan_theta(f3 v) {       /* Frame::tanTheta, frame.h */
    const float temp = 1.f - v.z * v.z;
    if (temp <= 0.0f) return 0.0f;
    return sqrtf(temp) / v.z;
}
__device__ __forceinline__ float nb_beckmann_D(f3 m, float alpha) {
    const float temp = nb_tan_theta(m) / alpha, ct = m.z, ct2 = ct * ct;
    return __expf(-temp * temp) / (NB_PI * alpha * alpha * ct2 * ct2);
}
__device__ __forceinline__ float nb_smith_G1(f3 v, f3 m, float alpha) {
    const float tanTheta = nb_tan_theta(v);
    if (tanTheta == 0.0f) return 1.0f;
    if (dot(m, v) * v.z <= 0.f) return 0.0f;
    const float a = 1.0f / (alpha * tanTheta);
    if (a >= 1.6f) return 1.0f;
    const float a2 = a * a;
    return (3.535f * a + 2.181f * a2) / (1.0f + 2.276f * a + 2.577f * a2);
}
__device__ __forceinline__ f3 nb_microfacet_eval(const nb_bsdf &b, f3 wi, f3 wo) {
    if (wo.z < 0.f) return mk3(0.f);
    const f3 wh = normalize(wi + wo);
    const float den = b.ks * nb_beckmann_D(wh, b.alpha) * nb_fresnel(dot(wh, wi), b.ext_ior, b.int_ior) *
                      nb_smith_G1(wi, wh, b.alpha) * nb_smith_G1(wo, wh, b.alpha);
    const float num = 4.f * wi.z * wo.z;
    const float spec = den / num;
    return mk3(b.albedo[0] * NB_INV_PI + spec, b.albedo[1] * NB_INV_PI + spec, b.albedo[2] * NB_INV_PI + spec);
}
__device__ __forceinline__ float nb_microfacet_pdf(const nb_bsdf &b, f3 wi, f3 wo) {
    if (wo.z <= 0.f) return 0.f;
    const f3 wh = normalize(wo + wi);
    const float part1 = b.ks * nb_beckmann_D(wh, b.alpha) * wh.z / (4.f * dot(wo, wh));
    const float part2 = (1.f - b.ks) * wo.z * NB_INV_PI;
    return part1 + part2;
}


/* ---- disney (src/bsdf/disney.cpp:111-205, helpers :275-318) -----------------------------------
 * Cold code, out of line (see nb_texture.cuh). `p` = the 10 clamped parameters in nb_bsdf order. */
__device__ __forceinline__ float nb_schlick_weight(float a) {          /* SchlickFresnel, disney.cpp:289-294 */
    const float m = fminf(fmaxf(1.f - a, 0.f), 1.f), m2 = m * m;
    return m2 * m2 * m;
}
__device__ __forceinline__ float nb_mixf(float a, float b, float t) { return a + (b - a) * t; }
__device__ __forceinline__ float nb_smith_ggx_aniso(float ndv, float vdx, float vdy, float ax, float ay) {
    return 1.f / (ndv + sqrtf(vdx * ax * vdx * ax + vdy * ay * vdy * ay + ndv * ndv));
}
__device__ __forceinline__ float nb_smith_ggx(float ndv, float alphaG) {
    const float a = alphaG * alphaG, b = ndv * ndv;
    return 1.f / (ndv + sqrtf(a + b - a * b));
}
static __device__ __noinline__ float4 nb_disney_eval_raw(const float *p, float br, float bg, float bb,
                                                         float lx, float ly, float lz, float vx, float vy, float vz) {
    const f3 L = mk3(lx, ly, lz), V = mk3(vx, vy, vz);                  /* L = bRec.wi, V = bRec.wo */
    const float NdotL = L.z, NdotV = V.z;
    if (NdotL < NB_EPS || NdotV < NB_EPS) return make_float4(0.f, 0.f, 0.f, 0.f);
    const float metallic = __ldg(p), subsurface = __ldg(p + 1), specular = __ldg(p + 2), roughness = __ldg(p + 3),
                specularTint = __ldg(p + 4), anisotropic = __ldg(p + 5), sheen = __ldg(p + 6), sheenTint = __ldg(p + 7),
                clearcoat = __ldg(p + 8), clearcoatGloss = __ldg(p + 9);
    const f3 H = normalize(L + V);
    const float NdotH = H.z, LdotH = dot(L, H);
    const f3 Cdlin = mk3(powf(br, 2.2f), powf(bg, 2.2f), powf(bb, 2.2f));        /* mon2lin */
    const float Cdlum = 0.3f * Cdlin.x + 0.6f * Cdlin.y + 0.1f * Cdlin.z;
    const f3 Ctint = Cdlum > 0.f ? Cdlin / Cdlum : mk3(1.f);
    const f3 one = mk3(1.f);
    const f3 tint = one + (Ctint - one) * specularTint;
    const f3 s0 = specular * 0.08f * tint;
    const f3 Cspec0 = s0 + (Cdlin - s0) * metallic;
    const f3 Csheen = one + (Ctint - one) * sheenTint;
    const float FL = nb_schlick_weight(NdotL), FV = nb_schlick_weight(NdotV);
    const float Fd90 = 0.5f + 2.f * LdotH * LdotH * roughness;
    const float Fd = nb_mixf(1.f, Fd90, FL) * nb_mixf(1.f, Fd90, FV);
    const float Fss90 = LdotH * LdotH * roughness;
    const float Fss = nb_mixf(1.f, Fss90, FL) * nb_mixf(1.f, Fss90, FV);
    const float ss = 1.25f * (Fss * (1.f / (NdotL + NdotV) - 0.5f) + 0.5f);
    const float aspect = sqrtf(1.f - anisotropic * 0.9f);
    const float ax = fmaxf(0.001f, roughness * roughness / aspect), ay = fmaxf(0.001f, roughness * roughness * aspect);
    const float hx = H.x / ax, hy = H.y / ay, q = hx * hx + hy * hy + NdotH * NdotH;
    const float Ds = 1.f / (NB_PI * ax * ay * q * q);                    /* GTR2_aniso */
    const float FH = nb_schlick_weight(LdotH);
    const f3 Fs = Cspec0 + (one - Cspec0) * FH;
    const float Gs = nb_smith_ggx_aniso(NdotL, L.x, L.y, ax, ay) * nb_smith_ggx_aniso(NdotV, V.x, V.y, ax, ay);
    const f3 Fsheen = FH * sheen * Csheen;
    const float a = nb_mixf(0.1f, 0.001f, clearcoatGloss);               /* GTR1(NdotH, a): a < 1 always */
    const float a2 = a * a, t = 1.f + (a2 - 1.f) * NdotH * NdotH;
    const float Dr = (a2 - 1.f) / (NB_PI * logf(a2) * t);
    const float Fr = nb_mixf(0.04f, 1.f, FH);
    const float Gr = nb_smith_ggx(NdotL, 0.25f) * nb_smith_ggx(NdotV, 0.25f);
    f3 c = (NB_INV_PI * nb_mixf(Fd, ss, subsurface) * Cdlin + Fsheen) * (1.f - metallic) + Gs * Ds * Fs + mk3(0.25f * clearcoat * Gr * Fr * Dr);
    const float lum = nb_luminance(c);
    if (lum > 1.f) c = c / lum;
    return make_float4(c.x, c.y, c.z, 0.f);
}
__device__ __forceinline__ f3 nb_disney_eval(const nb_bsdf &b, f3 wi, f3 wo, const float *params) {
    const float4 r = nb_disney_eval_raw(params + __float_as_uint(b.alpha), b.albedo[0], b.albedo[1], b.albedo[2], wi.x, wi.y, wi.z, wo.x, wo.y, wo.z);
    return mk3(r.x, r.y, r.z);
}

/* BSDF::eval for measure ESolidAngle */
template <bool EXT = true>
__device__ __forceinline__ f3 nb_bsdf_eval(const nb_bsdf &b, f3 wi, f3 wo, const float *params) {
    if (EXT && b.type == NB_BSDF_DISNEY) return nb_disney_eval(b, wi, wo, params);
    switch (b.type) {
        case NB_BSDF_DIFFUSE:                                  /* diffuse.cpp:94-103 */
            if (wi.z <= 0.f || wo.z <= 0.f) return mk3(0.f);
            return mk3(b.albedo[0], b.albedo[1], b.albedo[2]) * NB_INV_PI;
        case NB_BSDF_MICROFACET: return nb_microfacet_eval(b, wi, wo);
        default: return mk3(0.f);                              /* mirror.cpp, dielectric.cpp: discrete */
    }
}
/* BSDF::pdf for measure ESolidAngle */
template <bool EXT = true>
__device__ __forceinline__ float nb_bsdf_pdf(const nb_bsdf &b, f3 wi, f3 wo) {
    switch (b.type) {
        case NB_BSDF_DISNEY:                                   /* disney.cpp:196-203: the same cosine density */
            if (!EXT) return 0.f;
        case NB_BSDF_DIFFUSE:                                  /* diffuse.cpp:106-120 */
            if (wi.z <= 0.f || wo.z <= 0.f) return 0.f;
            return NB_INV_PI * wo.z;
        case NB_BSDF_MICROFACET: return nb_microfacet_pdf(b, wi, wo);
        default: return 0.f;
    }
}

/* BSDF::sample. Returns the weight; sets wo, `discrete` (measure == EDiscrete) and `pdf`
 * (= bsdf->pdf(bRec) for the sampled direction, needed by the MIS probe, path_mis.cpp:126). */
template <bool EXT = true>
__device__ __forceinline__ f3 nb_bsdf_sample(const nb_bsdf &b, f3 wi, float2 s, f3 &wo, bool &discrete, float &pdf, const float *params) {
    discrete = false; pdf = 0.f; wo = mk3(0.f, 0.f, 1.f);
    if (EXT && b.type == NB_BSDF_DISNEY) {                      /* disney.cpp:171-194: cosine lobe, returns eval() as is */
        if (wi.z <= 0.f) return mk3(0.f);
        wo = square_to_cosine_hemisphere(s);
        pdf = wo.z <= 0.f ? 0.f : NB_INV_PI * wo.z;
        if (pdf < NB_EPS) return mk3(0.f);
        return nb_disney_eval(b, wi, wo, params);
    }
    switch (b.type) {
        case NB_BSDF_DIFFUSE: {                                /* diffuse.cpp:123-140 */
            if (wi.z <= 0.f) return mk3(0.f);
            wo = square_to_cosine_hemisphere(s);
            pdf = wo.z <= 0.f ? 0.f : NB_INV_PI * wo.z;
            return mk3(b.albedo[0], b.albedo[1], b.albedo[2]);
        }
        case NB_BSDF_MIRROR: {                                 /* mirror.cpp:41-57 */
            if (wi.z <= 0.f) return mk3(0.f);
            wo = mk3(-wi.x, -wi.y, wi.z);
            discrete = true;
            return mk3(1.f);
        }
        case NB_BSDF_DIELECTRIC: {                             /* dielectric.cpp:51-102 */
            const float cosI = wi.z;
            const float F = nb_fresnel(cosI, b.ext_ior, b.int_ior);
            discrete = true;
            if (s.x < F) { wo = mk3(-wi.x, -wi.y, wi.z); return mk3(1.f); }
            float nz = 1.f, eta;
            if (cosI < 0.f) { nz = -1.f; eta = b.int_ior / b.ext_ior; } else eta = b.ext_ior / b.int_ior;
            const float wn = wi.z * nz;                         /* wi . normal */
            const f3 wt1 = mk3(-eta * wi.x, -eta * wi.y, -eta * (wi.z - wn * nz));
            const float c2 = -sqrtf(1.f - eta * eta * (1.f - wn * wn));
            wo = mk3(wt1.x, wt1.y, wt1.z + c2 * nz);
            return mk3(1.f / eta / eta);
        }
        case NB_BSDF_MICROFACET: {                             /* microfacet.cpp:122-148 */
            if (wi.z < 0.f) return mk3(0.f);
            if (s.y < b.ks) {
                s.y /= b.ks;
                const f3 wh = square_to_beckmann(s, b.alpha);
                wo = 2.f * (dot(wi, wh) * wh) - wi;
            } else {
                s.y = (s.y - b.ks) / (1.f - b.ks);
                wo = square_to_cosine_hemisphere(s);
            }
            if (wo.z <= 0.f) return mk3(0.f);
            pdf = nb_microfacet_pdf(b, wi, wo);
            return nb_microfacet_eval(b, wi, wo) / pdf * wo.z;
        }
        default: return mk3(0.f);
    }
}

/* ---- emitters -------------------------------------------------------------------------------- */
struct NbEmitterSample {       /* EmitterQueryRecord after Emitter::sample, emitter.h:37-67 */
    f3 wi;                     /* ref -> light */
    f3 ray_o, ray_d;           /* shadow ray (starts at the light) */
    float ray_mint, ray_maxt;
    float pdf;                 /* emitter->pdf(eqr) */
};

/* AreaEmitter::pdf (arealight.cpp:105-125) */
__device__ __forceinline__ float nb_area_pdf(const nb_shape &sh, f3 ref, f3 p, f3 n, f3 wi) {
    const float c = dot(n, -wi);
    if (c < 0.f) return 0.f;
    /* Mesh::pdfSurface (mesh.cpp:72-75) / Sphere::pdfSurface (sphere.cpp:134-137) */
    const float prob = sh.type == NB_SHAPE_SPHERE ? (1.f / sh.radius) * (1.f / sh.radius) * NB_INV_FOURPI : sh.pdf_norm;
    const f3 dlt = p - ref;
    return prob * dot(dlt, dlt) / fabsf(c);
}

/* Emitter::pdf for the emitter of a shape the path just hit: AreaEmitter::pdf, or VolumeEmitter::pdf
 * (volumelight.cpp:79-83: pdfVolume * squared distance) when a volume light sits on the shape itself */
template <bool EXT = true>
__device__ __forceinline__ float nb_hit_pdf(const NbDeviceScene &sc, const nb_shape &sh, f3 ref, f3 p, f3 n, f3 wi) {
    if (EXT) {
        const nb_emitter *e = sc.emitters + sh.emitter;
        if (e->type == NB_EMITTER_VOLUME) { const f3 dlt = p - ref; return e->position[0] * dot(dlt, dlt); }
    }
    return nb_area_pdf(sh, ref, p, n, wi);
}

/* DiscretePDF::sample (dpdf.h:124-129): lower_bound on cdf[0..n], index clamp */
__device__ __forceinline__ uint32_t nb_cdf_sample(const float *cdf, uint32_t n_entries, float x) {
    uint32_t lo = 0, hi = n_entries + 1;         /* cdf has n_entries+1 values */
    while (lo < hi) { const uint32_t mid = (lo + hi) >> 1; if (__ldg(cdf + mid) < x) lo = mid + 1; else hi = mid; }
    const int idx = (int) lo - 1;
    const uint32_t i = idx < 0 ? 0u : (uint32_t) idx;
    return i < n_entries - 1 ? i : n_entries - 1;
}

/* EnvMap::sample (environmentmap.cpp:73-97) */
__device__ __forceinline__ f3 nb_env_sample(const NbDeviceScene &sc, const nb_emitter *emp, f3 ref, float2 s, NbEmitterSample &out) {
    const nb_texture &t = sc.textures[emp->shape];
    f3 v;
    if (t.width == 1 && t.height == 1) v = square_to_uniform_sphere(s);
    else {
        const uint32_t elem = nb_cdf_sample(sc.area_cdf + __float_as_uint(emp->position[0]), t.width * t.height, s.x);
        const float i = (float) (int) (elem / t.width) / (float) t.height, j = (float) (int) (elem % t.width) / (float) t.width;
        v = nb_spherical_direction(j * NB_PI, i * 2.0f * NB_PI);
    }
    const f3 p = v * 1.f / NB_EPS;
    const f3 dlt = p - ref;
    out.wi = normalize(dlt);
    out.ray_o = p; out.ray_d = -out.wi; out.ray_mint = NB_EPS; out.ray_maxt = length(dlt) - NB_EPS;
    const float4 ep = nb_env_eval_pdf(sc.textures, sc.texels, emp, out.wi.x, out.wi.y, out.wi.z);
    out.pdf = ep.w;
    if (out.pdf < NB_EPS) return mk3(0.f);
    return mk3(ep.x, ep.y, ep.z) / out.pdf;
}


/* SpotLight::{sample,eval,falloff} (src/emitters/spotlight.cpp:53-72, 167-187) and DirectionalLight::{sample,
 * pdf,eval} (src/emitters/directionalLight.cpp:90-133; Warp::squareToUniformSphereCap{,Pdf} warp.cpp:58-72).
 * Cold code, out of line. p = the emitter's nb_scene_desc.params block. */
struct NbF8 { float4 a, b; };    /* { value rgb, pdf } { wi xyz, shadow-ray maxt }: returned by value so the caller's record stays in registers */
static __device__ __noinline__ NbF8 nb_spot_dir_sample(const float *p, const nb_emitter *emp, float rx, float ry, float rz,
                                                       float sx, float sy) {
    const nb_emitter em = *emp;
    NbEmitterSample out;
    const f3 ref = mk3(rx, ry, rz), dir = mk3(__ldg(p), __ldg(p + 1), __ldg(p + 2));
    const NbFrame coord = frame_from_normal(dir);                         /* m_coord = Frame(m_direction) */
    const f3 rad = mk3(em.radiance[0], em.radiance[1], em.radiance[2]);
    f3 ret;
    if (em.type == NB_EMITTER_SPOT) {
        const f3 pos = mk3(em.position[0], em.position[1], em.position[2]);
        const float cosStart = __ldg(p + 3), cosTotal = __ldg(p + 4);
        const f3 dlt = pos - ref;
        out.wi = normalize(dlt);
        out.ray_o = pos; out.ray_d = -out.wi; out.ray_mint = NB_EPS; out.ray_maxt = length(dlt) - NB_EPS;
        out.pdf = 1.f;
        const float ct = normalize(coord.to_local(-out.wi)).z;             /* falloff(-lRec.wi) */
        float fall;
        if (ct < cosTotal) fall = 0.f;
        else if (ct >= cosStart) fall = 1.f;
        else { const float d = (ct - cosTotal) / (cosStart - cosTotal); fall = (d * d) * (d * d); }
        const f3 i = rad / (1.f - .5f * (cosTotal + cosStart));
        const f3 rp = ref - pos;
        ret = i * fall / dot(rp, rp);
    } else {
        const float cosMax = __ldg(p + 3);
        const float z = sx * (1.0f - cosMax) + cosMax, r = sqrtf(1.0f - z * z), th = sy * 2.0f * NB_PI;
        float sn, cs; __sincosf(th, &sn, &cs);
        const f3 w = normalize(coord.to_world(mk3(r * cs, r * sn, z)));
        out.wi = -w;
        out.ray_o = ref; out.ray_d = out.wi; out.ray_mint = NB_EPS; out.ray_maxt = INFINITY;   /* Ray3f(ref, wi) defaults */
        const f3 lv = coord.to_local(w);
        out.pdf = (fabsf(dot(lv, lv) - 1.0f) < NB_EPS && lv.z <= cosMax) ? 0.0f : 1.0f / (2.0f * NB_PI * (1.0f - cosMax));
        ret = out.pdf < NB_EPS ? mk3(0.f) : rad / out.pdf;
    }
    NbF8 r;
    r.a = make_float4(ret.x, ret.y, ret.z, out.pdf);
    r.b = make_float4(out.wi.x, out.wi.y, out.wi.z, out.ray_maxt);
    return r;
}


/* VolumeEmitter::sample (src/emitters/volumelight.cpp:52-77) over Shape::sampleVolume (sphere.cpp:139-143 with
 * Warp::squareToUniformSphereVolume warp.cpp:88-92; any other shape: its bounding box, shape.cpp:97-101). The
 * reference takes the third coordinate of the volume sample from a function-local static sampler shared by all
 * render threads; here it is a hash of the two sample coordinates the integrator hands over (documented deviation:
 * same distribution, reproducible). */
static __device__ __noinline__ NbF8 nb_volume_light_sample(const nb_shape *shp, const float *box, const nb_emitter *emp,
                                                           float rx, float ry, float rz, float sx, float sy) {
    const nb_emitter em = *emp;
    const nb_shape sh = *shp;
    uint32_t h = __float_as_uint(sx) * 0x9E3779B1u ^ (__float_as_uint(sy) * 0x85EBCA77u + 0x165667B1u);
    h ^= h >> 15; h *= 0x2C1B3C6Du; h ^= h >> 12; h *= 0x297A2D39u; h ^= h >> 15;
    const float sz = __uint_as_float((h >> 9) | 0x3f800000u) - 1.0f;
    f3 p;
    if (sh.type == NB_SHAPE_SPHERE) {
        float2 s2; s2.x = sx; s2.y = sy;
        p = mk3(sh.center[0], sh.center[1], sh.center[2]) + sh.radius * (cbrtf(sz) * square_to_uniform_sphere(s2));
    } else p = mk3(__ldg(box) + __ldg(box + 3) * sx, __ldg(box + 1) + __ldg(box + 4) * sy, __ldg(box + 2) + __ldg(box + 5) * sz);
    const f3 ref = mk3(rx, ry, rz), dlt = p - ref;
    const f3 wi = normalize(dlt);
    const float d2 = dot(dlt, dlt), pdf = em.position[0] * d2;
    NbF8 r;
    const f3 v = pdf == 0.f ? mk3(0.f) : mk3(em.radiance[0], em.radiance[1], em.radiance[2]) / pdf;
    r.a = make_float4(v.x, v.y, v.z, pdf);
    r.b = make_float4(wi.x, wi.y, wi.z, sqrtf(d2) - NB_EPS);
    return r;
}

/* Emitter::sample for area (arealight.cpp:75-103 -> mesh.cpp:50-71 / sphere.cpp:126-132) and
 * point lights (pointlight.cpp:47-63). Returns radiance / pdf. */
template <bool EXT = true>
__device__ __forceinline__ f3 nb_emitter_sample(const NbDeviceScene &sc, uint32_t emitter, f3 ref, float2 s,
                                                NbEmitterSample &out) {
    const nb_emitter em = sc.emitters[emitter];
    if (em.type == NB_EMITTER_POINT) {
        const f3 pos = mk3(em.position[0], em.position[1], em.position[2]);
        const f3 dlt = ref - pos;
        const float dist = length(dlt);
        out.ray_o = pos; out.ray_d = normalize(dlt); out.ray_mint = NB_EPS; out.ray_maxt = dist - NB_EPS;
        out.wi = normalize(pos - ref);
        out.pdf = 1.f;
        return mk3(em.radiance[0], em.radiance[1], em.radiance[2]) / dot(dlt, dlt);
    }
    if (EXT && em.type == NB_EMITTER_ENVMAP) return nb_env_sample(sc, sc.emitters + emitter, ref, s, out);
    if (EXT && em.type == NB_EMITTER_VOLUME) {
        const NbF8 r = nb_volume_light_sample(sc.shapes + em.shape, sc.params + __float_as_uint(em.position[1]), sc.emitters + emitter,
                                              ref.x, ref.y, ref.z, s.x, s.y);
        out.wi = mk3(r.b.x, r.b.y, r.b.z); out.pdf = r.a.w; out.ray_mint = NB_EPS; out.ray_maxt = r.b.w;
        out.ray_o = ref; out.ray_d = out.wi;                        /* Ray3f(ref, wi, Epsilon, dist - Epsilon), volumelight.cpp:64 */
        return mk3(r.a.x, r.a.y, r.a.z);
    }
    if (EXT && em.type >= NB_EMITTER_SPOT) {
        const NbF8 r = nb_spot_dir_sample(sc.params + em.shape, sc.emitters + emitter, ref.x, ref.y, ref.z, s.x, s.y);
        out.wi = mk3(r.b.x, r.b.y, r.b.z); out.pdf = r.a.w; out.ray_mint = NB_EPS; out.ray_maxt = r.b.w;
        if (em.type == NB_EMITTER_SPOT) { out.ray_o = mk3(em.position[0], em.position[1], em.position[2]); out.ray_d = -out.wi; }
        else { out.ray_o = ref; out.ray_d = out.wi; }
        return mk3(r.a.x, r.a.y, r.a.z);
    }
    const nb_shape sh = sc.shapes[em.shape];
    f3 p, n;
    if (sh.type == NB_SHAPE_SPHERE) {
        const f3 q = square_to_uniform_sphere(s);
        p = mk3(sh.center[0], sh.center[1], sh.center[2]) + sh.radius * q;
        n = q;
    } else {
        const float *cdf = sc.area_cdf + sh.cdf_offset;
        const uint32_t tri = nb_cdf_sample(cdf, sh.prim_count, s.x);
        const float c0 = __ldg(cdf + tri), c1 = __ldg(cdf + tri + 1);
        s.x = (s.x - c0) / (c1 - c0);                           /* sampleReuse, dpdf.h:157-162 */
        const f3 bc = square_to_uniform_triangle(s);
        const uint32_t *f = sc.faces + ((size_t) sh.face_offset + tri) * 3;
        const size_t i0 = sh.vertex_offset + __ldg(f), i1 = sh.vertex_offset + __ldg(f + 1), i2 = sh.vertex_offset + __ldg(f + 2);
        const f3 p0 = ld3(sc.positions, i0), p1 = ld3(sc.positions, i1), p2 = ld3(sc.positions, i2);
        p = bc.x * p0 + bc.y * p1 + bc.z * p2;
        if (sh.flags & NB_SHAPE_HAS_NORMALS)
            n = normalize(bc.x * ld3(sc.normals, i0) + bc.y * ld3(sc.normals, i1) + bc.z * ld3(sc.normals, i2));
        else
            n = normalize(cross(p1 - p0, p2 - p0));
    }
    const f3 dlt = p - ref;
    out.wi = normalize(dlt);
    out.ray_o = p; out.ray_d = -out.wi; out.ray_mint = NB_EPS; out.ray_maxt = length(dlt) - NB_EPS;
    out.pdf = nb_area_pdf(sh, ref, p, n, out.wi);
    if (fabsf(out.pdf) < NB_EPS) return mk3(0.f);
    /* eval (arealight.cpp:58-73): front side only; pdf > 0 already implies it */
    return mk3(em.radiance[0], em.radiance[1], em.radiance[2]) / out.pdf;
}
