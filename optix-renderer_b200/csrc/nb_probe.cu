/* nb_probe.cu — parity entry for the medium row: Medium::sampleFreePath,
 * Medium::getTransmittance and NvdbVolume::getDensity evaluated on a fixed ray set, so the device
 * functions of nb_volume.cuh can be held against the oracle one by one (density: per point;
 * free path and transmittance: as means over `samples` draws), not only through rendered images. */
#include "nb_internal.h"
#include "nb_volume.cuh"
#include "nb_shading.cuh"

__global__ void nb_k_medium_probe(NbDeviceScene sc, int32_t medium, const float4 *rays, uint64_t n, uint32_t samples, uint64_t seed, float4 *out) {
    const uint64_t i = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float4 a = rays[2 * i], b = rays[2 * i + 1];
    const f3 o = mk3(a.x, a.y, a.z), d = mk3(a.w, b.x, b.y);
    const float maxt = b.w;
    const nb_medium &m = sc.media[medium];
    NbRng rng; rng.seed(seed, i);
    float dens = 0.f;
    if (m.type >= NB_MEDIUM_HETEROGENEOUS) dens = nb_grid_density(sc.density, m, nb_grid_point(m, o));
    float hit = 0.f, len = 0.f, tr = 0.f;
    const f3 to = o + d * maxt;
    for (uint32_t k = 0; k < samples; ++k) {
        const float t = nb_free_path(sc, medium, o, d, rng);
        if (t < maxt) { hit += 1.f; len += t; }
        tr += nb_transmittance(sc, medium, o, to, rng).x;
    }
    out[i] = make_float4(dens, hit / (float) samples, hit > 0.f ? len / hit : 0.f, tr / (float) samples);
}

/* Emitter::sample on fixed (reference point, sample) pairs: the row of every emitter plugin */
__global__ void nb_k_emitter_probe(NbDeviceScene sc, uint32_t emitter, const float *in, uint64_t n, float *out) {
    const uint64_t i = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const f3 ref = mk3(in[5 * i], in[5 * i + 1], in[5 * i + 2]);
    float2 s; s.x = in[5 * i + 3]; s.y = in[5 * i + 4];
    const nb_emitter em = sc.emitters[emitter];
    NbEmitterSample es;
    const f3 v = nb_emitter_sample(sc, emitter, ref, s, es);
    f3 e = mk3(0.f);
    if (em.type == NB_EMITTER_ENVMAP) e = nb_env_eval(sc, emitter, es.wi);
    else if (em.type == NB_EMITTER_POINT || em.type == NB_EMITTER_SPOT) e = v;
    else if (em.type == NB_EMITTER_VOLUME) e = mk3(em.radiance[0], em.radiance[1], em.radiance[2]);
    else e = v * es.pdf;
    float *o = out + 12 * i;
    o[0] = es.wi.x; o[1] = es.wi.y; o[2] = es.wi.z; o[3] = v.x; o[4] = v.y; o[5] = v.z;
    o[6] = es.pdf; o[7] = es.ray_maxt; o[8] = e.x; o[9] = e.y; o[10] = e.z; o[11] = 0.f;
}

extern "C" int nb_emitter_probe(nb_ctx *ctx, uint32_t emitter, const float *in, uint64_t n, float *out) {
    if (!ctx) return NB_E_INVALID;
    if (!ctx->scene_ready) return nb_fail(ctx, NB_E_STATE, "nb_emitter_probe: upload a scene first");
    if (emitter >= ctx->dev.n_emitters) return nb_fail(ctx, NB_E_INVALID, "nb_emitter_probe: emitter %u out of range", emitter);
    if (n == 0) return NB_OK;
    if (!in || !out) return nb_fail(ctx, NB_E_INVALID, "nb_emitter_probe: null argument");
    NB_CUDA(ctx, cudaSetDevice(ctx->device));
    float *d_in = nullptr, *d_out = nullptr;
    NB_CUDA(ctx, cudaMalloc((void **) &d_in, n * 20));
    cudaError_t e = cudaMalloc((void **) &d_out, n * 48);
    if (e == cudaSuccess) e = cudaMemcpyAsync(d_in, in, n * 20, cudaMemcpyHostToDevice, ctx->stream);
    if (e == cudaSuccess) {
        nb_k_emitter_probe<<<(unsigned) ((n + 127) / 128), 128, 0, ctx->stream>>>(ctx->dev, emitter, d_in, n, d_out);
        e = cudaMemcpyAsync(out, d_out, n * 48, cudaMemcpyDeviceToHost, ctx->stream);
    }
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    cudaFree(d_in); cudaFree(d_out);
    if (e != cudaSuccess) return nb_fail(ctx, NB_E_CUDA, "nb_emitter_probe: %s", cudaGetErrorString(e));
    return NB_OK;
}

extern "C" int nb_medium_probe(nb_ctx *ctx, int32_t medium, const nb_ray *rays, uint64_t n, uint32_t samples, uint64_t seed, float *out) {
    if (!ctx) return NB_E_INVALID;
    if (!ctx->scene_ready) return nb_fail(ctx, NB_E_STATE, "nb_medium_probe: upload a scene first");
    if (medium < 0 || medium >= (int32_t) ctx->dev.n_media) return nb_fail(ctx, NB_E_INVALID, "nb_medium_probe: medium %d out of range", medium);
    if (n == 0) return NB_OK;
    if (!rays || !out || samples == 0) return nb_fail(ctx, NB_E_INVALID, "nb_medium_probe: null argument");
    NB_CUDA(ctx, cudaSetDevice(ctx->device));
    float4 *d_rays = nullptr, *d_out = nullptr;
    NB_CUDA(ctx, cudaMalloc((void **) &d_rays, n * 32));
    cudaError_t e = cudaMalloc((void **) &d_out, n * 16);
    if (e == cudaSuccess) e = cudaMemcpyAsync(d_rays, rays, n * 32, cudaMemcpyHostToDevice, ctx->stream);
    if (e == cudaSuccess) {
        nb_k_medium_probe<<<(unsigned) ((n + 127) / 128), 128, 0, ctx->stream>>>(ctx->dev, medium, d_rays, n, samples, seed, d_out);
        e = cudaMemcpyAsync(out, d_out, n * 16, cudaMemcpyDeviceToHost, ctx->stream);
    }
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    cudaFree(d_rays); cudaFree(d_out);
    if (e != cudaSuccess) return nb_fail(ctx, NB_E_CUDA, "nb_medium_probe: %s", cudaGetErrorString(e));
    return NB_OK;
}

/* BSDF::eval / pdf / sample on fixed (wi, wo, sample, uv) records: the row of every BSDF plugin */
__global__ void nb_k_bsdf_probe(NbDeviceScene sc, uint32_t index, const float *in, uint64_t n, float *out) {
    const uint64_t i = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float *r = in + 10 * i;
    const f3 wi = mk3(r[0], r[1], r[2]), wo = mk3(r[3], r[4], r[5]);
    float2 s; s.x = r[6]; s.y = r[7];
    float2 uv; uv.x = r[8]; uv.y = r[9];
    nb_bsdf b = sc.bsdfs[index];
    nb_bsdf_resolve<true>(sc, b, uv);
    const f3 e = nb_bsdf_eval<true>(b, wi, wo, sc.params);
    const float pdf = nb_bsdf_pdf<true>(b, wi, wo);
    f3 swo; bool discrete; float spdf;
    const f3 w = nb_bsdf_sample<true>(b, wi, s, swo, discrete, spdf, sc.params);
    float *o = out + 12 * i;
    o[0] = e.x; o[1] = e.y; o[2] = e.z; o[3] = pdf; o[4] = w.x; o[5] = w.y; o[6] = w.z;
    o[7] = swo.x; o[8] = swo.y; o[9] = swo.z; o[10] = discrete ? 1.f : 0.f; o[11] = spdf;
}

extern "C" int nb_bsdf_probe(nb_ctx *ctx, uint32_t bsdf, const float *in, uint64_t n, float *out) {
    if (!ctx) return NB_E_INVALID;
    if (!ctx->scene_ready) return nb_fail(ctx, NB_E_STATE, "nb_bsdf_probe: upload a scene first");
    if (bsdf >= ctx->dev.n_bsdfs) return nb_fail(ctx, NB_E_INVALID, "nb_bsdf_probe: bsdf %u out of range", bsdf);
    if (n == 0) return NB_OK;
    if (!in || !out) return nb_fail(ctx, NB_E_INVALID, "nb_bsdf_probe: null argument");
    NB_CUDA(ctx, cudaSetDevice(ctx->device));
    float *d_in = nullptr, *d_out = nullptr;
    NB_CUDA(ctx, cudaMalloc((void **) &d_in, n * 40));
    cudaError_t e = cudaMalloc((void **) &d_out, n * 48);
    if (e == cudaSuccess) e = cudaMemcpyAsync(d_in, in, n * 40, cudaMemcpyHostToDevice, ctx->stream);
    if (e == cudaSuccess) {
        nb_k_bsdf_probe<<<(unsigned) ((n + 127) / 128), 128, 0, ctx->stream>>>(ctx->dev, bsdf, d_in, n, d_out);
        e = cudaMemcpyAsync(out, d_out, n * 48, cudaMemcpyDeviceToHost, ctx->stream);
    }
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    cudaFree(d_in); cudaFree(d_out);
    if (e != cudaSuccess) return nb_fail(ctx, NB_E_CUDA, "nb_bsdf_probe: %s", cudaGetErrorString(e));
    return NB_OK;
}

/* ---- pcg32 parity entry (ext/pcg32/pcg32.h:51-66, 101-110): the DEVICE generator's stream, word for word ---- */
__global__ void nb_k_rng_probe(unsigned long long initstate, unsigned long long initseq, uint32_t n, uint32_t *out_u, float *out_f) {
    if (blockIdx.x != 0 || threadIdx.x != 0) return;
    NbRng a; a.seed(initstate, initseq);
    for (uint32_t i = 0; i < n; ++i) out_u[i] = a.next_uint();
    NbRng b; b.seed(initstate, initseq);
    for (uint32_t i = 0; i < n; ++i) out_f[i] = b.next_float();
}

extern "C" int nb_rng_probe(nb_ctx *ctx, uint64_t initstate, uint64_t initseq, uint32_t n, uint32_t *out_u32, float *out_f32) {
    if (!ctx) return NB_E_INVALID;
    if (n == 0) return NB_OK;
    if (!out_u32 || !out_f32) return nb_fail(ctx, NB_E_INVALID, "nb_rng_probe: null argument");
    if (n > (1u << 20)) return nb_fail(ctx, NB_E_INVALID, "nb_rng_probe: at most 2^20 values per call");
    NB_CUDA(ctx, cudaSetDevice(ctx->device));
    uint32_t *d = nullptr;
    NB_CUDA(ctx, cudaMalloc((void **) &d, (size_t) n * 8));
    nb_k_rng_probe<<<1, 32, 0, ctx->stream>>>(initstate, initseq, n, d, (float *) (d + n));
    cudaError_t e = cudaMemcpyAsync(out_u32, d, (size_t) n * 4, cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(out_f32, d + n, (size_t) n * 4, cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    cudaFree(d);
    if (e != cudaSuccess) return nb_fail(ctx, NB_E_CUDA, "nb_rng_probe: %s", cudaGetErrorString(e));
    return NB_OK;
}
