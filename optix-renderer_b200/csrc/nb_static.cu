/* nb_static.cu — nb_trace_rays_static: the parity trace entry for callers that may not allocate.
 *
 * No counterpart in the reference; added for callers that cannot call
 * cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it). It computes exactly what
 * nb_trace_rays computes (nb_bvh.cu), but the device staging that entry cudaMallocs per call (packed rays, hit
 * records, shadow verdicts) lives here in ONE file-scope __device__ array of fixed size, and the host-side
 * repacking and slot -> (shape, prim) translation move into the kernel, so the host path needs no heap either.
 *
 *   - `nb_static_scratch` below is a host-side stand-in in host code, not a device address: its address is taken
 *     once per device with cudaGetSymbolAddress (that device current) and only that pointer is given to the kernel.
 *   - every GPU that loads the library holds the array (NB_STATIC_BYTES, 3.5 MiB) until exit.
 *   - all calls on a device share the array and it keeps the last call's contents. Nothing in it has to start at
 *     zero: a call overwrites the first n ray records and the kernel writes every field of the first n hit
 *     records before they are copied out; records past n are stale and never read.
 *   - calls that overlap on the same GPU (two contexts / two host threads) race on the array: serialise them.
 */
#include "nb_trace.cuh"
#include <mutex>

#define NB_STATIC_RAY_BYTES ((size_t) NB_TRACE_STATIC_MAX_RAYS * sizeof(nb_ray))      /* 2 MiB: nb_ray records as the caller holds them */
#define NB_STATIC_HIT_BYTES ((size_t) NB_TRACE_STATIC_MAX_RAYS * sizeof(nb_hit))      /* 1.5 MiB: finished nb_hit records */
#define NB_STATIC_BYTES (NB_STATIC_RAY_BYTES + NB_STATIC_HIT_BYTES)
#define NB_STATIC_MAX_DEVICES 64

static __device__ __align__(256) unsigned char nb_static_scratch[NB_STATIC_BYTES];

static_assert(sizeof(nb_ray) == 32 && sizeof(nb_hit) == 24, "record layout");
static_assert(NB_STATIC_RAY_BYTES % 256 == 0, "hit region alignment");

/* Variant of nb_k_trace_hits (nb_trace.cu): same queries, but it reads nb_ray records as the ABI lays them out
 * ({o.xyz, d.x} {d.yz, mint, maxt}) and writes finished nb_hit records, taking (shape, local prim) from the
 * primitive slot instead of the host-side tables nb_trace_rays uses. */
template <bool FLAT>
__global__ void __launch_bounds__(256)
nb_k_trace_hits_static(const float4 *__restrict__ nodes, const float4 *__restrict__ slots, uint32_t n_nodes, uint32_t n_slots,
                       const float4 *rays, uint32_t n, int any_hit, bool ordered, float2 *out) {
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const float4 q0 = rays[2 * i], q1 = rays[2 * i + 1];
        const float3 o = make_float3(q0.x, q0.y, q0.z), d = make_float3(q0.w, q1.x, q1.y);
        const float mint = q1.z, maxt = q1.w;
        NbHit hit;
        nb_query<false, FLAT>(nodes, slots, n_nodes, n_slots, o, d, mint, maxt, hit, ordered);
        int32_t shape = -1, prim = -1;
        if (hit.slot != 0xffffffffu) {
            const float4 s2 = __ldg(slots + 3 * (size_t) hit.slot + 2);
            shape = (int32_t) (__float_as_uint(s2.y) & NB_SLOT_SHAPE_MASK);
            prim = (int32_t) __float_as_uint(s2.z);
        }
        uint32_t occluded = 0u;
        if (any_hit) {
            NbHit dummy;
            occluded = nb_query<true, FLAT>(nodes, slots, n_nodes, n_slots, o, d, mint, maxt, dummy, ordered) ? 1u : 0u;
        }
        out[3 * (size_t) i] = make_float2(hit.t, hit.u);
        out[3 * (size_t) i + 1] = make_float2(hit.v, __int_as_float(shape));
        out[3 * (size_t) i + 2] = make_float2(__int_as_float(prim), __uint_as_float(occluded));
    }
}

/* Device address of nb_static_scratch on the CURRENT device, looked up once per device. */
static std::mutex g_static_mutex;
static unsigned char *g_static_ptr[NB_STATIC_MAX_DEVICES];

static cudaError_t nb_static_address(int device, unsigned char **ptr) {
    std::lock_guard<std::mutex> lock(g_static_mutex);
    if (!g_static_ptr[device]) {
        void *p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, nb_static_scratch);
        if (e != cudaSuccess) return e;
        g_static_ptr[device] = (unsigned char *) p;
    }
    *ptr = g_static_ptr[device];
    return cudaSuccess;
}

extern "C" int nb_trace_rays_static(nb_ctx *ctx, const nb_ray *rays, uint64_t n, int any_hit, nb_hit *out) {
    if (!ctx) return NB_E_INVALID;
    if (!ctx->bvh_ready) return nb_fail(ctx, NB_E_STATE, "nb_trace_rays_static: no BVH (call nb_build_bvh)");
    if (n > NB_TRACE_STATIC_MAX_RAYS)
        return nb_fail(ctx, NB_E_INVALID, "nb_trace_rays_static: %llu rays exceed NB_TRACE_STATIC_MAX_RAYS = %u "
                       "(fixed scratch, this entry never allocates): split the batch or call nb_trace_rays",
                       (unsigned long long) n, (unsigned) NB_TRACE_STATIC_MAX_RAYS);
    if (n == 0) return NB_OK;
    if (!rays || !out) return nb_fail(ctx, NB_E_INVALID, "nb_trace_rays_static: NULL buffer");
    if (ctx->device < 0 || ctx->device >= NB_STATIC_MAX_DEVICES)
        return nb_fail(ctx, NB_E_UNSUPPORTED, "nb_trace_rays_static: device %d beyond the %d this entry keeps an address for",
                       ctx->device, NB_STATIC_MAX_DEVICES);
    NB_CUDA(ctx, cudaSetDevice(ctx->device));
    unsigned char *scratch = nullptr;
    NB_CUDA(ctx, nb_static_address(ctx->device, &scratch));
    float4 *d_rays = (float4 *) scratch;
    float2 *d_hits = (float2 *) (scratch + NB_STATIC_RAY_BYTES);
    NB_CUDA(ctx, cudaMemcpyAsync(d_rays, rays, (size_t) n * sizeof(nb_ray), cudaMemcpyHostToDevice, ctx->stream));
    const unsigned blocks = (unsigned) ((n + 255) / 256);
    const NbDeviceScene &sc = ctx->dev;
    if (sc.flat)
        nb_k_trace_hits_static<true><<<blocks, 256, 0, ctx->stream>>>(sc.nodes, sc.slots, sc.n_nodes, sc.n_slots, d_rays,
                                                                      (uint32_t) n, any_hit, false, d_hits);
    else
        nb_k_trace_hits_static<false><<<blocks, 256, 0, ctx->stream>>>(sc.nodes, sc.slots, sc.n_nodes, sc.n_slots, d_rays,
                                                                       (uint32_t) n, any_hit, sc.ordered != 0, d_hits);
    NB_CUDA(ctx, cudaGetLastError());
    NB_CUDA(ctx, cudaMemcpyAsync(out, d_hits, (size_t) n * sizeof(nb_hit), cudaMemcpyDeviceToHost, ctx->stream));
    NB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return NB_OK;
}
