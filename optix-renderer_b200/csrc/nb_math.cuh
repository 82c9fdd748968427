/* nb_math.cuh — small device math used by the shading kernels: float3 algebra, Frame
 * (include/nori/frame.h:33-150, coordinateSystem src/utils/common.cpp:293-306), pcg32
 * (ext/pcg32/pcg32.h:51-110) and the sampling warps (src/utils/warp.cpp). Shading is judged
 * statistically, so FMA contraction and fast intrinsics are allowed here (not in nb_trace.cuh). */
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#define NB_EPS 1e-4f
#define NB_PI 3.14159265358979323846f
#define NB_INV_PI 0.31830988618379067154f
#define NB_INV_FOURPI 0.07957747154594766788f

typedef float3 f3;
__device__ __forceinline__ f3 mk3(float x, float y, float z) { return make_float3(x, y, z); }
__device__ __forceinline__ f3 mk3(float s) { return make_float3(s, s, s); }
__device__ __forceinline__ f3 operator+(f3 a, f3 b) { return mk3(a.x + b.x, a.y + b.y, a.z + b.z); }
__device__ __forceinline__ f3 operator-(f3 a, f3 b) { return mk3(a.x - b.x, a.y - b.y, a.z - b.z); }
__device__ __forceinline__ f3 operator-(f3 a) { return mk3(-a.x, -a.y, -a.z); }
__device__ __forceinline__ f3 operator*(f3 a, f3 b) { return mk3(a.x * b.x, a.y * b.y, a.z * b.z); }
__device__ __forceinline__ f3 operator*(f3 a, float s) { return mk3(a.x * s, a.y * s, a.z * s); }
__device__ __forceinline__ f3 operator*(float s, f3 a) { return mk3(a.x * s, a.y * s, a.z * s); }
__device__ __forceinline__ f3 operator/(f3 a, float s) { return mk3(a.x / s, a.y / s, a.z / s); }
__device__ __forceinline__ f3 &operator+=(f3 &a, f3 b) { a.x += b.x; a.y += b.y; a.z += b.z; return a; }
__device__ __forceinline__ float dot(f3 a, f3 b) { return a.x * b.x + (a.y * b.y + a.z * b.z); }
__device__ __forceinline__ f3 cross(f3 a, f3 b) {
    return mk3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}
__device__ __forceinline__ float max3(f3 a) { return fmaxf(a.x, fmaxf(a.y, a.z)); }
/* Eigen normalized(): v / sqrt(squaredNorm) when the norm is > 0 */
__device__ __forceinline__ f3 normalize(f3 a) {
    const float z = dot(a, a);
    if (z > 0.f) { const float inv = rsqrtf(z); return a * inv; }
    return a;
}
__device__ __forceinline__ float length(f3 a) { return sqrtf(dot(a, a)); }
/* Eigen isZero(prec) on an array: every |c| <= prec */
__device__ __forceinline__ bool is_zero(f3 a, float prec) {
    return fabsf(a.x) <= prec && fabsf(a.y) <= prec && fabsf(a.z) <= prec;
}
__device__ __forceinline__ bool is_black(f3 a) { return a.x == 0.f && a.y == 0.f && a.z == 0.f; }

struct NbFrame {
    f3 s, t, n;
    __device__ __forceinline__ f3 to_local(f3 v) const { return mk3(dot(v, s), dot(v, t), dot(v, n)); }
    __device__ __forceinline__ f3 to_world(f3 v) const { return s * v.x + t * v.y + n * v.z; }
};
/* Frame(n): coordinateSystem(n, s, t), common.cpp:293-306 */
__device__ __forceinline__ NbFrame frame_from_normal(f3 a) {
    NbFrame f; f.n = a;
    f3 c;
    if (fabsf(a.x) > fabsf(a.y)) {
        const float inv = rsqrtf(a.x * a.x + a.z * a.z);
        c = mk3(a.z * inv, 0.f, -a.x * inv);
    } else {
        const float inv = rsqrtf(a.y * a.y + a.z * a.z);
        c = mk3(0.f, a.z * inv, -a.y * inv);
    }
    f.t = c;
    f.s = cross(c, a);
    return f;
}

/* ---- pcg32 (ext/pcg32/pcg32.h) -------------------------------------------------------------- */
#define NB_PCG32_MULT 0x5851f42d4c957f2dULL
struct NbRng {
    uint64_t state, inc;
    __device__ __forceinline__ uint32_t next_uint() {
        const uint64_t old = state;
        state = old * NB_PCG32_MULT + inc;
        const uint32_t xorshifted = (uint32_t) (((old >> 18u) ^ old) >> 27u);
        const uint32_t rot = (uint32_t) (old >> 59u);
        return (xorshifted >> rot) | (xorshifted << ((0u - rot) & 31u));
    }
    __device__ __forceinline__ void seed(uint64_t initstate, uint64_t initseq) {
        state = 0u; inc = (initseq << 1u) | 1u;
        next_uint(); state += initstate; next_uint();
    }
    __device__ __forceinline__ float next_float() {
        return __uint_as_float((next_uint() >> 9) | 0x3f800000u) - 1.0f;
    }
};

/* ---- warps (src/utils/warp.cpp) ------------------------------------------------------------ */
__device__ __forceinline__ float2 square_to_uniform_disk(float2 s) {            /* warp.cpp:48-52 */
    const float rho = sqrtf(s.x), theta = s.y * 2.0f * NB_PI;
    float sn, cs; __sincosf(theta, &sn, &cs);
    return make_float2(rho * cs, rho * sn);
}
__device__ __forceinline__ f3 square_to_cosine_hemisphere(float2 s) {           /* warp.cpp:111-122 */
    const float2 d = square_to_uniform_disk(s);
    return mk3(d.x, d.y, sqrtf(1.f - (d.x * d.x + d.y * d.y)));
}
__device__ __forceinline__ f3 square_to_uniform_sphere(float2 s) {              /* warp.cpp:74-82 */
    const float z = 2.0f * s.x - 1.0f, r = sqrtf(1.0f - z * z), sigma = 2.0f * NB_PI * s.y;
    float sn, cs; __sincosf(sigma, &sn, &cs);
    return normalize(mk3(r * cs, r * sn, z));
}
__device__ __forceinline__ f3 square_to_uniform_triangle(float2 s) {            /* warp.cpp:162-166 */
    const float su1 = sqrtf(s.x), u = 1.f - su1, v = s.y * su1;
    return mk3(u, v, 1.f - u - v);
}
__device__ __forceinline__ f3 square_to_beckmann(float2 s, float alpha) {       /* warp.cpp:131-150 */
    float lg = __logf(1.f - s.x);
    if (isinf(lg)) lg = 0.f;
    const float tan2 = -alpha * alpha * lg, phi = s.y * 2.f * NB_PI;
    const float ct = 1.f / sqrtf(1.f + tan2), st = sqrtf(1.f - ct * ct);
    float sn, cs; __sincosf(phi, &sn, &cs);
    f3 r = mk3(st * cs, st * sn, ct);
    if (r.z < 0.f) r = -r;
    return r;
}
