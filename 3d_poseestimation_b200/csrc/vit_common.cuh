// Device helpers shared by the MyViT forward (vit_ops.cu) and training (vit_train.cu) kernels.
#pragma once
#include <math_constants.h>

#include "common.cuh"

namespace lifter {

// Ampere-style asynchronous 16-byte copies global -> shared (no register staging): a warp that stages its whole working set
// issues every copy before waiting once, so one (pose, head) has all its loads in flight instead of one per loop trip.
__device__ __forceinline__ void cp_async16(uint32_t smem_addr, const void* gptr) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" :: "r"(smem_addr), "l"(gptr) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() {
    asm volatile("cp.async.commit_group;\n\tcp.async.wait_group 0;" ::: "memory");
}
namespace {

constexpr int VIT_MAX_TOKENS = 32;

__device__ __forceinline__ float wsum(float v) {
#pragma unroll
    for (int s = 16; s >= 1; s >>= 1) v += __shfl_xor_sync(0xffffffffu, v, s);
    return v;
}

// A warp owns token rows of width Hd = NV * 256: lane l holds columns v * 256 + l * 8 + [0, 8), v < NV.  Each warp walks
// VIT_RPW consecutive rows with everything per-column (LayerNorm gains / biases, Linear weights) loaded ONCE into
// registers: re-reading 8-16 parameter vectors per row made the first version of these kernels L1-bandwidth bound.
constexpr int VIT_RPW = 8;

template <int NV>
struct RowParams { float g[NV][8], b[NV][8]; };

template <int NV>
__device__ __forceinline__ void load_row_params(RowParams<NV>& p, const float* __restrict__ g, const float* __restrict__ b, int lane) {
#pragma unroll
    for (int v = 0; v < NV; ++v) {
        const int c0 = v * 256 + lane * 8;
        const float4 g0 = *reinterpret_cast<const float4*>(g + c0), g1 = *reinterpret_cast<const float4*>(g + c0 + 4);
        const float4 b0 = *reinterpret_cast<const float4*>(b + c0), b1 = *reinterpret_cast<const float4*>(b + c0 + 4);
        p.g[v][0] = g0.x; p.g[v][1] = g0.y; p.g[v][2] = g0.z; p.g[v][3] = g0.w;
        p.g[v][4] = g1.x; p.g[v][5] = g1.y; p.g[v][6] = g1.z; p.g[v][7] = g1.w;
        p.b[v][0] = b0.x; p.b[v][1] = b0.y; p.b[v][2] = b0.z; p.b[v][3] = b0.w;
        p.b[v][4] = b1.x; p.b[v][5] = b1.y; p.b[v][6] = b1.z; p.b[v][7] = b1.w;
    }
}

template <int NV>
__device__ __forceinline__ void warp_layernorm(float (&x)[NV][8], const RowParams<NV>& p, float eps) {
    constexpr float inv = 1.0f / (NV * 256);
    float s = 0.0f;
#pragma unroll
    for (int v = 0; v < NV; ++v)
#pragma unroll
        for (int e = 0; e < 8; ++e) s += x[v][e];
    const float mean = wsum(s) * inv;
    float q = 0.0f;
#pragma unroll
    for (int v = 0; v < NV; ++v)
#pragma unroll
        for (int e = 0; e < 8; ++e) { const float d = x[v][e] - mean; q = fmaf(d, d, q); }
    const float rstd = rsqrtf(wsum(q) * inv + eps);                  // biased variance, torch.nn.LayerNorm
#pragma unroll
    for (int v = 0; v < NV; ++v)
#pragma unroll
        for (int e = 0; e < 8; ++e) x[v][e] = fmaf((x[v][e] - mean) * rstd, p.g[v][e], p.b[v][e]);
}

template <int NV>
__device__ __forceinline__ void store_row16(void* out, long row, int lane, const float (&x)[NV][8], int fmt) {
    uint4* dst = reinterpret_cast<uint4*>(static_cast<unsigned short*>(out) + row * (NV * 256));
#pragma unroll
    for (int v = 0; v < NV; ++v)
        dst[v * 32 + lane] = make_uint4(pack16(x[v][0], x[v][1], fmt), pack16(x[v][2], x[v][3], fmt),
                                        pack16(x[v][4], x[v][5], fmt), pack16(x[v][6], x[v][7], fmt));
}


template <int NV>
__device__ __forceinline__ void load_row16(const void* in, long row, int lane, float (&x)[NV][8], int fmt) {
    const uint4* src = reinterpret_cast<const uint4*>(static_cast<const unsigned short*>(in) + row * (NV * 256));
#pragma unroll
    for (int v = 0; v < NV; ++v) {
        const uint4 r = src[v * 32 + lane];
        const uint32_t wd[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            const float2 f = unpack16(wd[e], fmt);
            x[v][2 * e] = f.x;
            x[v][2 * e + 1] = f.y;
        }
    }
}

// fp32 rows of the residual stream (eval path): lane l holds columns v * 256 + l * 8 + [0, 8) as two float4
template <int NV>
__device__ __forceinline__ void store_row32(float* out, long row, int lane, const float (&x)[NV][8]) {
    float4* dst = reinterpret_cast<float4*>(out + row * (NV * 256));
#pragma unroll
    for (int v = 0; v < NV; ++v) {
        dst[v * 64 + lane * 2] = make_float4(x[v][0], x[v][1], x[v][2], x[v][3]);
        dst[v * 64 + lane * 2 + 1] = make_float4(x[v][4], x[v][5], x[v][6], x[v][7]);
    }
}

#define VIT_DISPATCH_NV(nv, CALL)                     \
    switch (nv) {                                     \
        case 1: { constexpr int NV = 1; CALL; break; } \
        case 2: { constexpr int NV = 2; CALL; break; } \
        case 3: { constexpr int NV = 3; CALL; break; } \
        default: { constexpr int NV = 4; CALL; break; } \
    }

}  // namespace
}  // namespace lifter
