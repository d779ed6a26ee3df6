#include "train_kernels.cuh"

#include "philox.cuh"

namespace lifter {

namespace {

__device__ __forceinline__ void unpack8(const uint4& q, int fmt, float (&v)[8]) {
    const uint32_t w[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
    for (int e = 0; e < 4; ++e) {
        const float2 f = unpack16(w[e], fmt);
        v[2 * e] = f.x;
        v[2 * e + 1] = f.y;
    }
}
__device__ __forceinline__ uint4 pack8(const float (&v)[8], int fmt) {
    return make_uint4(pack16(v[0], v[1], fmt), pack16(v[2], v[3], fmt), pack16(v[4], v[5], fmt),
                      pack16(v[6], v[7], fmt));
}
__device__ __forceinline__ void load8f(const float* p, float (&v)[8]) {
    const float4 a = *reinterpret_cast<const float4*>(p), b = *reinterpret_cast<const float4*>(p + 4);
    v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
}

// ---------------------------------------------------------------------------------- row reductions
// block = 32 columns x 32 row-lanes; every thread sums its rows in order, then a fixed-order tree over lanes
template <typename OutT>
__global__ void reduce_rows_kernel(const float* __restrict__ a, const float* __restrict__ b, int P, int n, int ld,
                                   OutT* __restrict__ out_a, OutT* __restrict__ out_b) {
    __shared__ double sh[2][32][33];
    const int col = blockIdx.x * 32 + threadIdx.x;
    const int lane_r = threadIdx.y;
    double sa = 0.0, sb = 0.0;
    if (col < n) {
        for (int p = lane_r; p < P; p += 32) {
            sa += static_cast<double>(a[static_cast<long>(p) * ld + col]);
            if (b != nullptr) sb += static_cast<double>(b[static_cast<long>(p) * ld + col]);
        }
    }
    sh[0][lane_r][threadIdx.x] = sa;
    sh[1][lane_r][threadIdx.x] = sb;
    __syncthreads();
    if (lane_r == 0 && col < n) {
        double ta = 0.0, tb = 0.0;
#pragma unroll 8
        for (int r = 0; r < 32; ++r) {
            ta += sh[0][r][threadIdx.x];
            tb += sh[1][r][threadIdx.x];
        }
        out_a[col] = static_cast<OutT>(ta);
        if (b != nullptr) out_b[col] = static_cast<OutT>(tb);
    }
}

// ---------------------------------------------------------------------------------- BN finalize (forward)
__global__ void bn_finalize_fwd_kernel(BnFwdFinalize a) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j == 0 && a.bn_mode == 0 && a.use_bn && a.num_batches_tracked != nullptr) *a.num_batches_tracked += 1;
    if (j >= a.H) return;
    if (!a.use_bn) {
        a.mean_z[j] = 0.0f; a.rstd[j] = 1.0f; a.scale[j] = 1.0f; a.shift[j] = a.bias[j];
        return;
    }
    float mean, rstd;
    if (a.bn_mode == 0) {
        const double m = a.sums[j] / a.count;
        double var = a.sums[a.H + j] / a.count - m * m;              // biased batch variance (normalisation)
        if (var < 0.0) var = 0.0;
        mean = static_cast<float>(m);
        rstd = static_cast<float>(1.0 / sqrt(var + static_cast<double>(a.eps)));
        if (a.running_mean != nullptr) {                             // torch: momentum 0.1, unbiased running var
            const double unbiased = a.count > 1.0 ? var * a.count / (a.count - 1.0) : var;
            a.running_mean[j] = (1.0f - a.momentum) * a.running_mean[j] +
                                a.momentum * static_cast<float>(m + static_cast<double>(a.bias[j]));
            a.running_var[j] = (1.0f - a.momentum) * a.running_var[j] + a.momentum * static_cast<float>(unbiased);
        }
    } else {                                                         // running statistics; z excludes the bias
        mean = a.running_mean[j] - a.bias[j];
        rstd = 1.0f / sqrtf(a.running_var[j] + a.eps);
    }
    const float s = a.gamma[j] * rstd;
    a.mean_z[j] = mean;
    a.rstd[j] = rstd;
    a.scale[j] = s;
    a.shift[j] = a.beta[j] - mean * s;
}

// ---------------------------------------------------------------------------------- BN apply + ReLU + dropout (+ residual)
__global__ void bn_apply_fwd_kernel(BnApplyFwd a, uint32_t threshold, float inv_keep) {
    const int groups = a.H >> 3;
    const long idx = static_cast<long>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (idx >= a.rows * groups) return;
    const long row = idx / groups;
    const int cg = static_cast<int>(idx - row * groups);
    const long off = row * a.H + cg * 8;
    float z[8], sc[8], sh[8];
    unpack8(*reinterpret_cast<const uint4*>(static_cast<const uint16_t*>(a.z) + off), a.fmt, z);
    load8f(a.scale + cg * 8, sc);
    load8f(a.shift + cg * 8, sh);
    uint32_t keep = 0xFFu;
    if (a.apply_dropout)
        keep = dropout_keep_bits8(static_cast<uint32_t>(row + a.row_offset), static_cast<uint32_t>(cg), a.site, a.step,
                                  a.seed, threshold);
    float out[8];
    uint32_t active = 0;
#pragma unroll
    for (int e = 0; e < 8; ++e) {
        const float y = fmaf(z[e], sc[e], sh[e]);
        const bool on = (y > 0.0f) && ((keep >> e) & 1u);
        active |= (on ? 1u : 0u) << e;
        out[e] = on ? y * inv_keep : 0.0f;
    }
    if (a.resid != nullptr) {
        float r[8];
        unpack8(*reinterpret_cast<const uint4*>(static_cast<const uint16_t*>(a.resid) + off), a.fmt, r);
#pragma unroll
        for (int e = 0; e < 8; ++e) out[e] += r[e];
    }
    *reinterpret_cast<uint4*>(static_cast<uint16_t*>(a.a) + off) = pack8(out, a.fmt);
    a.bits[idx] = static_cast<uint8_t>(active);
}

// ---------------------------------------------------------------------------------- backward passes
// Shared tiling: TPR = H/8 threads cover a row, RPB = 256/TPR rows per iteration, EW_ITERS iterations per block.
template <bool WRITE_DZ>
__global__ void __launch_bounds__(EW_THREADS) bwd_pass_kernel(BwdCommon c, const float* __restrict__ coef,
                                                              uint16_t* __restrict__ dz, float* __restrict__ psum0,
                                                              float* __restrict__ psum1) {
    __shared__ float sh[2][EW_THREADS][8];
    const int tpr = c.H >> 3;
    const int rpb = EW_THREADS / tpr;
    const int cg = threadIdx.x % tpr, rl = threadIdx.x / tpr;
    const long base = static_cast<long>(blockIdx.x) * rpb * EW_ITERS;
    float mean[8], rstd[8], k[8], c1[8], c2[8];
    load8f(c.mean_z + cg * 8, mean);
    load8f(c.rstd + cg * 8, rstd);
    if (WRITE_DZ) {
        load8f(coef + cg * 8, k);
        load8f(coef + c.H + cg * 8, c1);
        load8f(coef + 2 * c.H + cg * 8, c2);
    }
    float acc0[8] = {0, 0, 0, 0, 0, 0, 0, 0}, acc1[8] = {0, 0, 0, 0, 0, 0, 0, 0};
#pragma unroll
    for (int it = 0; it < EW_ITERS; ++it) {
        const long row = base + it * rpb + rl;
        if (row < c.rows) {
            const long off = row * c.H + cg * 8;
            float g[8], z[8];
            unpack8(*reinterpret_cast<const uint4*>(static_cast<const uint16_t*>(c.g) + off), c.fmt, g);
            unpack8(*reinterpret_cast<const uint4*>(static_cast<const uint16_t*>(c.z) + off), c.fmt, z);
            const uint32_t active = c.bits[row * tpr + cg];
            if (WRITE_DZ) {
                float o[8];
#pragma unroll
                for (int e = 0; e < 8; ++e) {
                    const float dy = ((active >> e) & 1u) ? g[e] * c.inv_keep : 0.0f;
                    const float zh = (z[e] - mean[e]) * rstd[e];
                    o[e] = k[e] * (dy - c1[e] - zh * c2[e]);
                }
                const uint4 q = pack8(o, c.fmt);
                *reinterpret_cast<uint4*>(dz + off) = q;
                float r[8];
                unpack8(q, c.fmt, r);                              // bias gradient of the values the GEMMs consume
#pragma unroll
                for (int e = 0; e < 8; ++e) acc0[e] += r[e];
            } else {
#pragma unroll
                for (int e = 0; e < 8; ++e) {
                    const float dy = ((active >> e) & 1u) ? g[e] * c.inv_keep : 0.0f;
                    acc0[e] += dy;
                    acc1[e] += dy * (z[e] - mean[e]) * rstd[e];
                }
            }
        }
    }
#pragma unroll
    for (int e = 0; e < 8; ++e) {
        sh[0][threadIdx.x][e] = acc0[e];
        sh[1][threadIdx.x][e] = acc1[e];
    }
    __syncthreads();
    if (rl == 0) {
        float t0[8] = {0, 0, 0, 0, 0, 0, 0, 0}, t1[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        for (int r = 0; r < rpb; ++r) {
#pragma unroll
            for (int e = 0; e < 8; ++e) {
                t0[e] += sh[0][r * tpr + cg][e];
                t1[e] += sh[1][r * tpr + cg][e];
            }
        }
        float* o0 = psum0 + static_cast<long>(blockIdx.x) * c.H + cg * 8;
        *reinterpret_cast<float4*>(o0) = make_float4(t0[0], t0[1], t0[2], t0[3]);
        *reinterpret_cast<float4*>(o0 + 4) = make_float4(t0[4], t0[5], t0[6], t0[7]);
        if (!WRITE_DZ) {
            float* o1 = psum1 + static_cast<long>(blockIdx.x) * c.H + cg * 8;
            *reinterpret_cast<float4*>(o1) = make_float4(t1[0], t1[1], t1[2], t1[3]);
            *reinterpret_cast<float4*>(o1 + 4) = make_float4(t1[4], t1[5], t1[6], t1[7]);
        }
    }
}

__global__ void bn_finalize_bwd_kernel(BnBwdFinalize a) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= a.H) return;
    float k = 1.0f, c1 = 0.0f, c2 = 0.0f;
    if (a.use_bn) {
        k = a.gamma[j] * a.rstd[j];
        if (a.bn_mode == 0) {
            c1 = static_cast<float>(a.sums[j] / a.count);
            c2 = static_cast<float>(a.sums[a.H + j] / a.count);
        }
    }
    a.coef[j] = k;
    a.coef[a.H + j] = c1;
    a.coef[2 * a.H + j] = c2;
}

__global__ void bn_param_grads_kernel(const double* __restrict__ s, int H, float* __restrict__ dgamma,
                                      float* __restrict__ dbeta) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= H) return;
    dbeta[j] = static_cast<float>(s[j]);
    dgamma[j] = static_cast<float>(s[H + j]);
}

__global__ void wgrad_reduce_kernel(const float* __restrict__ slabs, int splits, long slab_elems, int ld, int rows,
                                    int cols, float* __restrict__ grad) {
    const long idx = static_cast<long>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (idx >= static_cast<long>(rows) * cols) return;
    const int r = static_cast<int>(idx / cols), c = static_cast<int>(idx - static_cast<long>(r) * cols);
    const long src = static_cast<long>(r) * ld + c;
    float s = 0.0f;
    for (int k = 0; k < splits; ++k) s += slabs[k * slab_elems + src];
    grad[idx] = s;
}

__global__ void __launch_bounds__(EW_THREADS) prep_grad_out_kernel(const float* __restrict__ dout, long rows, int O,
                                                                   int Op, int fmt, uint16_t* __restrict__ d16,
                                                                   float* __restrict__ psum) {
    // Op == 64: 8 column groups per row, 32 rows per iteration, EW_ITERS iterations per block
    __shared__ float sh[EW_THREADS][8];
    const int tpr = Op >> 3, rpb = EW_THREADS / tpr;
    const int cg = threadIdx.x % tpr, rl = threadIdx.x / tpr;
    const long base = static_cast<long>(blockIdx.x) * rpb * EW_ITERS;
    float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
#pragma unroll
    for (int it = 0; it < EW_ITERS; ++it) {
        const long row = base + it * rpb + rl;
        if (row < rows) {
            float v[8];
#pragma unroll
            for (int e = 0; e < 8; ++e) {
                const int c = cg * 8 + e;
                v[e] = c < O ? dout[row * O + c] : 0.0f;
            }
            const uint4 q = pack8(v, fmt);
            *reinterpret_cast<uint4*>(d16 + row * Op + cg * 8) = q;
#pragma unroll
            for (int e = 0; e < 8; ++e) acc[e] += v[e];
        }
    }
#pragma unroll
    for (int e = 0; e < 8; ++e) sh[threadIdx.x][e] = acc[e];
    __syncthreads();
    if (rl == 0) {
        float t[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        for (int r = 0; r < rpb; ++r)
#pragma unroll
            for (int e = 0; e < 8; ++e) t[e] += sh[r * tpr + cg][e];
        float* o = psum + static_cast<long>(blockIdx.x) * Op + cg * 8;
#pragma unroll
        for (int e = 0; e < 8; ++e) o[e] = t[e];
    }
}

}  // namespace

int ew_rows_per_block(int H) { return (EW_THREADS / (H / 8)) * EW_ITERS; }
int ew_partial_rows(int64_t rows, int H) {
    const int rpb = ew_rows_per_block(H);
    return static_cast<int>((rows + rpb - 1) / rpb);
}

int reduce_rows_f64(const float* a, const float* b, int P, int n, int ld, double* out_a, double* out_b, cudaStream_t st) {
    reduce_rows_kernel<double><<<ceil_div(n, 32), dim3(32, 32), 0, st>>>(a, b, P, n, ld, out_a, out_b);
    LIFTER_CHECK_CUDA(cudaGetLastError());
    return LIFTER_OK;
}
int reduce_rows_f32(const float* a, int P, int n, int ld, float* out, cudaStream_t st) {
    reduce_rows_kernel<float><<<ceil_div(n, 32), dim3(32, 32), 0, st>>>(a, nullptr, P, n, ld, out, nullptr);
    LIFTER_CHECK_CUDA(cudaGetLastError());
    return LIFTER_OK;
}

int bn_finalize_fwd(const BnFwdFinalize& a, cudaStream_t st) {
    bn_finalize_fwd_kernel<<<ceil_div(a.H, 256), 256, 0, st>>>(a);
    LIFTER_CHECK_CUDA(cudaGetLastError());
    return LIFTER_OK;
}

int bn_apply_fwd(const BnApplyFwd& a, cudaStream_t st) {
    LIFTER_REQUIRE(a.H % 8 == 0, "bn_apply: H must be a multiple of 8");
    const uint32_t thr = dropout_threshold(a.p);
    const float inv_keep = (a.apply_dropout && a.p < 1.0f) ? 1.0f / (1.0f - a.p) : (a.apply_dropout ? 0.0f : 1.0f);
    const long n = a.rows * (a.H / 8);
    bn_apply_fwd_kernel<<<static_cast<unsigned>((n + 255) / 256), 256, 0, st>>>(a, thr, inv_keep);
    LIFTER_CHECK_CUDA(cudaGetLastError());
    return LIFTER_OK;
}

int bwd_stats(const BwdCommon& c, float* psum_dy, float* psum_dyz, cudaStream_t st) {
    LIFTER_REQUIRE(c.H >= 64 && c.H <= 2048 && (c.H & (c.H - 1)) == 0, "bwd_stats: unsupported H=%d", c.H);
    bwd_pass_kernel<false><<<ew_partial_rows(c.rows, c.H), EW_THREADS, 0, st>>>(c, nullptr, nullptr, psum_dy, psum_dyz);
    LIFTER_CHECK_CUDA(cudaGetLastError());
    return LIFTER_OK;
}

int bn_finalize_bwd(const BnBwdFinalize& a, cudaStream_t st) {
    bn_finalize_bwd_kernel<<<ceil_div(a.H, 256), 256, 0, st>>>(a);
    LIFTER_CHECK_CUDA(cudaGetLastError());
    return LIFTER_OK;
}

int bn_param_grads(const double* local_sums, int H, float* dgamma, float* dbeta, cudaStream_t st) {
    bn_param_grads_kernel<<<ceil_div(H, 256), 256, 0, st>>>(local_sums, H, dgamma, dbeta);
    LIFTER_CHECK_CUDA(cudaGetLastError());
    return LIFTER_OK;
}

int bwd_dz(const BwdCommon& c, const float* coef, void* dz, float* psum_dz, cudaStream_t st) {
    LIFTER_REQUIRE(c.H >= 64 && c.H <= 2048 && (c.H & (c.H - 1)) == 0, "bwd_dz: unsupported H=%d", c.H);
    bwd_pass_kernel<true><<<ew_partial_rows(c.rows, c.H), EW_THREADS, 0, st>>>(c, coef, static_cast<uint16_t*>(dz),
                                                                            psum_dz, nullptr);
    LIFTER_CHECK_CUDA(cudaGetLastError());
    return LIFTER_OK;
}

int wgrad_reduce(const float* slabs, int splits, int slab_rows, int ld, int rows, int cols, float* grad,
                 cudaStream_t st) {
    const long n = static_cast<long>(rows) * cols;
    wgrad_reduce_kernel<<<static_cast<unsigned>((n + 255) / 256), 256, 0, st>>>(
        slabs, splits, static_cast<long>(slab_rows) * ld, ld, rows, cols, grad);
    LIFTER_CHECK_CUDA(cudaGetLastError());
    return LIFTER_OK;
}

int prep_grad_out(const float* dout, int64_t rows, int O, int Op, int fmt, void* d16, float* psum, cudaStream_t st) {
    LIFTER_REQUIRE(Op == 64, "prep_grad_out: padded output width must be 64");
    prep_grad_out_kernel<<<ew_partial_rows(rows, Op), EW_THREADS, 0, st>>>(dout, rows, O, Op, fmt,
                                                                       static_cast<uint16_t*>(d16), psum);
    LIFTER_CHECK_CUDA(cudaGetLastError());
    return LIFTER_OK;
}

}  // namespace lifter
