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
    griddep_launch();
    griddep_wait();
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
__device__ __forceinline__ void bn_finalize_fwd_col(const BnFwdFinalize& a, int j, double sum, double sumsq) {
    if (!a.use_bn) {
        a.mean_z[j] = 0.0f; a.rstd[j] = 1.0f; a.scale[j] = 1.0f; a.shift[j] = a.bias[j];
        return;
    }
    float mean, rstd;
    if (a.bn_mode == 0) {
        const double m = sum / a.count;
        double var = sumsq / a.count - m * m;                        // biased batch variance (normalisation)
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

__global__ void bn_finalize_fwd_kernel(BnFwdFinalize a) {
    griddep_launch();
    griddep_wait();
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j == 0 && a.bn_mode == 0 && a.use_bn && a.num_batches_tracked != nullptr) *a.num_batches_tracked += 1;
    if (j >= a.H) return;
    const bool batch = a.use_bn && a.bn_mode == 0;
    bn_finalize_fwd_col(a, j, batch ? a.sums[j] : 0.0, batch ? a.sums[a.H + j] : 0.0);
}

// single-GPU path: second stage of the column reduction and the finalize in one launch
// (block = 32 columns x 32 row lanes, like reduce_rows_kernel)
__global__ void bn_reduce_finalize_fwd_kernel(const float* __restrict__ psum, const float* __restrict__ psq, int P,
                                              BnFwdFinalize a, double* __restrict__ sums_out) {
    griddep_launch();
    griddep_wait();
    __shared__ double sh[2][32][33];
    const int col = blockIdx.x * 32 + threadIdx.x;
    const int lane_r = threadIdx.y;
    double sa = 0.0, sb = 0.0;
    if (col < a.H) {
        for (int p = lane_r; p < P; p += 32) {
            sa += static_cast<double>(psum[static_cast<long>(p) * a.H + col]);
            sb += static_cast<double>(psq[static_cast<long>(p) * a.H + col]);
        }
    }
    sh[0][lane_r][threadIdx.x] = sa;
    sh[1][lane_r][threadIdx.x] = sb;
    __syncthreads();
    if (lane_r == 0 && col < a.H) {
        double ta = 0.0, tb = 0.0;
#pragma unroll 8
        for (int r = 0; r < 32; ++r) {
            ta += sh[0][r][threadIdx.x];
            tb += sh[1][r][threadIdx.x];
        }
        sums_out[col] = ta;
        sums_out[a.H + col] = tb;
        if (col == 0 && a.num_batches_tracked != nullptr) *a.num_batches_tracked += 1;
        bn_finalize_fwd_col(a, col, ta, tb);
    }
}

// ---------------------------------------------------------------------------------- BN apply + ReLU + dropout (+ residual)
__global__ void __launch_bounds__(256) bn_apply_fwd_kernel(BnApplyFwd a, uint32_t threshold, float inv_keep) {
    griddep_launch();
    griddep_wait();
    const int groups = a.H >> 3;
    const long total = a.rows * groups;
    const long stride = static_cast<long>(gridDim.x) * blockDim.x;
    const uint16_t* zp = static_cast<const uint16_t*>(a.z);
    const uint16_t* rp = static_cast<const uint16_t*>(a.resid);
    for (long i0 = static_cast<long>(blockIdx.x) * blockDim.x + threadIdx.x; i0 < total; i0 += stride * EW_UNROLL) {
        uint4 zq[EW_UNROLL], rq[EW_UNROLL];
#pragma unroll
        for (int u = 0; u < EW_UNROLL; ++u) {                       // all loads of the group first
            const long idx = i0 + u * stride;
            zq[u] = make_uint4(0, 0, 0, 0);
            rq[u] = make_uint4(0, 0, 0, 0);
            if (idx < total) {
                zq[u] = *reinterpret_cast<const uint4*>(zp + idx * 8);
                if (rp != nullptr) rq[u] = *reinterpret_cast<const uint4*>(rp + idx * 8);
            }
        }
#pragma unroll
        for (int u = 0; u < EW_UNROLL; ++u) {
            const long idx = i0 + u * stride;
            if (idx >= total) break;
            const long row = idx / groups;
            const int cg = static_cast<int>(idx - row * groups);
            float z[8], sc[8], sh[8];
            unpack8(zq[u], a.fmt, z);
            load8f(a.scale + cg * 8, sc);
            load8f(a.shift + cg * 8, sh);
            uint32_t keep = 0xFFu;
            if (a.apply_dropout)
                keep = dropout_keep_bits8(static_cast<uint32_t>(row + a.row_offset), static_cast<uint32_t>(cg), a.site,
                                          a.step, a.seed, threshold);
            float out[8];
            uint32_t active = 0;
#pragma unroll
            for (int e = 0; e < 8; ++e) {
                const float y = fmaf(z[e], sc[e], sh[e]);
                const bool on = (y > 0.0f) && ((keep >> e) & 1u);
                active |= (on ? 1u : 0u) << e;
                out[e] = on ? y * inv_keep : 0.0f;
            }
            if (rp != nullptr) {
                float r[8];
                unpack8(rq[u], a.fmt, r);
#pragma unroll
                for (int e = 0; e < 8; ++e) out[e] += r[e];
            }
            *reinterpret_cast<uint4*>(static_cast<uint16_t*>(a.a) + idx * 8) = pack8(out, a.fmt);
            a.bits[idx] = static_cast<uint8_t>(active);
        }
    }
}

// ---------------------------------------------------------------------------------- backward passes
// Shared tiling: TPR = H/8 threads cover a row (16-byte vectors), RPB = 256/TPR rows per iteration.  A block owns
// `rows_per_block` consecutive rows and walks them EW_UNROLL iterations at a time: all loads of a group are issued
// before any arithmetic (2 x EW_UNROLL 16-byte loads in flight per thread), sums stay in registers, and one
// partial row per block leaves through a fixed-order shared-memory reduction (deterministic).
template <bool WRITE_DZ>
__global__ void __launch_bounds__(EW_THREADS) bwd_pass_kernel(BwdCommon c, int rows_per_block,
                                                              const float* __restrict__ coef,
                                                              uint16_t* __restrict__ dz, float* __restrict__ psum0,
                                                              float* __restrict__ psum1) {
    griddep_launch();
    griddep_wait();
    __shared__ float sh[2][EW_THREADS][8];
    const int tpr = c.H >> 3;
    const int rpb = EW_THREADS / tpr;
    const int cg = threadIdx.x % tpr, rl = threadIdx.x / tpr;
    const long base = static_cast<long>(blockIdx.x) * rows_per_block;
    const long end = min(base + rows_per_block, static_cast<long>(c.rows));
    float mean[8], rstd[8], k[8], c1[8], c2[8];
    load8f(c.mean_z + cg * 8, mean);
    load8f(c.rstd + cg * 8, rstd);
    if (WRITE_DZ) {
        load8f(coef + cg * 8, k);
        load8f(coef + c.H + cg * 8, c1);
        load8f(coef + 2 * c.H + cg * 8, c2);
    }
    const uint16_t* gp = static_cast<const uint16_t*>(c.g);
    const uint16_t* zp = static_cast<const uint16_t*>(c.z);
    float acc0[8] = {0, 0, 0, 0, 0, 0, 0, 0}, acc1[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    for (long r0 = base + rl; r0 < end; r0 += static_cast<long>(rpb) * EW_UNROLL) {
        uint4 gq[EW_UNROLL], zq[EW_UNROLL];
        uint32_t act[EW_UNROLL];
#pragma unroll
        for (int u = 0; u < EW_UNROLL; ++u) {
            const long row = r0 + static_cast<long>(u) * rpb;
            if (row < end) {
                const long off = row * c.H + cg * 8;
                gq[u] = *reinterpret_cast<const uint4*>(gp + off);
                zq[u] = *reinterpret_cast<const uint4*>(zp + off);
                act[u] = c.bits[row * tpr + cg];
            } else {
                gq[u] = make_uint4(0, 0, 0, 0);
                zq[u] = make_uint4(0, 0, 0, 0);
                act[u] = 0;
            }
        }
#pragma unroll
        for (int u = 0; u < EW_UNROLL; ++u) {
            const long row = r0 + static_cast<long>(u) * rpb;
            float g[8], z[8];
            unpack8(gq[u], c.fmt, g);
            unpack8(zq[u], c.fmt, z);
            if (WRITE_DZ) {
                if (row < end) {
                    float o[8];
#pragma unroll
                    for (int e = 0; e < 8; ++e) {
                        const float dy = ((act[u] >> e) & 1u) ? g[e] * c.inv_keep : 0.0f;
                        const float zh = (z[e] - mean[e]) * rstd[e];
                        o[e] = k[e] * (dy - c1[e] - zh * c2[e]);
                    }
                    const uint4 q = pack8(o, c.fmt);
                    *reinterpret_cast<uint4*>(dz + row * c.H + cg * 8) = q;
                    float r[8];
                    unpack8(q, c.fmt, r);                          // bias gradient of the values the GEMMs consume
#pragma unroll
                    for (int e = 0; e < 8; ++e) acc0[e] += r[e];
                }
            } else {
#pragma unroll
                for (int e = 0; e < 8; ++e) {                      // rows past `end` carry act == 0: they add nothing
                    const float dy = ((act[u] >> e) & 1u) ? g[e] * c.inv_keep : 0.0f;
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
    griddep_launch();
    griddep_wait();
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

// single-GPU path: reduce the pass-1 partials, write dgamma/dbeta and the pass-2 coefficients in one launch
__global__ void bn_reduce_finalize_bwd_kernel(const float* __restrict__ pdy, const float* __restrict__ pdyz, int P,
                                              BnBwdFinalize a, double* __restrict__ sums_out,
                                              float* __restrict__ dgamma, float* __restrict__ dbeta) {
    griddep_launch();
    griddep_wait();
    __shared__ double sh[2][32][33];
    const int col = blockIdx.x * 32 + threadIdx.x;
    const int lane_r = threadIdx.y;
    double sa = 0.0, sb = 0.0;
    if (col < a.H) {
        for (int p = lane_r; p < P; p += 32) {
            sa += static_cast<double>(pdy[static_cast<long>(p) * a.H + col]);
            sb += static_cast<double>(pdyz[static_cast<long>(p) * a.H + col]);
        }
    }
    sh[0][lane_r][threadIdx.x] = sa;
    sh[1][lane_r][threadIdx.x] = sb;
    __syncthreads();
    if (lane_r == 0 && col < a.H) {
        double ta = 0.0, tb = 0.0;
#pragma unroll 8
        for (int r = 0; r < 32; ++r) {
            ta += sh[0][r][threadIdx.x];
            tb += sh[1][r][threadIdx.x];
        }
        sums_out[col] = ta;
        sums_out[a.H + col] = tb;
        dbeta[col] = static_cast<float>(ta);
        dgamma[col] = static_cast<float>(tb);
        const float k = a.gamma[col] * a.rstd[col];
        a.coef[col] = k;
        a.coef[a.H + col] = a.bn_mode == 0 ? static_cast<float>(ta / a.count) : 0.0f;
        a.coef[2 * a.H + col] = a.bn_mode == 0 ? static_cast<float>(tb / a.count) : 0.0f;
    }
}

__global__ void bn_param_grads_kernel(const double* __restrict__ s, int H, float* __restrict__ dgamma,
                                      float* __restrict__ dbeta) {
    griddep_launch();
    griddep_wait();
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= H) return;
    dbeta[j] = static_cast<float>(s[j]);
    dgamma[j] = static_cast<float>(s[H + j]);
}

__global__ void wgrad_reduce_kernel(const float* __restrict__ slabs, int splits, long slab_elems, int ld, int rows,
                                    int cols, float* __restrict__ grad) {
    griddep_launch();
    griddep_wait();
    const long idx = static_cast<long>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (idx >= static_cast<long>(rows) * cols) return;
    const int r = static_cast<int>(idx / cols), c = static_cast<int>(idx - static_cast<long>(r) * cols);
    const long src = static_cast<long>(r) * ld + c;
    float s = 0.0f;
    for (int k = 0; k < splits; ++k) s += slabs[k * slab_elems + src];
    grad[idx] = s;
}

__global__ void __launch_bounds__(EW_THREADS) prep_grad_out_kernel(const float* __restrict__ dout, long rows, int O,
                                                                   int Op, int fmt, int rows_per_block,
                                                                   uint16_t* __restrict__ d16,
                                                                   float* __restrict__ psum) {
    griddep_launch();
    griddep_wait();
    // Op == 64: 8 column groups per row, 32 rows per iteration
    __shared__ float sh[EW_THREADS][8];
    const int tpr = Op >> 3, rpb = EW_THREADS / tpr;
    const int cg = threadIdx.x % tpr, rl = threadIdx.x / tpr;
    const long base = static_cast<long>(blockIdx.x) * rows_per_block;
    const long end = min(base + rows_per_block, rows);
    float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    for (long row = base + rl; row < end; row += rpb) {
        float v[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) {
            const int c = cg * 8 + e;
            v[e] = c < O ? dout[row * O + c] : 0.0f;
        }
        *reinterpret_cast<uint4*>(d16 + row * Op + cg * 8) = pack8(v, fmt);
#pragma unroll
        for (int e = 0; e < 8; ++e) acc[e] += v[e];
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

// rows owned by one block: a multiple of the rows one unrolled iteration covers, sized so that at most
// 4 x SM-count blocks (= partial rows) exist
int ew_rows_per_block(int64_t rows, int H) {
    const int64_t step = static_cast<int64_t>(EW_THREADS / (H / 8)) * EW_UNROLL;
    const int64_t steps = (rows + step - 1) / step;
    const int64_t max_blocks = static_cast<int64_t>(num_sms()) * 4;
    const int64_t steps_per_block = (steps + max_blocks - 1) / max_blocks;
    return static_cast<int>(steps_per_block * step);
}
int ew_partial_rows(int64_t rows, int H) {
    const int rpb = ew_rows_per_block(rows, H);
    return static_cast<int>((rows + rpb - 1) / rpb);
}

int reduce_rows_f64(const float* a, const float* b, int P, int n, int ld, double* out_a, double* out_b, cudaStream_t st) {
    LIFTER_CHECK_CUDA(launch_pdl(reduce_rows_kernel<double>, dim3(ceil_div(n, 32)), dim3(32, 32), 0, st, a, b, P, n, ld, out_a, out_b));
    return LIFTER_OK;
}
int reduce_rows_f32(const float* a, int P, int n, int ld, float* out, cudaStream_t st) {
    LIFTER_CHECK_CUDA(launch_pdl(reduce_rows_kernel<float>, dim3(ceil_div(n, 32)), dim3(32, 32), 0, st, a, nullptr, P, n, ld, out, nullptr));
    return LIFTER_OK;
}

int bn_finalize_fwd(const BnFwdFinalize& a, cudaStream_t st) {
    LIFTER_CHECK_CUDA(launch_pdl(bn_finalize_fwd_kernel, dim3(ceil_div(a.H, 256)), dim3(256), 0, st, a));
    return LIFTER_OK;
}

int bn_reduce_finalize_fwd(const float* psum, const float* psq, int P, const BnFwdFinalize& a, double* sums_out,
                           cudaStream_t st) {
    LIFTER_CHECK_CUDA(launch_pdl(bn_reduce_finalize_fwd_kernel, dim3(ceil_div(a.H, 32)), dim3(32, 32), 0, st, psum, psq, P, a,
                                 sums_out));
    return LIFTER_OK;
}

int bn_reduce_finalize_bwd(const float* pdy, const float* pdyz, int P, const BnBwdFinalize& a, double* sums_out,
                           float* dgamma, float* dbeta, cudaStream_t st) {
    LIFTER_CHECK_CUDA(launch_pdl(bn_reduce_finalize_bwd_kernel, dim3(ceil_div(a.H, 32)), dim3(32, 32), 0, st, pdy, pdyz, P, a,
                                 sums_out, dgamma, dbeta));
    return LIFTER_OK;
}

int bn_apply_fwd(const BnApplyFwd& a, cudaStream_t st) {
    LIFTER_REQUIRE(a.H % 8 == 0, "bn_apply: H must be a multiple of 8");
    const uint32_t thr = dropout_threshold(a.p);
    const float inv_keep = (a.apply_dropout && a.p < 1.0f) ? 1.0f / (1.0f - a.p) : (a.apply_dropout ? 0.0f : 1.0f);
    const long n = a.rows * (a.H / 8);
    long blocks = (n + 256 * EW_UNROLL - 1) / (256 * EW_UNROLL);
    const long cap = static_cast<long>(num_sms()) * 16;
    if (blocks > cap) blocks = cap;
    LIFTER_CHECK_CUDA(launch_pdl(bn_apply_fwd_kernel, dim3(static_cast<unsigned>(blocks)), dim3(256), 0, st, a, thr, inv_keep));
    return LIFTER_OK;
}

int bwd_stats(const BwdCommon& c, float* psum_dy, float* psum_dyz, cudaStream_t st) {
    LIFTER_REQUIRE(c.H >= 64 && c.H <= 2048 && (c.H & (c.H - 1)) == 0, "bwd_stats: unsupported H=%d", c.H);
    LIFTER_CHECK_CUDA(launch_pdl(bwd_pass_kernel<false>, dim3(ew_partial_rows(c.rows, c.H)), dim3(EW_THREADS), 0, st, c,
                                 ew_rows_per_block(c.rows, c.H), nullptr, nullptr, psum_dy, psum_dyz));
    return LIFTER_OK;
}

int bn_finalize_bwd(const BnBwdFinalize& a, cudaStream_t st) {
    LIFTER_CHECK_CUDA(launch_pdl(bn_finalize_bwd_kernel, dim3(ceil_div(a.H, 256)), dim3(256), 0, st, a));
    return LIFTER_OK;
}

int bn_param_grads(const double* local_sums, int H, float* dgamma, float* dbeta, cudaStream_t st) {
    LIFTER_CHECK_CUDA(launch_pdl(bn_param_grads_kernel, dim3(ceil_div(H, 256)), dim3(256), 0, st, local_sums, H, dgamma, dbeta));
    return LIFTER_OK;
}

int bwd_dz(const BwdCommon& c, const float* coef, void* dz, float* psum_dz, cudaStream_t st) {
    LIFTER_REQUIRE(c.H >= 64 && c.H <= 2048 && (c.H & (c.H - 1)) == 0, "bwd_dz: unsupported H=%d", c.H);
    LIFTER_CHECK_CUDA(launch_pdl(bwd_pass_kernel<true>, dim3(ew_partial_rows(c.rows, c.H)), dim3(EW_THREADS), 0, st, c,
                                 ew_rows_per_block(c.rows, c.H), coef, static_cast<uint16_t*>(dz),
                                                                            psum_dz, nullptr));
    return LIFTER_OK;
}

int wgrad_reduce(const float* slabs, int splits, int slab_rows, int ld, int rows, int cols, float* grad,
                 cudaStream_t st) {
    const long n = static_cast<long>(rows) * cols;
    LIFTER_CHECK_CUDA(launch_pdl(wgrad_reduce_kernel, dim3(static_cast<unsigned>((n + 255) / 256)), dim3(256), 0, st, 
        slabs, splits, static_cast<long>(slab_rows) * ld, ld, rows, cols, grad));
    return LIFTER_OK;
}

int prep_grad_out(const float* dout, int64_t rows, int O, int Op, int fmt, void* d16, float* psum, cudaStream_t st) {
    LIFTER_REQUIRE(Op == 64, "prep_grad_out: padded output width must be 64");
    LIFTER_CHECK_CUDA(launch_pdl(prep_grad_out_kernel, dim3(ew_partial_rows(rows, Op)), dim3(EW_THREADS), 0, st, dout, rows, O, Op, fmt,
                                 ew_rows_per_block(rows, Op),
                                                                       static_cast<uint16_t*>(d16), psum));
    return LIFTER_OK;
}

}  // namespace lifter
