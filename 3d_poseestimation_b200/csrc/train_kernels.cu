#include "train_kernels.cuh"

#include "bn_device.cuh"
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
// streaming 16-byte accesses: these tensors are touched once per kernel, keep them out of L1
__device__ __forceinline__ uint4 ldg_stream(const void* p) {
    uint4 r;
    asm volatile("ld.global.L1::no_allocate.v4.b32 {%0, %1, %2, %3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}

// ---------------------------------------------------------------------------------- column reductions
// Second stage of every column reduction: P partial rows -> one value per column, fixed order (deterministic),
// fp64 accumulation.  Block = 8 columns x 128 row-lanes (a warp reads 4 rows x 32 B: whole sectors), one block per
// 8 columns, so every thread issues only P/128 loads per array and the chain of dependent L2 round trips that made
// the first version take ~10 us is gone; the row-lanes are combined by a fixed-pairing tree in shared memory.
constexpr int RED_COLS = 8, RED_LANES = 128;

__device__ __forceinline__ void reduce_partials_2(const float* __restrict__ a, const float* __restrict__ b, int P, int n,
                                                  int ld, double (&sh)[2][RED_LANES][RED_COLS], double& ta, double& tb) {
    const int col = blockIdx.x * RED_COLS + threadIdx.x;
    const int lane_r = threadIdx.y;
    double sa = 0.0, sb = 0.0;
    if (col < n) {
        for (int p = lane_r; p < P; p += RED_LANES) {
            sa += static_cast<double>(ldg_f32(a + static_cast<long>(p) * ld + col));
            if (b != nullptr) sb += static_cast<double>(ldg_f32(b + static_cast<long>(p) * ld + col));
        }
    }
    sh[0][lane_r][threadIdx.x] = sa;
    sh[1][lane_r][threadIdx.x] = sb;
    __syncthreads();
#pragma unroll
    for (int s = RED_LANES / 2; s >= 1; s >>= 1) {
        if (lane_r < s) {
            sh[0][lane_r][threadIdx.x] += sh[0][lane_r + s][threadIdx.x];
            sh[1][lane_r][threadIdx.x] += sh[1][lane_r + s][threadIdx.x];
        }
        __syncthreads();
    }
    ta = sh[0][0][threadIdx.x];
    tb = sh[1][0][threadIdx.x];
}

template <typename OutT>
__global__ void __launch_bounds__(RED_COLS * RED_LANES, 1) reduce_rows_kernel(const float* __restrict__ a, const float* __restrict__ b, int P, int n, int ld,
                                   OutT* __restrict__ out_a, OutT* __restrict__ out_b) {
    griddep_launch();
    griddep_wait();
    __shared__ double sh[2][RED_LANES][RED_COLS];
    double ta, tb;
    reduce_partials_2(a, b, P, n, ld, sh, ta, tb);
    const int col = blockIdx.x * RED_COLS + threadIdx.x;
    if (threadIdx.y == 0 && col < n) {
        out_a[col] = static_cast<OutT>(ta);
        if (b != nullptr) out_b[col] = static_cast<OutT>(tb);
    }
}

// ---------------------------------------------------------------------------------- BN finalize (forward)
__global__ void bn_finalize_fwd_kernel(BnFwdFinalize a) {
    griddep_launch();
    griddep_wait();
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j == 0 && a.bn_mode == 0 && a.use_bn && a.num_batches_tracked != nullptr) *a.num_batches_tracked += 1;
    if (j >= a.H) return;
    const bool batch = a.use_bn && a.bn_mode == 0;
    bn_finalize_fwd_col(a, j, batch ? a.sums[j] : 0.0, batch ? a.sums[a.H + j] : 0.0, bn_column_inputs(a, j));
}

// single-GPU path: second stage of the column reduction and the finalize in one launch
__global__ void __launch_bounds__(RED_COLS * RED_LANES, 1) bn_reduce_finalize_fwd_kernel(const float* __restrict__ psum, const float* __restrict__ psq, int P,
                                              BnFwdFinalize a, double* __restrict__ sums_out) {
    griddep_launch();
    griddep_wait();
    __shared__ double sh[2][RED_LANES][RED_COLS];
    double ta, tb;
    const int col = blockIdx.x * RED_COLS + threadIdx.x;
    BnColumnInputs ci = {0.0f, 1.0f, 0.0f, 0.0f, 1.0f};
    if (threadIdx.y == 0 && col < a.H) ci = bn_column_inputs(a, col);
    reduce_partials_2(psum, psq, P, a.H, a.H, sh, ta, tb);
    if (a.ex.world > 1 && threadIdx.y < 4) exchange_column_sums_warp(a.ex, a.H, col, col < a.H, threadIdx.y, ta, tb);   // warp 0
    if (threadIdx.y == 0 && col < a.H) {
        sums_out[col] = ta;
        sums_out[a.H + col] = tb;
        if (col == 0 && a.num_batches_tracked != nullptr) *a.num_batches_tracked += 1;
        bn_finalize_fwd_col(a, col, ta, tb, ci);
    }
}

// ---------------------------------------------------------------------------------- streaming tiling
// Shared by the three big elementwise kernels: TPR = H/8 threads cover a row (16-byte vectors), RPB = 256/TPR rows
// per iteration; a block owns `rows_per_block` consecutive rows and walks them EW_UNROLL iterations at a time with
// all loads of a group issued before any arithmetic.  A thread keeps the same 8 columns for its whole life, so
// every per-column coefficient is loaded once into registers.

// BN affine + ReLU + Philox dropout (+ block residual): a = drop(relu(z * scale + shift)) [+ resid], and one bit per
// element "ReLU-active AND kept" for the backward pass.  1/(1-p) is folded into scale/shift (it is positive, so the
// ReLU decision is unchanged).
//
// A thread owns 8 columns and walks AP_ROWS = 16 CONSECUTIVE rows per group: that is the footprint of one Philox call
// with one-bit lanes (p = 0.5, philox.cuh), so the generator costs one call per 128 elements instead of one per 8 and
// the kernel is bound by its loads and stores, not by instruction issue (round 1: 13.7 us per 8,192 x 1,024 layer at
// 50 % issue utilisation).  The rows of a group are loaded AP_SUB at a time, the next sub-batch in flight while the
// current one is computed; a warp instruction still touches 32 x 16 contiguous bytes of one row.
constexpr int AP_ROWS = 16;
constexpr int AP_SUB = 4;
__device__ __forceinline__ uint4 and4(const uint4& a, const uint4& m) {
    return make_uint4(a.x & m.x, a.y & m.y, a.z & m.z, a.w & m.w);
}

template <int FMT>
__device__ __forceinline__ void apply_rows(const BnApplyFwd& a, long base, long end, const DropoutLanes& lanes, float inv_keep,
                                           const PhiloxKeys& keys, const float* __restrict__ scale,
                                           const float* __restrict__ shift) {
    const int tpr = a.H >> 3;
    const int rpb = EW_THREADS / tpr;
    const int cg = threadIdx.x % tpr, rl = threadIdx.x / tpr;
    float sc[8], sh[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) {                   // past L1: another block of this launch may have written them
        sc[e] = __ldcg(scale + cg * 8 + e) * inv_keep;
        sh[e] = __ldcg(shift + cg * 8 + e) * inv_keep;
    }
    const uint32_t dstep = (a.step_dev != nullptr ? *a.step_dev : a.step) * a.step_mul + a.step_add;
    const bool has_resid = a.resid != nullptr, dropout = a.apply_dropout != 0;
    const uint32_t rmask = (1u << lanes.rows_shift) - 1u;
    const long H = a.H;

    for (long g0 = base + static_cast<long>(rl) * AP_ROWS; g0 < end; g0 += static_cast<long>(rpb) * AP_ROWS) {
        const int n = static_cast<int>(min(static_cast<long>(AP_ROWS), end - g0));
        const uint16_t* zp = static_cast<const uint16_t*>(a.z) + g0 * H + cg * 8;
        const uint16_t* rp = has_resid ? static_cast<const uint16_t*>(a.resid) + g0 * H + cg * 8 : nullptr;
        uint16_t* ap = static_cast<uint16_t*>(a.a) + g0 * H + cg * 8;
        uint8_t* bp = a.bits + g0 * tpr + cg;
        const uint32_t grow0 = static_cast<uint32_t>(g0 + a.row_offset);          // global row of this group's first row
        uint4 zq[2][AP_SUB], rq[2][AP_SUB];
        auto load = [&](int s, uint4 (&zb)[AP_SUB], uint4 (&rb)[AP_SUB]) {
#pragma unroll
            for (int u = 0; u < AP_SUB; ++u) {
                const int r = s * AP_SUB + u;
                zb[u] = make_uint4(0, 0, 0, 0);
                rb[u] = make_uint4(0, 0, 0, 0);
                if (r < n) {
                    zb[u] = ldg_stream(zp + r * H);
                    if (has_resid) rb[u] = ldg_stream(rp + r * H);
                }
            }
        };
        Philox4 rnd = {0u, 0u, 0u, 0u};
        load(0, zq[0], rq[0]);
#pragma unroll
        for (int s = 0; s < AP_ROWS / AP_SUB; ++s) {
            if ((s + 1) * AP_SUB < n) load(s + 1, zq[(s + 1) & 1], rq[(s + 1) & 1]);
#pragma unroll
            for (int u = 0; u < AP_SUB; ++u) {
                const int r = s * AP_SUB + u;
                if (r >= n) break;
                const uint32_t grow = grow0 + static_cast<uint32_t>(r);
                uint32_t keep = 0xFFu;
                if (dropout) {
                    if (r == 0 || (grow & rmask) == 0u)                      // first row of a call's footprint
                        rnd = philox4x32_10(grow >> lanes.rows_shift, static_cast<uint32_t>(cg), a.site, dstep, keys);
                    keep = dropout_keep8(rnd, grow, lanes.bits, lanes.threshold);
                }
                float z[8], out[8];
                unpack8(zq[s & 1][u], FMT, z);
                uint32_t active = 0;
#pragma unroll
                for (int e = 0; e < 8; ++e) {
                    const float y = fmaf(z[e], sc[e], sh[e]);
                    const bool on = ((keep >> e) & 1u) && (y > 0.0f);
                    active |= on ? (1u << e) : 0u;
                    out[e] = on ? y : 0.0f;
                }
                if (has_resid) {
                    float rr[8];
                    unpack8(rq[s & 1][u], FMT, rr);
#pragma unroll
                    for (int e = 0; e < 8; ++e) out[e] += rr[e];
                }
                *reinterpret_cast<uint4*>(ap + r * H) = pack8(out, FMT);
                bp[static_cast<long>(r) * tpr] = static_cast<uint8_t>(active);
            }
        }
    }
}

template <int FMT>
__global__ void __launch_bounds__(EW_THREADS, 2) bn_apply_fwd_kernel(BnApplyFwd a, int rows_per_block, DropoutLanes lanes,
                                                                  float inv_keep, const PhiloxKeys keys) {
    griddep_launch();
    griddep_wait();
    const long base = static_cast<long>(blockIdx.x) * rows_per_block;
    const long end = min(base + rows_per_block, static_cast<long>(a.rows));
    apply_rows<FMT>(a, base, end, lanes, inv_keep, keys, a.scale, a.shift);
}

// ---- warp-unit variant ---------------------------------------------------------------------------------------------
// The block-per-32-rows kernel above launches 256 blocks of 256 threads for 8,192 x 1,024 (two resident per SM at 124
// registers): 0.86 waves, so some SMs carry two blocks and some one, and an SM never holds more than 16 warps (ncu:
// 21 % of the warp slots, 28 % issue utilisation, 20 instructions per element).  Here the unit of work is ONE WARP:
// 16 consecutive rows (one Philox call per 8 columns) x 32 column groups = 256 columns; blocks of two warps, so the same
// layer is 1,024 small blocks that the hardware spreads evenly (6.9 per SM, all resident at once).  The keep mask is
// applied to the packed 16-bit words through a 256-entry table (one LDS + four ANDs per eight elements), ReLU is a max,
// and only the "active" bit costs a compare per element.
constexpr int APW_THREADS = 64;

// APW_SUB rows of loads in flight together (x 2 buffers) at MINB resident blocks per SM: <4, 8> = 16 warps, deep prefetch,
// 128 registers; <2, 12> = 24 warps, shallow prefetch, 80 registers
template <int FMT, int APW_SUB, int MINB>
__global__ void __launch_bounds__(APW_THREADS, MINB) bn_apply_fwd_warp_kernel(BnApplyFwd a, int chunks, int cpw,
                                                                        DropoutLanes lanes, float inv_keep,
                                                                        const PhiloxKeys keys) {
    griddep_launch();
    __shared__ uint4 lut[256];
    for (int b = threadIdx.x; b < 256; b += APW_THREADS) {
        uint32_t w[4];
#pragma unroll
        for (int i = 0; i < 4; ++i)
            w[i] = (((b >> (2 * i)) & 1) ? 0x0000FFFFu : 0u) | (((b >> (2 * i + 1)) & 1) ? 0xFFFF0000u : 0u);
        lut[b] = make_uint4(w[0], w[1], w[2], w[3]);
    }
    __syncthreads();
    griddep_wait();
    const int lane = threadIdx.x & 31;
    const long wu = static_cast<long>(blockIdx.x) * (APW_THREADS / 32) + (threadIdx.x >> 5);
    const int rpw = 32 / cpw;                                     // 16-row groups a warp walks side by side (1 when H >= 256)
    const long rg = wu / chunks;
    const int cg = static_cast<int>(wu - rg * chunks) * cpw + lane % cpw;
    const long g0 = (rg * rpw + lane / cpw) * AP_ROWS;
    if (g0 >= a.rows) return;
    const int tpr = a.H >> 3;
    const long H = a.H;
    const int n = static_cast<int>(min(static_cast<long>(AP_ROWS), a.rows - g0));
    float sc[8], sh[8];
    load8f(a.scale + cg * 8, sc);
    load8f(a.shift + cg * 8, sh);
#pragma unroll
    for (int e = 0; e < 8; ++e) { sc[e] *= inv_keep; sh[e] *= inv_keep; }
    const uint32_t dstep = (a.step_dev != nullptr ? *a.step_dev : a.step) * a.step_mul + a.step_add;
    const bool has_resid = a.resid != nullptr, dropout = a.apply_dropout != 0;
    const uint32_t rmask = (1u << lanes.rows_shift) - 1u;
    const uint16_t* zp = static_cast<const uint16_t*>(a.z) + g0 * H + cg * 8;
    const uint16_t* rp = has_resid ? static_cast<const uint16_t*>(a.resid) + g0 * H + cg * 8 : nullptr;
    uint16_t* ap = static_cast<uint16_t*>(a.a) + g0 * H + cg * 8;
    uint8_t* bp = a.bits + g0 * tpr + cg;
    const uint32_t grow0 = static_cast<uint32_t>(g0 + a.row_offset);
    uint4 zq[2][APW_SUB], rq[2][APW_SUB];
    auto load = [&](int s, uint4 (&zb)[APW_SUB], uint4 (&rb)[APW_SUB]) {
#pragma unroll
        for (int u = 0; u < APW_SUB; ++u) {
            const int r = s * APW_SUB + u;
            zb[u] = make_uint4(0, 0, 0, 0);
            rb[u] = make_uint4(0, 0, 0, 0);
            if (r < n) {
                zb[u] = ldg_stream(zp + r * H);
                if (has_resid) rb[u] = ldg_stream(rp + r * H);
            }
        }
    };
    Philox4 rnd = {0u, 0u, 0u, 0u};
    load(0, zq[0], rq[0]);
#pragma unroll
    for (int s = 0; s < AP_ROWS / APW_SUB; ++s) {
        if ((s + 1) * APW_SUB < n) load(s + 1, zq[(s + 1) & 1], rq[(s + 1) & 1]);
#pragma unroll
        for (int u = 0; u < APW_SUB; ++u) {
            const int r = s * APW_SUB + u;
            if (r >= n) break;
            const uint32_t grow = grow0 + static_cast<uint32_t>(r);
            uint32_t keep = 0xFFu;
            if (dropout) {
                if (r == 0 || (grow & rmask) == 0u)
                    rnd = philox4x32_10(grow >> lanes.rows_shift, static_cast<uint32_t>(cg), a.site, dstep, keys);
                keep = dropout_keep8(rnd, grow, lanes.bits, lanes.threshold);
            }
            float z[8];
            unpack8(zq[s & 1][u], FMT, z);
            uint32_t pos = 0;
#pragma unroll
            for (int e = 0; e < 8; ++e) {
                const float y = fmaf(z[e], sc[e], sh[e]);
                pos |= (y > 0.0f) ? (1u << e) : 0u;
                z[e] = fmaxf(y, 0.0f);
            }
            const uint32_t active = pos & keep;
            if (has_resid) {
                // the dropped elements must be zero BEFORE the residual is added: mask in fp32 only where needed
                float rr[8];
                unpack8(rq[s & 1][u], FMT, rr);
#pragma unroll
                for (int e = 0; e < 8; ++e) z[e] = ((keep >> e) & 1u) ? z[e] + rr[e] : rr[e];
                *reinterpret_cast<uint4*>(ap + r * H) = pack8(z, FMT);
            } else {
                *reinterpret_cast<uint4*>(ap + r * H) = and4(pack8(z, FMT), lut[keep]);
            }
            bp[static_cast<long>(r) * tpr] = static_cast<uint8_t>(active);
        }
    }
}

// ---- the column sums of the GEMM epilogue -> batch statistics -> apply, in ONE launch --------------------------------
// Blocks first share out the second stage of the column reduction (8 columns each: fp64 sums over the GEMM's partial
// rows, SyncBN exchange, scale / shift, running statistics), meet at a grid barrier, then every block applies
// BN + ReLU + dropout to its rows.  Against finalize + apply as two launches this drops one kernel boundary per layer
// from the step's critical path (~3 us each at this size).
constexpr int FRED_LANES = EW_THREADS / RED_COLS;        // 32 row-lanes x 8 columns per block

__device__ __forceinline__ void reduce_partials_block(const float* __restrict__ a, const float* __restrict__ b, int P,
                                                      int n, int ld, int col, int lane_r,
                                                      double (&sh)[2][FRED_LANES][RED_COLS], double& ta, double& tb) {
    double sa = 0.0, sb = 0.0;
    if (col < n) {
        for (int p = lane_r; p < P; p += FRED_LANES) {
            sa += static_cast<double>(ldg_f32(a + static_cast<long>(p) * ld + col));
            sb += static_cast<double>(ldg_f32(b + static_cast<long>(p) * ld + col));
        }
    }
    const int c = threadIdx.x % RED_COLS;
    sh[0][lane_r][c] = sa;
    sh[1][lane_r][c] = sb;
    __syncthreads();
#pragma unroll
    for (int s = FRED_LANES / 2; s >= 1; s >>= 1) {
        if (lane_r < s) {
            sh[0][lane_r][c] += sh[0][lane_r + s][c];
            sh[1][lane_r][c] += sh[1][lane_r + s][c];
        }
        __syncthreads();
    }
    ta = sh[0][0][c];
    tb = sh[1][0][c];
    __syncthreads();                                     // sh is reused by the caller's next column group
}

template <int FMT>
__global__ void __launch_bounds__(EW_THREADS, 2) bn_fwd_fused_kernel(const float* __restrict__ psum,
                                                                  const float* __restrict__ psq, int P, BnFwdFinalize f,
                                                                  double* __restrict__ sums_out, BnApplyFwd a,
                                                                  int rows_per_block, DropoutLanes lanes, float inv_keep,
                                                                  const PhiloxKeys keys, unsigned* __restrict__ barrier) {
    griddep_launch();
    griddep_wait();
    __shared__ double sh[2][FRED_LANES][RED_COLS];
    const int c = threadIdx.x % RED_COLS, lane_r = threadIdx.x / RED_COLS;
    for (int cb = blockIdx.x; cb * RED_COLS < f.H; cb += gridDim.x) {
        const int col = cb * RED_COLS + c;
        BnColumnInputs ci = {0.0f, 1.0f, 0.0f, 0.0f, 1.0f};
        if (lane_r == 0 && col < f.H) ci = bn_column_inputs(f, col);
        double ta, tb;
        reduce_partials_block(psum, psq, P, f.H, f.H, col, lane_r, sh, ta, tb);
        if (f.ex.world > 1 && lane_r < 4) exchange_column_sums_warp(f.ex, f.H, col, col < f.H, lane_r, ta, tb);
        if (lane_r == 0 && col < f.H) {
            sums_out[col] = ta;
            sums_out[f.H + col] = tb;
            if (col == 0 && f.num_batches_tracked != nullptr) *f.num_batches_tracked += 1;
            bn_finalize_fwd_col(f, col, ta, tb, ci);
        }
    }
    grid_barrier_once(barrier, gridDim.x);
    const long base = static_cast<long>(blockIdx.x) * rows_per_block;
    const long end = min(base + rows_per_block, static_cast<long>(a.rows));
    apply_rows<FMT>(a, base, end, lanes, inv_keep, keys, a.scale, a.shift);
}

// ---------------------------------------------------------------------------------- backward passes
// The saved bit per element selects dy = g * inv_keep or 0.  A 256-entry table in shared memory expands a byte of
// eight bits into four 32-bit AND-masks for the four packed 16-bit pairs, so masking costs one LDS + four LOPs
// per eight elements.
__device__ __forceinline__ void fill_mask_lut(uint4* lut) {
    for (int b = threadIdx.x; b < 256; b += blockDim.x) {
        uint32_t w[4];
#pragma unroll
        for (int i = 0; i < 4; ++i)
            w[i] = (((b >> (2 * i)) & 1) ? 0x0000FFFFu : 0u) | (((b >> (2 * i + 1)) & 1) ? 0xFFFF0000u : 0u);
        lut[b] = make_uint4(w[0], w[1], w[2], w[3]);
    }
}

// pass 1: partial column sums of gm = (g masked by the active bit) and of gm * z (raw z).  The finalize turns them
// into sum dy = ik * S1 and sum dy * zhat = ik * rstd * (S2 - mean * S1) in fp64.
template <int FMT>
__global__ void __launch_bounds__(EW_THREADS, 2) bwd_stats_kernel(BwdCommon c, int rows_per_block,
                                                               float* __restrict__ psum0, float* __restrict__ psum1) {
    griddep_launch();
    __shared__ uint4 lut[256];
    __shared__ float sh[2][EW_THREADS][8];
    fill_mask_lut(lut);
    __syncthreads();
    griddep_wait();
    const int tpr = c.H >> 3;
    const int rpb = EW_THREADS / tpr;
    const int cg = threadIdx.x % tpr, rl = threadIdx.x / tpr;
    const long base = static_cast<long>(blockIdx.x) * rows_per_block;
    const long end = min(base + rows_per_block, static_cast<long>(c.rows));
    const uint16_t* gp = static_cast<const uint16_t*>(c.g);
    const uint16_t* zp = static_cast<const uint16_t*>(c.z);
    float acc0[8] = {0, 0, 0, 0, 0, 0, 0, 0}, acc1[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    const long step = static_cast<long>(rpb) * EW_UNROLL;
    auto load = [&](long r0, uint4 (&gq)[EW_UNROLL], uint4 (&zq)[EW_UNROLL], uint32_t (&act)[EW_UNROLL]) {
#pragma unroll
        for (int u = 0; u < EW_UNROLL; ++u) {
            const long row = r0 + static_cast<long>(u) * rpb;
            gq[u] = make_uint4(0, 0, 0, 0);
            zq[u] = make_uint4(0, 0, 0, 0);
            act[u] = 0;
            if (row < end) {
                const long off = row * c.H + cg * 8;
                gq[u] = ldg_stream(gp + off);
                zq[u] = ldg_stream(zp + off);
                act[u] = c.bits[row * tpr + cg];
            }
        }
    };
    auto compute = [&](const uint4 (&gq)[EW_UNROLL], const uint4 (&zq)[EW_UNROLL], const uint32_t (&act)[EW_UNROLL]) {
#pragma unroll
        for (int u = 0; u < EW_UNROLL; ++u) {                       // rows past `end` carry act == 0: they add nothing
            float g[8], z[8];
            unpack8(and4(gq[u], lut[act[u]]), FMT, g);
            unpack8(zq[u], FMT, z);
#pragma unroll
            for (int e = 0; e < 8; ++e) {
                acc0[e] += g[e];
                acc1[e] = fmaf(g[e], z[e], acc1[e]);
            }
        }
    };
    uint4 g0[EW_UNROLL], z0[EW_UNROLL], g1[EW_UNROLL], z1[EW_UNROLL];
    uint32_t a0[EW_UNROLL], a1[EW_UNROLL];
    long r = base + rl;
    if (r < end) load(r, g0, z0, a0);
    while (r < end) {
        if (r + step < end) load(r + step, g1, z1, a1);
        compute(g0, z0, a0);
        r += step;
        if (r >= end) break;
        if (r + step < end) load(r + step, g0, z0, a0);
        compute(g1, z1, a1);
        r += step;
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
        float* o1 = psum1 + static_cast<long>(blockIdx.x) * c.H + cg * 8;
        *reinterpret_cast<float4*>(o0) = make_float4(t0[0], t0[1], t0[2], t0[3]);
        *reinterpret_cast<float4*>(o0 + 4) = make_float4(t0[4], t0[5], t0[6], t0[7]);
        *reinterpret_cast<float4*>(o1) = make_float4(t1[0], t1[1], t1[2], t1[3]);
        *reinterpret_cast<float4*>(o1 + 4) = make_float4(t1[4], t1[5], t1[6], t1[7]);
    }
}

// pass 2: dz = A * gm + Bz * z + Cc per column (coef = [3][H], see bn_bwd_coefficients), 16-bit output.
// WITH_BIAS also leaves partial column sums of the stored dz (bias gradient when BatchNorm does not cancel it).
template <bool WITH_BIAS, int FMT>
__global__ void __launch_bounds__(EW_THREADS, 2) bwd_dz_kernel(BwdCommon c, int rows_per_block,
                                                            const float* __restrict__ coef, uint16_t* __restrict__ dz,
                                                            float* __restrict__ psum) {
    griddep_launch();
    __shared__ uint4 lut[256];
    __shared__ float sh[WITH_BIAS ? EW_THREADS : 1][8];
    fill_mask_lut(lut);
    __syncthreads();
    griddep_wait();
    const int tpr = c.H >> 3;
    const int rpb = EW_THREADS / tpr;
    const int cg = threadIdx.x % tpr, rl = threadIdx.x / tpr;
    const long base = static_cast<long>(blockIdx.x) * rows_per_block;
    const long end = min(base + rows_per_block, static_cast<long>(c.rows));
    float ka[8], kz[8], kc[8];
    load8f(coef + cg * 8, ka);
    load8f(coef + c.H + cg * 8, kz);
    load8f(coef + 2 * c.H + cg * 8, kc);
    const uint16_t* gp = static_cast<const uint16_t*>(c.g);
    const uint16_t* zp = static_cast<const uint16_t*>(c.z);
    float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    const long step = static_cast<long>(rpb) * EW_UNROLL;
    auto load = [&](long r0, uint4 (&gq)[EW_UNROLL], uint4 (&zq)[EW_UNROLL], uint32_t (&act)[EW_UNROLL]) {
#pragma unroll
        for (int u = 0; u < EW_UNROLL; ++u) {
            const long row = r0 + static_cast<long>(u) * rpb;
            gq[u] = make_uint4(0, 0, 0, 0);
            zq[u] = make_uint4(0, 0, 0, 0);
            act[u] = 0;
            if (row < end) {
                const long off = row * c.H + cg * 8;
                gq[u] = ldg_stream(gp + off);
                zq[u] = ldg_stream(zp + off);
                act[u] = c.bits[row * tpr + cg];
            }
        }
    };
    auto compute = [&](long r0, const uint4 (&gq)[EW_UNROLL], const uint4 (&zq)[EW_UNROLL],
                       const uint32_t (&act)[EW_UNROLL]) {
#pragma unroll
        for (int u = 0; u < EW_UNROLL; ++u) {
            const long row = r0 + static_cast<long>(u) * rpb;
            if (row >= end) break;
            float g[8], z[8], o[8];
            unpack8(and4(gq[u], lut[act[u]]), FMT, g);
            unpack8(zq[u], FMT, z);
#pragma unroll
            for (int e = 0; e < 8; ++e) o[e] = fmaf(ka[e], g[e], fmaf(kz[e], z[e], kc[e]));
            const uint4 q = pack8(o, FMT);
            *reinterpret_cast<uint4*>(dz + row * c.H + cg * 8) = q;
            if (WITH_BIAS) {
                float rr[8];
                unpack8(q, FMT, rr);                             // bias gradient of the values the GEMMs consume
#pragma unroll
                for (int e = 0; e < 8; ++e) acc[e] += rr[e];
            }
        }
    };
    uint4 g0[EW_UNROLL], z0[EW_UNROLL], g1[EW_UNROLL], z1[EW_UNROLL];
    uint32_t a0[EW_UNROLL], a1[EW_UNROLL];
    long r = base + rl;
    if (r < end) load(r, g0, z0, a0);
    while (r < end) {
        if (r + step < end) load(r + step, g1, z1, a1);
        compute(r, g0, z0, a0);
        r += step;
        if (r >= end) break;
        if (r + step < end) load(r + step, g0, z0, a0);
        compute(r, g1, z1, a1);
        r += step;
    }
    if (WITH_BIAS) {
#pragma unroll
        for (int e = 0; e < 8; ++e) sh[threadIdx.x][e] = acc[e];
        __syncthreads();
        if (rl == 0) {
            float t[8] = {0, 0, 0, 0, 0, 0, 0, 0};
            for (int r = 0; r < rpb; ++r)
#pragma unroll
                for (int e = 0; e < 8; ++e) t[e] += sh[r * tpr + cg][e];
            float* o = psum + static_cast<long>(blockIdx.x) * c.H + cg * 8;
            *reinterpret_cast<float4*>(o) = make_float4(t[0], t[1], t[2], t[3]);
            *reinterpret_cast<float4*>(o + 4) = make_float4(t[4], t[5], t[6], t[7]);
        }
    }
}

// ---------------------------------------------------------------------------------- BN finalize (backward)
// coefficients of pass 2 from the GLOBAL sums:  dz = k (dy - c1 - zhat c2),  dy = ik gm,  zhat = (z - mean) rstd
//   =>  dz = A gm + Bz z + Cc   with  A = k ik,  Bz = -k c2 rstd,  Cc = -k c1 - Bz mean
__global__ void bn_finalize_bwd_kernel(BnBwdFinalize a) {
    griddep_launch();
    griddep_wait();
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= a.H) return;
    const bool batch = a.use_bn && a.bn_mode == 0;
    bn_bwd_coefficients(a, j, batch ? a.sums[j] : 0.0, batch ? a.sums[a.H + j] : 0.0, bn_bwd_column(a, j));
}

// second stage of pass 1: LOCAL sums of dy and dy * zhat (fp64, what a data-parallel run all-reduces) and the
// BatchNorm parameter gradients from them; FUSED also writes the pass-2 coefficients (single GPU: local == global)
template <bool FUSED>
__global__ void __launch_bounds__(RED_COLS * RED_LANES, 1) bn_bwd_reduce_kernel(const float* __restrict__ pg, const float* __restrict__ pgz, int P,
                                     BnBwdFinalize a, double* __restrict__ sums_out, float* __restrict__ dgamma,
                                     float* __restrict__ dbeta) {
    griddep_launch();
    griddep_wait();
    __shared__ double sh[2][RED_LANES][RED_COLS];
    double s1, s2;
    const int col = blockIdx.x * RED_COLS + threadIdx.x;
    BnBwdColumn cv = {1.0f, 0.0f, 1.0f};
    if (threadIdx.y == 0 && col < a.H) cv = bn_bwd_column(a, col);
    reduce_partials_2(pg, pgz, P, a.H, a.H, sh, s1, s2);
    double sum_dy = 0.0, sum_dyz = 0.0;
    if (threadIdx.y == 0 && col < a.H) {
        const double ik = static_cast<double>(a.inv_keep);
        sum_dy = ik * s1;
        sum_dyz = ik * static_cast<double>(cv.rstd) * (s2 - static_cast<double>(cv.mean) * s1);
        dbeta[col] = static_cast<float>(sum_dy);                            // parameter gradients: LOCAL sums
        dgamma[col] = static_cast<float>(sum_dyz);
    }
    if (FUSED && a.ex.world > 1 && threadIdx.y < 4) exchange_column_sums_warp(a.ex, a.H, col, col < a.H, threadIdx.y, sum_dy, sum_dyz);
    if (threadIdx.y == 0 && col < a.H) {
        sums_out[col] = sum_dy;
        sums_out[a.H + col] = sum_dyz;
        if (FUSED) bn_bwd_coefficients(a, col, sum_dy, sum_dyz, cv);
    }
}

// ---- BatchNorm backward of one layer in ONE launch: pass 1 -> column reduction -> coefficients -> pass 2 ----------------
// The three launches it replaces (bwd_stats, bn_bwd_reduce<fused>, bwd_dz) are serial links of the step's critical
// path; here they are phases of one grid separated by two grid barriers.  Phase 1: every block leaves one partial row of
// sum gm and sum gm z over its rows.  Phase 2: the blocks share out the columns (8 each): fp64 sums over the partial
// rows, dgamma / dbeta, SyncBN exchange, the pass-2 coefficients.  Phase 3: dz = A gm + Bz z + Cc over the same rows --
// the second read of (g, z, bits) comes from L2, where phase 1 just left them.
template <int FMT>
__global__ void __launch_bounds__(EW_THREADS, 2) bn_bwd_fused_kernel(BwdCommon c, int rows_per_block,
                                                                  float* __restrict__ psum0, float* __restrict__ psum1,
                                                                  BnBwdFinalize f, double* __restrict__ sums_out,
                                                                  float* __restrict__ dgamma, float* __restrict__ dbeta,
                                                                  uint16_t* __restrict__ dz, unsigned* __restrict__ barrier) {
    griddep_launch();
    __shared__ uint4 lut[256];
    __shared__ union { float part[2][EW_THREADS][8]; double red[2][FRED_LANES][RED_COLS]; } sm;
    fill_mask_lut(lut);
    __syncthreads();
    griddep_wait();
    const int tpr = c.H >> 3;
    const int rpb = EW_THREADS / tpr;
    const int cg = threadIdx.x % tpr, rl = threadIdx.x / tpr;
    const long base = static_cast<long>(blockIdx.x) * rows_per_block;
    const long end = min(base + rows_per_block, static_cast<long>(c.rows));
    const uint16_t* gp = static_cast<const uint16_t*>(c.g);
    const uint16_t* zp = static_cast<const uint16_t*>(c.z);
    const long step = static_cast<long>(rpb) * EW_UNROLL;
    auto load = [&](long r0, uint4 (&gq)[EW_UNROLL], uint4 (&zq)[EW_UNROLL], uint32_t (&act)[EW_UNROLL]) {
#pragma unroll
        for (int u = 0; u < EW_UNROLL; ++u) {
            const long row = r0 + static_cast<long>(u) * rpb;
            gq[u] = make_uint4(0, 0, 0, 0);
            zq[u] = make_uint4(0, 0, 0, 0);
            act[u] = 0;
            if (row < end) {
                const long off = row * c.H + cg * 8;
                gq[u] = ldg_stream(gp + off);
                zq[u] = ldg_stream(zp + off);
                act[u] = c.bits[row * tpr + cg];
            }
        }
    };
    uint4 g0[EW_UNROLL], z0[EW_UNROLL], g1[EW_UNROLL], z1[EW_UNROLL];
    uint32_t a0[EW_UNROLL], a1[EW_UNROLL];

    // ---------------- phase 1: partial column sums
    {
        float acc0[8] = {0, 0, 0, 0, 0, 0, 0, 0}, acc1[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        auto compute = [&](const uint4 (&gq)[EW_UNROLL], const uint4 (&zq)[EW_UNROLL], const uint32_t (&act)[EW_UNROLL]) {
#pragma unroll
            for (int u = 0; u < EW_UNROLL; ++u) {                   // rows past `end` carry act == 0: they add nothing
                float g[8], z[8];
                unpack8(and4(gq[u], lut[act[u]]), FMT, g);
                unpack8(zq[u], FMT, z);
#pragma unroll
                for (int e = 0; e < 8; ++e) {
                    acc0[e] += g[e];
                    acc1[e] = fmaf(g[e], z[e], acc1[e]);
                }
            }
        };
        long r = base + rl;
        if (r < end) load(r, g0, z0, a0);
        while (r < end) {
            if (r + step < end) load(r + step, g1, z1, a1);
            compute(g0, z0, a0);
            r += step;
            if (r >= end) break;
            if (r + step < end) load(r + step, g0, z0, a0);
            compute(g1, z1, a1);
            r += step;
        }
#pragma unroll
        for (int e = 0; e < 8; ++e) {
            sm.part[0][threadIdx.x][e] = acc0[e];
            sm.part[1][threadIdx.x][e] = acc1[e];
        }
        __syncthreads();
        if (rl == 0) {
            float t0[8] = {0, 0, 0, 0, 0, 0, 0, 0}, t1[8] = {0, 0, 0, 0, 0, 0, 0, 0};
            for (int q = 0; q < rpb; ++q) {
#pragma unroll
                for (int e = 0; e < 8; ++e) {
                    t0[e] += sm.part[0][q * tpr + cg][e];
                    t1[e] += sm.part[1][q * tpr + cg][e];
                }
            }
            float* o0 = psum0 + static_cast<long>(blockIdx.x) * c.H + cg * 8;
            float* o1 = psum1 + static_cast<long>(blockIdx.x) * c.H + cg * 8;
            *reinterpret_cast<float4*>(o0) = make_float4(t0[0], t0[1], t0[2], t0[3]);
            *reinterpret_cast<float4*>(o0 + 4) = make_float4(t0[4], t0[5], t0[6], t0[7]);
            *reinterpret_cast<float4*>(o1) = make_float4(t1[0], t1[1], t1[2], t1[3]);
            *reinterpret_cast<float4*>(o1 + 4) = make_float4(t1[4], t1[5], t1[6], t1[7]);
        }
    }
    grid_barrier_once(barrier, gridDim.x);

    // ---------------- phase 2: column totals, parameter gradients, exchange, coefficients
    {
        const int cc = threadIdx.x % RED_COLS, lane_r = threadIdx.x / RED_COLS;
        for (int cb = blockIdx.x; cb * RED_COLS < f.H; cb += gridDim.x) {
            const int col = cb * RED_COLS + cc;
            BnBwdColumn cv = {1.0f, 0.0f, 1.0f};
            if (lane_r == 0 && col < f.H) cv = bn_bwd_column(f, col);
            double s1, s2;
            reduce_partials_block(psum0, psum1, static_cast<int>(gridDim.x), f.H, f.H, col, lane_r, sm.red, s1, s2);
            double sum_dy = 0.0, sum_dyz = 0.0;
            if (lane_r == 0 && col < f.H) {
                const double ik = static_cast<double>(f.inv_keep);
                sum_dy = ik * s1;
                sum_dyz = ik * static_cast<double>(cv.rstd) * (s2 - static_cast<double>(cv.mean) * s1);
                dbeta[col] = static_cast<float>(sum_dy);                        // parameter gradients: LOCAL sums
                dgamma[col] = static_cast<float>(sum_dyz);
            }
            if (f.ex.world > 1 && lane_r < 4) exchange_column_sums_warp(f.ex, f.H, col, col < f.H, lane_r, sum_dy, sum_dyz);
            if (lane_r == 0 && col < f.H) {
                sums_out[col] = sum_dy;
                sums_out[f.H + col] = sum_dyz;
                bn_bwd_coefficients(f, col, sum_dy, sum_dyz, cv);
            }
        }
    }
    grid_barrier_once(barrier + 1, gridDim.x);

    // ---------------- phase 3: dz
    {
        float ka[8], kz[8], kc[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) {                               // written by other blocks of this launch: past L1
            ka[e] = __ldcg(f.coef + cg * 8 + e);
            kz[e] = __ldcg(f.coef + c.H + cg * 8 + e);
            kc[e] = __ldcg(f.coef + 2 * c.H + cg * 8 + e);
        }
        auto compute = [&](long r0, const uint4 (&gq)[EW_UNROLL], const uint4 (&zq)[EW_UNROLL],
                           const uint32_t (&act)[EW_UNROLL]) {
#pragma unroll
            for (int u = 0; u < EW_UNROLL; ++u) {
                const long row = r0 + static_cast<long>(u) * rpb;
                if (row >= end) break;
                float g[8], z[8], o[8];
                unpack8(and4(gq[u], lut[act[u]]), FMT, g);
                unpack8(zq[u], FMT, z);
#pragma unroll
                for (int e = 0; e < 8; ++e) o[e] = fmaf(ka[e], g[e], fmaf(kz[e], z[e], kc[e]));
                *reinterpret_cast<uint4*>(dz + row * c.H + cg * 8) = pack8(o, FMT);
            }
        };
        long r = base + rl;
        if (r < end) load(r, g0, z0, a0);
        while (r < end) {
            if (r + step < end) load(r + step, g1, z1, a1);
            compute(r, g0, z0, a0);
            r += step;
            if (r >= end) break;
            if (r + step < end) load(r + step, g0, z0, a0);
            compute(r, g1, z1, a1);
            r += step;
        }
    }
}

// ---------------------------------------------------------------------------------- gradient packing
// sum of the split-K slabs -> dense fp32 gradient [rows][cols]; fixed order, four columns per thread when the
// shapes allow 16-byte accesses
__global__ void wgrad_reduce_kernel(const float* __restrict__ slabs, int splits, long slab_elems, int ld, int rows,
                                    int cols, float* __restrict__ grad) {
    griddep_launch();
    griddep_wait();
    const long idx = static_cast<long>(blockIdx.x) * blockDim.x + threadIdx.x;
    if ((cols & 3) == 0 && (ld & 3) == 0) {
        const int c4 = cols >> 2;
        if (idx >= static_cast<long>(rows) * c4) return;
        const int r = static_cast<int>(idx / c4), c = static_cast<int>(idx - static_cast<long>(r) * c4) * 4;
        const long src = static_cast<long>(r) * ld + c;
        float4 s = *reinterpret_cast<const float4*>(slabs + src);
        for (int k = 1; k < splits; ++k) {
            const float4 v = *reinterpret_cast<const float4*>(slabs + k * slab_elems + src);
            s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
        }
        *reinterpret_cast<float4*>(grad + static_cast<long>(r) * cols + c) = s;
        return;
    }
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
                                                                   float* __restrict__ psum,
                                                                   unsigned* __restrict__ zero_words, int n_zero) {
    griddep_launch();
    griddep_wait();
    if (blockIdx.x == 0 && static_cast<int>(threadIdx.x) < n_zero) zero_words[threadIdx.x] = 0u;   // the pass's barrier words
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

// rows owned by one block of a column-reducing kernel: a multiple of the rows one unrolled iteration covers, sized
// so that at most 2 x SM-count blocks (= partial rows) exist
static int rows_per_block_for(int64_t rows, int H, int blocks_per_sm, bool whole_groups = true) {
    const int64_t rpb = EW_THREADS / (H / 8);
    // whole_groups: a multiple of the rows one unrolled iteration covers (the unchecked fast path of the apply kernel);
    // otherwise a multiple of the rows of ONE iteration, which fills the resident block slots more evenly
    const int64_t step = whole_groups ? rpb * EW_UNROLL : rpb;
    const int64_t steps = (rows + step - 1) / step;
    const int64_t max_blocks = static_cast<int64_t>(num_sms()) * blocks_per_sm;
    const int64_t steps_per_block = (steps + max_blocks - 1) / max_blocks;
    return static_cast<int>(steps_per_block * step);
}
int ew_rows_per_block(int64_t rows, int H) { return rows_per_block_for(rows, H, 2, false); }
int ew_partial_rows(int64_t rows, int H) {
    const int rpb = ew_rows_per_block(rows, H);
    return static_cast<int>((rows + rpb - 1) / rpb);
}

int reduce_rows_f64(const float* a, const float* b, int P, int n, int ld, double* out_a, double* out_b, cudaStream_t st) {
    LIFTER_CHECK_CUDA(launch_pdl("reduce_rows_kernel<double>", reduce_rows_kernel<double>, dim3(ceil_div(n, RED_COLS)), dim3(RED_COLS, RED_LANES), 0, st, a, b, P, n, ld, out_a, out_b));
    return LIFTER_OK;
}
int reduce_rows_f32(const float* a, int P, int n, int ld, float* out, cudaStream_t st) {
    LIFTER_CHECK_CUDA(launch_pdl("reduce_rows_kernel<float>", reduce_rows_kernel<float>, dim3(ceil_div(n, RED_COLS)), dim3(RED_COLS, RED_LANES), 0, st, a, nullptr, P, n, ld, out, nullptr));
    return LIFTER_OK;
}

int bn_finalize_fwd(const BnFwdFinalize& a, cudaStream_t st) {
    LIFTER_CHECK_CUDA(launch_pdl("bn_finalize_fwd_kernel", bn_finalize_fwd_kernel, dim3(ceil_div(a.H, 256)), dim3(256), 0, st, a));
    return LIFTER_OK;
}

int bn_reduce_finalize_fwd(const float* psum, const float* psq, int P, const BnFwdFinalize& a, double* sums_out,
                           cudaStream_t st) {
    LIFTER_CHECK_CUDA(launch_pdl("bn_reduce_finalize_fwd_kernel", bn_reduce_finalize_fwd_kernel, dim3(ceil_div(a.H, RED_COLS)), dim3(RED_COLS, RED_LANES), 0, st, psum, psq, P, a,
                                 sums_out));
    return LIFTER_OK;
}

static int check_h(int H, const char* what) {
    LIFTER_REQUIRE(H >= 64 && H <= 2048 && (H & (H - 1)) == 0, "%s: unsupported H=%d", what, H);
    return LIFTER_OK;
}

int bn_apply_fwd(const BnApplyFwd& a, cudaStream_t st) {
    LIFTER_PROPAGATE(check_h(a.H, "bn_apply"));
    const DropoutLanes lanes = dropout_lanes(a.p);
    const PhiloxKeys keys = philox_round_keys(a.seed);
    const float inv_keep = (a.apply_dropout && a.p < 1.0f) ? 1.0f / (1.0f - a.p) : (a.apply_dropout ? 0.0f : 1.0f);
    static const bool block_variant = [] { const char* e = getenv("LIFTER_APPLY_BLOCK"); return e != nullptr && e[0] == '1'; }();
    if (!block_variant) {                         // one warp per 16 rows x 256 columns, two warps per block
        const int tpr = a.H / 8;
        const int cpw = tpr < 32 ? tpr : 32, chunks = tpr / cpw, rpw = 32 / cpw;
        const int64_t units = ((a.rows + static_cast<int64_t>(rpw) * AP_ROWS - 1) / (static_cast<int64_t>(rpw) * AP_ROWS)) * chunks;
        const unsigned blocks = static_cast<unsigned>((units + APW_THREADS / 32 - 1) / (APW_THREADS / 32));
        static const bool shallow = [] { const char* e = getenv("LIFTER_APPLY_SHALLOW"); return e != nullptr && e[0] == '1'; }();
        auto kern = a.fmt == FMT_BF16 ? (shallow ? bn_apply_fwd_warp_kernel<FMT_BF16, 2, 12> : bn_apply_fwd_warp_kernel<FMT_BF16, 4, 8>)
                                      : (shallow ? bn_apply_fwd_warp_kernel<FMT_F16, 2, 12> : bn_apply_fwd_warp_kernel<FMT_F16, 4, 8>);
        LIFTER_CHECK_CUDA(launch_pdl("bn_apply_fwd_warp_kernel", kern, dim3(blocks), dim3(APW_THREADS), 0, st, a, chunks, cpw, lanes, inv_keep, keys));
        return LIFTER_OK;
    }
    // (LIFTER_APPLY_BLOCK=1: the block-per-32-rows kernel, kept for A/B timing)
    // a block owns whole 16-row groups; up to four blocks per SM's worth of them, filled evenly
    const int64_t group = static_cast<int64_t>(EW_THREADS / (a.H / 8)) * AP_ROWS;
    const int64_t groups = (a.rows + group - 1) / group;
    const int64_t max_blocks = static_cast<int64_t>(num_sms()) * 4;
    const int64_t per_block = (groups + max_blocks - 1) / max_blocks;
    const int rpb = static_cast<int>(per_block * group);
    const unsigned blocks = static_cast<unsigned>((a.rows + rpb - 1) / rpb);
    if (a.fmt == FMT_BF16)
        LIFTER_CHECK_CUDA(launch_pdl("bn_apply_fwd_kernel", bn_apply_fwd_kernel<FMT_BF16>, dim3(blocks), dim3(EW_THREADS), 0, st, a, rpb, lanes, inv_keep, keys));
    else
        LIFTER_CHECK_CUDA(launch_pdl("bn_apply_fwd_kernel", bn_apply_fwd_kernel<FMT_F16>, dim3(blocks), dim3(EW_THREADS), 0, st, a, rpb, lanes, inv_keep, keys));
    return LIFTER_OK;
}

// Blocks of a grid-barrier kernel that fit on the device at once (occupancy query, cached per device and kernel).
template <typename K>
static int resident_blocks(K kernel, int slot) {
    static int cached[2][64] = {};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return num_sms();
    if (cached[slot][dev] == 0) {
        int per_sm = 0;
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, EW_THREADS, 0) != cudaSuccess || per_sm < 1) per_sm = 1;
        if (per_sm > 2) per_sm = 2;
        cached[slot][dev] = per_sm * num_sms();
    }
    return cached[slot][dev];
}

// rows per block of a barrier kernel: whole `group`-row units, at most `max_blocks` blocks, every block non-empty
static int barrier_rows_per_block(int64_t rows, int64_t group, int max_blocks) {
    const int64_t groups = (rows + group - 1) / group;
    const int64_t per_block = (groups + max_blocks - 1) / max_blocks;
    return static_cast<int>(per_block * group);
}

int bn_fwd_fused(const float* psum, const float* psq, int P, const BnFwdFinalize& f, double* sums_out,
                 const BnApplyFwd& a, unsigned* barrier, cudaStream_t st) {
    LIFTER_PROPAGATE(check_h(a.H, "bn_fwd_fused"));
    LIFTER_REQUIRE(barrier != nullptr, "bn_fwd_fused: null barrier word");
    const DropoutLanes lanes = dropout_lanes(a.p);
    const PhiloxKeys keys = philox_round_keys(a.seed);
    const float inv_keep = (a.apply_dropout && a.p < 1.0f) ? 1.0f / (1.0f - a.p) : (a.apply_dropout ? 0.0f : 1.0f);
    const int64_t group = static_cast<int64_t>(EW_THREADS / (a.H / 8)) * AP_ROWS;
    if (a.fmt == FMT_BF16) {
        const int rpb = barrier_rows_per_block(a.rows, group, resident_blocks(bn_fwd_fused_kernel<FMT_BF16>, 0));
        const unsigned blocks = static_cast<unsigned>((a.rows + rpb - 1) / rpb);
        LIFTER_CHECK_CUDA(launch_pdl("bn_fwd_fused_kernel", bn_fwd_fused_kernel<FMT_BF16>, dim3(blocks), dim3(EW_THREADS), 0, st,
                                     psum, psq, P, f, sums_out, a, rpb, lanes, inv_keep, keys, barrier));
    } else {
        const int rpb = barrier_rows_per_block(a.rows, group, resident_blocks(bn_fwd_fused_kernel<FMT_F16>, 1));
        const unsigned blocks = static_cast<unsigned>((a.rows + rpb - 1) / rpb);
        LIFTER_CHECK_CUDA(launch_pdl("bn_fwd_fused_kernel", bn_fwd_fused_kernel<FMT_F16>, dim3(blocks), dim3(EW_THREADS), 0, st,
                                     psum, psq, P, f, sums_out, a, rpb, lanes, inv_keep, keys, barrier));
    }
    return LIFTER_OK;
}

int bn_bwd_fused_partial_rows(int64_t rows, int H) {
    // upper bound over the two operand formats (the scratch buffers are sized before the format is known)
    const int cap = 2 * num_sms();
    const int64_t group = EW_THREADS / (H / 8);
    const int rpb = barrier_rows_per_block(rows, group, cap);
    return static_cast<int>((rows + rpb - 1) / rpb);
}

int bn_bwd_fused(const BwdCommon& c, float* psum_g, float* psum_gz, const BnBwdFinalize& f, double* sums_out,
                 float* dgamma, float* dbeta, void* dz, unsigned* barrier, cudaStream_t st) {
    LIFTER_PROPAGATE(check_h(c.H, "bn_bwd_fused"));
    LIFTER_REQUIRE(barrier != nullptr, "bn_bwd_fused: null barrier words");
    const int64_t group = EW_THREADS / (c.H / 8);
    uint16_t* d = static_cast<uint16_t*>(dz);
    if (c.fmt == FMT_BF16) {
        const int rpb = barrier_rows_per_block(c.rows, group, resident_blocks(bn_bwd_fused_kernel<FMT_BF16>, 0));
        const unsigned blocks = static_cast<unsigned>((c.rows + rpb - 1) / rpb);
        LIFTER_CHECK_CUDA(launch_pdl("bn_bwd_fused_kernel", bn_bwd_fused_kernel<FMT_BF16>, dim3(blocks), dim3(EW_THREADS), 0, st,
                                     c, rpb, psum_g, psum_gz, f, sums_out, dgamma, dbeta, d, barrier));
    } else {
        const int rpb = barrier_rows_per_block(c.rows, group, resident_blocks(bn_bwd_fused_kernel<FMT_F16>, 1));
        const unsigned blocks = static_cast<unsigned>((c.rows + rpb - 1) / rpb);
        LIFTER_CHECK_CUDA(launch_pdl("bn_bwd_fused_kernel", bn_bwd_fused_kernel<FMT_F16>, dim3(blocks), dim3(EW_THREADS), 0, st,
                                     c, rpb, psum_g, psum_gz, f, sums_out, dgamma, dbeta, d, barrier));
    }
    return LIFTER_OK;
}

int bwd_stats(const BwdCommon& c, float* psum_g, float* psum_gz, cudaStream_t st) {
    LIFTER_PROPAGATE(check_h(c.H, "bwd_stats"));
    const dim3 grid(ew_partial_rows(c.rows, c.H));
    const int rpb = ew_rows_per_block(c.rows, c.H);
    if (c.fmt == FMT_BF16)
        LIFTER_CHECK_CUDA(launch_pdl("bwd_stats_kernel", bwd_stats_kernel<FMT_BF16>, grid, dim3(EW_THREADS), 0, st, c, rpb, psum_g, psum_gz));
    else
        LIFTER_CHECK_CUDA(launch_pdl("bwd_stats_kernel", bwd_stats_kernel<FMT_F16>, grid, dim3(EW_THREADS), 0, st, c, rpb, psum_g, psum_gz));
    return LIFTER_OK;
}

int bn_finalize_bwd(const BnBwdFinalize& a, cudaStream_t st) {
    LIFTER_CHECK_CUDA(launch_pdl("bn_finalize_bwd_kernel", bn_finalize_bwd_kernel, dim3(ceil_div(a.H, 256)), dim3(256), 0, st, a));
    return LIFTER_OK;
}

int bn_bwd_reduce(const float* pg, const float* pgz, int P, const BnBwdFinalize& a, double* sums_out, float* dgamma,
                  float* dbeta, bool fused, cudaStream_t st) {
    if (fused)
        LIFTER_CHECK_CUDA(launch_pdl("bn_bwd_reduce_kernel<fused>", bn_bwd_reduce_kernel<true>, dim3(ceil_div(a.H, RED_COLS)), dim3(RED_COLS, RED_LANES), 0, st, pg, pgz, P, a, sums_out, dgamma, dbeta));
    else
        LIFTER_CHECK_CUDA(launch_pdl("bn_bwd_reduce_kernel", bn_bwd_reduce_kernel<false>, dim3(ceil_div(a.H, RED_COLS)), dim3(RED_COLS, RED_LANES), 0, st, pg, pgz, P, a, sums_out, dgamma, dbeta));
    return LIFTER_OK;
}

int bwd_dz(const BwdCommon& c, const float* coef, void* dz, float* psum_dz, cudaStream_t st) {
    LIFTER_PROPAGATE(check_h(c.H, "bwd_dz"));
    const dim3 grid(ew_partial_rows(c.rows, c.H));
    const int rpb = ew_rows_per_block(c.rows, c.H);
    uint16_t* d = static_cast<uint16_t*>(dz);
    if (psum_dz != nullptr) {
        if (c.fmt == FMT_BF16)
            LIFTER_CHECK_CUDA(launch_pdl("bwd_dz_kernel<bias>", bwd_dz_kernel<true, FMT_BF16>, grid, dim3(EW_THREADS), 0, st, c, rpb, coef, d, psum_dz));
        else
            LIFTER_CHECK_CUDA(launch_pdl("bwd_dz_kernel<bias>", bwd_dz_kernel<true, FMT_F16>, grid, dim3(EW_THREADS), 0, st, c, rpb, coef, d, psum_dz));
    } else {
        if (c.fmt == FMT_BF16)
            LIFTER_CHECK_CUDA(launch_pdl("bwd_dz_kernel", bwd_dz_kernel<false, FMT_BF16>, grid, dim3(EW_THREADS), 0, st, c, rpb, coef, d, psum_dz));
        else
            LIFTER_CHECK_CUDA(launch_pdl("bwd_dz_kernel", bwd_dz_kernel<false, FMT_F16>, grid, dim3(EW_THREADS), 0, st, c, rpb, coef, d, psum_dz));
    }
    return LIFTER_OK;
}

int wgrad_reduce(const float* slabs, int splits, int slab_rows, int ld, int rows, int cols, float* grad,
                 cudaStream_t st) {
    const bool vec = (cols & 3) == 0 && (ld & 3) == 0;
    const long n = static_cast<long>(rows) * (vec ? cols / 4 : cols);
    LIFTER_CHECK_CUDA(launch_pdl("wgrad_reduce_kernel", wgrad_reduce_kernel, dim3(static_cast<unsigned>((n + 255) / 256)), dim3(256), 0, st,
        slabs, splits, static_cast<long>(slab_rows) * ld, ld, rows, cols, grad));
    return LIFTER_OK;
}

int prep_grad_out(const float* dout, int64_t rows, int O, int Op, int fmt, void* d16, float* psum, cudaStream_t st,
                  unsigned* zero_words, int n_zero) {
    LIFTER_REQUIRE(Op == 64, "prep_grad_out: padded output width must be 64");
    LIFTER_CHECK_CUDA(launch_pdl("prep_grad_out_kernel", prep_grad_out_kernel, dim3(ew_partial_rows(rows, Op)), dim3(EW_THREADS), 0, st, dout, rows, O, Op, fmt,
                                 ew_rows_per_block(rows, Op),
                                                                       static_cast<uint16_t*>(d16), psum, zero_words,
                                                                       zero_words != nullptr ? n_zero : 0));
    return LIFTER_OK;
}

}  // namespace lifter
