// MyViT lifter, training path (phase1_lifting/baselineModel.py:228-362 under loss.backward(),
// the loop body train_1.py:73-100 with model_lift = MyViT).  The transformer has no dropout and no BatchNorm, so the
// train-mode forward is the eval arithmetic with every GEMM input kept for backward; backward re-uses the tcgen05 GEMM
// for all data gradients (dX = dZ W: MN-major B) and weight gradients (dW = dZ^T X: both operands MN-major, split-K
// slabs) and adds the kernels between them:
//   vit_gelu_fwd / vit_gelu_bwd     m = GELU(z1);  dz1 = dm * GELU'(z1)   (z1 is saved: GELU is not invertible)
//   vit_relu_bwd                    head: dz = dr * (r > 0)
//   vit_ln_bwd                      one or two LayerNorms back to back, + the residual gradient, + d gain / d bias partials
//   vit_attention_bwd               dq, dk, dv of softmax(q k^T / sqrt d) v per (pose, head), probabilities recomputed
//   vit_colsum16 / vit_embed_bwd    bias gradients; d linear_mapper
// Every column reduction is two-stage and fixed-order (per-block partial rows, then reduce_rows_f32): deterministic.
#include "vit_ops.cuh"

#include "gemm_sm100.cuh"
#include "train_kernels.cuh"
#include "vit_common.cuh"

namespace lifter {

// split-K factor of a weight-gradient GEMM: enough slabs to give every SM a tile (as lifter_train.cu does)
static int wgrad_splits(int M, int N, int K) {
    const int bn = N <= 64 ? 64 : (N <= 128 ? 128 : 256);
    const int tiles = ceil_div(M, 128) * ceil_div(N, bn);
    int s = num_sms() / tiles;
    if (s < 1) s = 1;
    // K is the token count here (139,264 at 8,192 poses): a 256 x 256 gradient has two tiles, and the cap of 16 slabs the
    // lifter's 8,192-row batches use would leave it on 32 of 148 SMs.  Up to 74 slabs while a slab keeps >= 2,048 rows.
    const int by_rows = K / 2048 > 16 ? K / 2048 : 16;
    const int cap = by_rows < 74 ? by_rows : 74;
    if (s > cap) s = cap;
    return gemm_effective_splits(K, s);
}

namespace {

// Phi(x), phi(x) of the standard normal with the erf polynomial of the forward epilogue (gemm_sm100.cu gelu_erf)
__device__ __forceinline__ void normal_cdf_pdf(float x, float& cdf, float& pdf) {
    const float ax = fabsf(x);
    float t, e;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(t) : "f"(fmaf(ax, 0.3275911f * 0.70710678118654752f, 1.0f)));
    float p = fmaf(0.5f * 1.061405429f, t, 0.5f * -1.453152027f);
    p = fmaf(p, t, 0.5f * 1.421413741f);
    p = fmaf(p, t, 0.5f * -0.284496736f);
    p = fmaf(p, t, 0.5f * 0.254829592f);
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(x * x * (-0.5f * 1.4426950408889634f)));
    const float tail = p * t * e;                                   // Phi(-|x|)
    cdf = x >= 0.0f ? 1.0f - tail : tail;
    pdf = e * 0.3989422804014327f;
}

// mode 0: out = GELU(a)   mode 1: out = b * GELU'(a)   mode 2: out = b * (a > 0).   16-bit, 8 elements per thread.
__global__ void __launch_bounds__(256) vit_pointwise_kernel(const uint4* __restrict__ a, const uint4* b, uint4* out, long n8,
                                                            int mode, int fmt) {       // out may be b (in place)
    griddep_launch();
    griddep_wait();
    const long stride = static_cast<long>(gridDim.x) * blockDim.x;
    for (long i = static_cast<long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n8; i += stride) {
        const uint4 ra = a[i];
        const uint4 rb = mode == 0 ? ra : b[i];
        const uint32_t wa[4] = {ra.x, ra.y, ra.z, ra.w}, wb[4] = {rb.x, rb.y, rb.z, rb.w};
        uint32_t wo[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            const float2 x = unpack16(wa[e], fmt), g = unpack16(wb[e], fmt);
            float2 y;
            if (mode == 2) {
                y.x = x.x > 0.0f ? g.x : 0.0f;
                y.y = x.y > 0.0f ? g.y : 0.0f;
            } else {
                float c0, p0, c1, p1;
                normal_cdf_pdf(x.x, c0, p0);
                normal_cdf_pdf(x.y, c1, p1);
                if (mode == 0) { y.x = x.x * c0; y.y = x.y * c1; }
                else { y.x = g.x * fmaf(x.x, p0, c0); y.y = g.y * fmaf(x.y, p1, c1); }
            }
            wo[e] = pack16(y.x, y.y, fmt);
        }
        out[i] = make_uint4(wo[0], wo[1], wo[2], wo[3]);
    }
}

// partial column sums of a 16-bit matrix [T][W] (W % 8 == 0, W <= 2048): block b sums rows b * rpb .. into part[b][W]
constexpr int CS_THREADS = 256;
__global__ void __launch_bounds__(CS_THREADS) vit_colsum16_kernel(const unsigned short* __restrict__ m, long T, int W, long rpb,
                                                                  float* __restrict__ part, int fmt) {
    griddep_launch();
    griddep_wait();
    __shared__ float sh[CS_THREADS][8];
    const int vpr = W / 8, G = CS_THREADS / vpr > 0 ? CS_THREADS / vpr : 1;      // row groups per block
    const int t = threadIdx.x, vec = t % vpr, g = t / vpr;
    float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    const long r0 = static_cast<long>(blockIdx.x) * rpb, r1 = min(T, r0 + rpb);
    if (g < G && t < G * vpr) {
        for (long r = r0 + g; r < r1; r += G) {
            const uint4 v = *reinterpret_cast<const uint4*>(m + r * W + vec * 8);
            const uint32_t wd[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                const float2 f = unpack16(wd[e], fmt);
                acc[2 * e] += f.x;
                acc[2 * e + 1] += f.y;
            }
        }
    }
#pragma unroll
    for (int e = 0; e < 8; ++e) sh[t][e] = acc[e];
    __syncthreads();
    if (t < vpr) {
        for (int k = 1; k < G; ++k)
#pragma unroll
            for (int e = 0; e < 8; ++e) acc[e] += sh[k * vpr + t][e];
        float* p = part + static_cast<long>(blockIdx.x) * W + t * 8;
#pragma unroll
        for (int e = 0; e < 8; ++e) p[e] = acc[e];
    }
}

}  // namespace
}  // namespace lifter

namespace lifter {
namespace {

// Sum one per-lane column accumulator over the 8 warps of a block (fixed order) and store the block's partial row.
template <int NV>
__device__ __forceinline__ void block_colsum_store(const float (&acc)[NV][8], float* sh /* [8][NV * 256] */, float* dst) {
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    __syncthreads();                                                  // previous round's readers are done with sh
#pragma unroll
    for (int v = 0; v < NV; ++v)
#pragma unroll
        for (int e = 0; e < 8; ++e) sh[wid * (NV * 256) + v * 256 + lane * 8 + e] = acc[v][e];
    __syncthreads();
    for (int c = threadIdx.x; c < NV * 256; c += blockDim.x) {
        float s = 0.0f;
#pragma unroll
        for (int w = 0; w < 8; ++w) s += sh[w * (NV * 256) + c];
        dst[c] = s;
    }
}

template <int NV>
__device__ __forceinline__ void ln_stats(const float (&x)[NV][8], float& mean, float& rstd, float eps) {
    constexpr float inv = 1.0f / (NV * 256);
    float s = 0.0f;
#pragma unroll
    for (int v = 0; v < NV; ++v)
#pragma unroll
        for (int e = 0; e < 8; ++e) s += x[v][e];
    mean = wsum(s) * inv;
    float q = 0.0f;
#pragma unroll
    for (int v = 0; v < NV; ++v)
#pragma unroll
        for (int e = 0; e < 8; ++e) { const float d = x[v][e] - mean; q = fmaf(d, d, q); }
    rstd = rsqrtf(wsum(q) * inv + eps);
}

// dy (in/out: becomes dx), xh = normalised input, g = gain.  dx = rstd (a - mean(a) - xh mean(a xh)), a = dy g;
// dg += dy xh, db += dy.
template <int NV>
__device__ __forceinline__ void ln_bwd_row(float (&dy)[NV][8], const float (&xh)[NV][8], const float (&g)[NV][8], float rstd,
                                           float (&dg)[NV][8], float (&db)[NV][8]) {
    constexpr float inv = 1.0f / (NV * 256);
    float s1 = 0.0f, s2 = 0.0f;
#pragma unroll
    for (int v = 0; v < NV; ++v)
#pragma unroll
        for (int e = 0; e < 8; ++e) {
            dg[v][e] = fmaf(dy[v][e], xh[v][e], dg[v][e]);
            db[v][e] += dy[v][e];
            dy[v][e] *= g[v][e];
            s1 += dy[v][e];
            s2 = fmaf(dy[v][e], xh[v][e], s2);
        }
    const float c1 = wsum(s1) * inv, c2 = wsum(s2) * inv;
#pragma unroll
    for (int v = 0; v < NV; ++v)
#pragma unroll
        for (int e = 0; e < 8; ++e) dy[v][e] = rstd * (dy[v][e] - c1 - xh[v][e] * c2);
}

// dx_out = dres + d/dx [ LN2(LN1(x)) ] . dy   (TWO == false: one LayerNorm; dres nullable).
// part [gridDim.x][4][Hd]: block partials of d gain1, d bias1, d gain2, d bias2.
// Each warp owns `rows_per_warp` consecutive token rows and keeps the raw 16-bit rows of the next LNB_DEPTH rows in flight
// in a private shared-memory ring (cp.async; lane l copies and later reads the same 16 bytes, so the ring needs no
// barrier).  Round 2: the first version prefetched ONE row ahead in registers -- 24 KB in flight per SM, 3.4 TB/s, 128
// registers -- and paid the 4-array block reduction once per 64 rows; the grid is now two blocks per SM.
template <int NV> struct LnbCfg { static constexpr int DEPTH = NV <= 2 ? 4 : 2, SMEM = 8 * DEPTH * 3 * NV * 512; };

template <int DEPTH_M1>
__device__ __forceinline__ void cp_async_wait_group() {
    asm volatile("cp.async.wait_group %0;" :: "n"(DEPTH_M1) : "memory");
}

template <int NV, bool TWO>
__global__ void __launch_bounds__(256, NV == 1 ? 2 : 1) vit_ln_bwd_kernel(const void* __restrict__ x, const void* __restrict__ dy_in,
                                                            const void* __restrict__ dres, long T, long rows_per_warp,
                                                            const float* __restrict__ g1, const float* __restrict__ b1,
                                                            const float* __restrict__ g2, const float* __restrict__ b2, float eps,
                                                            void* __restrict__ dx_out, float* __restrict__ part, int fmt) {
    griddep_launch();
    griddep_wait();
    extern __shared__ __align__(16) unsigned char lnb_smem[];
    constexpr int Hd = NV * 256, DEPTH = LnbCfg<NV>::DEPTH, ROWB = NV * 512;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const long t0 = min(T, (static_cast<long>(blockIdx.x) * 8 + wid) * rows_per_warp);
    const long t1 = min(T, t0 + rows_per_warp);
    unsigned char* ring = lnb_smem + wid * (DEPTH * 3 * ROWB) + lane * 16;
    const uint32_t ring_addr = smem_u32(ring);
    const char* srcs[3] = {static_cast<const char*>(x), static_cast<const char*>(dy_in), static_cast<const char*>(dres)};
    auto issue = [&](long row, int slot) {                            // the three raw rows of `row` -> ring slot
        if (row < t1) {
#pragma unroll
            for (int a = 0; a < 3; ++a) {
                if (a == 2 && dres == nullptr) break;
#pragma unroll
                for (int v = 0; v < NV; ++v)
                    cp_async16(ring_addr + (slot * 3 + a) * ROWB + v * 512, srcs[a] + row * (Hd * 2L) + v * 512 + lane * 16);
            }
        }
        asm volatile("cp.async.commit_group;" ::: "memory");            // one group per row, empty past the end: uniform counting
    };
    auto read_raw = [&](int slot, int a, float (&f)[NV][8]) {
#pragma unroll
        for (int v = 0; v < NV; ++v) {
            const uint4 r = *reinterpret_cast<const uint4*>(ring + (slot * 3 + a) * ROWB + v * 512);
            const uint32_t wd[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                const float2 p = unpack16(wd[e], fmt);
                f[v][2 * e] = p.x;
                f[v][2 * e + 1] = p.y;
            }
        }
    };
#pragma unroll
    for (int k = 0; k < DEPTH; ++k) issue(t0 + k, k);
    RowParams<NV> p1, p2;
    load_row_params<NV>(p1, g1, b1, lane);
    if constexpr (TWO) load_row_params<NV>(p2, g2, b2, lane);
    float dg1[NV][8], db1[NV][8], dg2[TWO ? NV : 1][8], db2[TWO ? NV : 1][8];
#pragma unroll
    for (int v = 0; v < NV; ++v)
#pragma unroll
        for (int e = 0; e < 8; ++e) dg1[v][e] = db1[v][e] = 0.0f;
#pragma unroll
    for (int v = 0; v < (TWO ? NV : 1); ++v)
#pragma unroll
        for (int e = 0; e < 8; ++e) dg2[v][e] = db2[v][e] = 0.0f;
    int slot = 0;
    for (long t = t0; t < t1; ++t) {
        cp_async_wait_group<DEPTH - 1>();                              // the group of row t has landed
        float xh1[NV][8], d[NV][8], r[NV][8];
        read_raw(slot, 0, xh1);
        read_raw(slot, 1, d);
        if (dres != nullptr) read_raw(slot, 2, r);
        issue(t + DEPTH, slot);                                        // the slot's values are in registers: refill it
        slot = slot + 1 == DEPTH ? 0 : slot + 1;
        float mean, rstd1;
        ln_stats<NV>(xh1, mean, rstd1, eps);
#pragma unroll
        for (int v = 0; v < NV; ++v)
#pragma unroll
            for (int e = 0; e < 8; ++e) xh1[v][e] = (xh1[v][e] - mean) * rstd1;
        if constexpr (TWO) {
            float xh2[NV][8];
#pragma unroll
            for (int v = 0; v < NV; ++v)
#pragma unroll
                for (int e = 0; e < 8; ++e) xh2[v][e] = fmaf(xh1[v][e], p1.g[v][e], p1.b[v][e]);
            float mean2, rstd2;
            ln_stats<NV>(xh2, mean2, rstd2, eps);
#pragma unroll
            for (int v = 0; v < NV; ++v)
#pragma unroll
                for (int e = 0; e < 8; ++e) xh2[v][e] = (xh2[v][e] - mean2) * rstd2;
            ln_bwd_row<NV>(d, xh2, p2.g, rstd2, dg2, db2);
        }
        ln_bwd_row<NV>(d, xh1, p1.g, rstd1, dg1, db1);
        if (dres != nullptr) {
#pragma unroll
            for (int v = 0; v < NV; ++v)
#pragma unroll
                for (int e = 0; e < 8; ++e) d[v][e] += r[v][e];
        }
        store_row16<NV>(dx_out, t, lane, d, fmt);
    }
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    float* sh = reinterpret_cast<float*>(lnb_smem);                    // the ring is dead: block reduction scratch [8][Hd]
    float* dst = part + static_cast<long>(blockIdx.x) * 4 * Hd;
    block_colsum_store<NV>(dg1, sh, dst);
    block_colsum_store<NV>(db1, sh, dst + Hd);
    if constexpr (TWO) {
        block_colsum_store<NV>(dg2, sh, dst + 2 * Hd);
        block_colsum_store<NV>(db2, sh, dst + 3 * Hd);
    }
}

// d linear_mapper: part [gridDim.x][1 + in_d][Hd]: block partials of sum_t dh[t][c] (bias) and sum_t dh[t][c] x[t][i]
// Same row walk as vit_ln_bwd_kernel: `rows_per_warp` consecutive rows per warp, the next EMB_DEPTH gradient rows in flight in
// a private cp.async ring, the next row's key point in registers, two blocks per SM (the first version loaded one row,
// used it, loaded the next: 55 us for a 71 MB read).
constexpr int EMB_DEPTH = 4;
template <int NV>
__global__ void __launch_bounds__(256) vit_embed_bwd_kernel(const void* __restrict__ dh, const float* __restrict__ x, long T,
                                                            long rows_per_warp, int in_d, float* __restrict__ part,
                                                            const float* __restrict__ w, float* __restrict__ dx, int fmt) {
    griddep_launch();
    griddep_wait();
    extern __shared__ __align__(16) unsigned char emb_smem[];
    constexpr int Hd = NV * 256, ROWB = NV * 512;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const long t0 = min(T, (static_cast<long>(blockIdx.x) * 8 + wid) * rows_per_warp);
    const long t1 = min(T, t0 + rows_per_warp);
    unsigned char* ring = emb_smem + wid * (EMB_DEPTH * ROWB) + lane * 16;
    const uint32_t ring_addr = smem_u32(ring);
    auto issue = [&](long row, int slot) {
        if (row < t1) {
#pragma unroll
            for (int v = 0; v < NV; ++v)
                cp_async16(ring_addr + slot * ROWB + v * 512, static_cast<const char*>(dh) + row * (Hd * 2L) + v * 512 + lane * 16);
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    };
    auto load_x = [&](long row, float (&xi)[4]) {
#pragma unroll
        for (int i = 0; i < 4; ++i) xi[i] = (i < in_d && row < t1) ? x[row * in_d + i] : 0.0f;
    };
#pragma unroll
    for (int k = 0; k < EMB_DEPTH; ++k) issue(t0 + k, k);
    float acc[5][NV][8];
#pragma unroll
    for (int k = 0; k < 5; ++k)
#pragma unroll
        for (int v = 0; v < NV; ++v)
#pragma unroll
            for (int e = 0; e < 8; ++e) acc[k][v][e] = 0.0f;
    float xn[4];
    load_x(t0, xn);
    int slot = 0;
    for (long t = t0; t < t1; ++t) {
        float xi[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) xi[i] = xn[i];
        load_x(t + 1, xn);
        cp_async_wait_group<EMB_DEPTH - 1>();
        float d[NV][8];
#pragma unroll
        for (int v = 0; v < NV; ++v) {
            const uint4 r = *reinterpret_cast<const uint4*>(ring + slot * ROWB + v * 512);
            const uint32_t wd[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                const float2 f = unpack16(wd[e], fmt);
                d[v][2 * e] = f.x;
                d[v][2 * e + 1] = f.y;
            }
        }
        issue(t + EMB_DEPTH, slot);
        slot = slot + 1 == EMB_DEPTH ? 0 : slot + 1;
#pragma unroll
        for (int v = 0; v < NV; ++v)
#pragma unroll
            for (int e = 0; e < 8; ++e) {
                acc[0][v][e] += d[v][e];
#pragma unroll
                for (int i = 0; i < 4; ++i) acc[1 + i][v][e] = fmaf(d[v][e], xi[i], acc[1 + i][v][e]);
            }
        if (dx != nullptr) {                                          // d loss / d key points: dx[t][i] = sum_c dh[t][c] W[c][i]
            for (int i = 0; i < in_d; ++i) {
                float sx = 0.0f;
#pragma unroll
                for (int v = 0; v < NV; ++v)
#pragma unroll
                    for (int e = 0; e < 8; ++e) sx = fmaf(d[v][e], w[(v * 256 + lane * 8 + e) * in_d + i], sx);
                sx = wsum(sx);
                if (lane == 0) dx[t * in_d + i] = sx;
            }
        }
    }
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    float* sh = reinterpret_cast<float*>(emb_smem);                    // the ring is dead: block reduction scratch [8][Hd]
    float* dst = part + static_cast<long>(blockIdx.x) * (1 + in_d) * Hd;
#pragma unroll
    for (int k = 0; k < 5; ++k)
        if (k <= in_d) block_colsum_store<NV>(acc[k], sh, dst + k * Hd);
}

}  // namespace
}  // namespace lifter

namespace lifter {
namespace {

// Backward of softmax(q k^T / sqrt d) v for one (pose, head) per warp on mma.sync m16n8k16 (fp32 accumulate), the layout
// of the forward kernel: q, k, v, dO slices staged with `tokens + 1` rows (every MMA-tile row past the token count aliases
// the zero row), all copies of a (pose, head) in flight at once (cp.async).  Four phases, each reading shared memory only:
//   1. per 16-row query tile: S = q k^T and dP = dO v^T as accumulator fragments, softmax and
//      dS = P (dP - rowsum(dP P)) / sqrt d in fp32 on the fragments; P and dS go to shared memory as 16-bit [query][key]
//      tiles (again tokens + 1 rows: the tile rows past the token count are the zero row);
//   2. dv = P^T dO   (A operand = the P tile TRANSPOSED by ldmatrix.trans)  -> staged over the v rows (dead after phase 1);
//   3. dk = dS^T q   (same, from the dS tile)                               -> staged over the dO rows (dead after phase 2);
//   4. dq = dS k     (A operand = the dS tile as stored)                    -> staged over the q rows (dead after phase 3);
// then dq | dk | dv leave with coalesced 16-byte stores.  Round-2 rewrite: the first version issued one global load per
// loop trip (1.2 TB/s, latency-bound), 3,260 instructions per (pose, head) -- runtime operand-format branches around every
// MMA / pack, 96 predicated 4-byte global stores -- and held 15.5 KB of shared memory per warp (12 warps per SM).
// qkv, dqkv 16-bit [T][3 Hd]; dO 16-bit [T][Hd].
constexpr int ATTB_WARPS = 4;
constexpr int ATTB_LDP = 40;                           // halfs per row of the P / dS tiles (32 keys + 8: conflict-free ldmatrix)

__device__ __forceinline__ void attb_ldsm_x4(uint32_t (&r)[4], uint32_t addr) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0, %1, %2, %3}, [%4];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}
__device__ __forceinline__ void attb_ldsm_x4_trans(uint32_t (&r)[4], uint32_t addr) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0, %1, %2, %3}, [%4];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}
template <int FMT>
__device__ __forceinline__ void attb_mma(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    if constexpr (FMT == FMT_BF16)
        asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, {%0, %1, %2, %3};"
                     : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3]) : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
    else
        asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, {%0, %1, %2, %3};"
                     : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3]) : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

template <int DH, int FMT>
__global__ void __launch_bounds__(ATTB_WARPS * 32) vit_attention_bwd_kernel(const void* __restrict__ qkv, const void* __restrict__ dO,
                                                                            long pose_heads, int tokens, int heads, float scale,
                                                                            void* __restrict__ dqkv) {
    griddep_launch();
    griddep_wait();
    constexpr int LDS = DH + 8, VPR = DH / 8, RPP = 32 / VPR;         // RPP: staged rows per warp-wide pass of 16-byte vectors
    extern __shared__ __align__(16) unsigned char attb_smem[];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const long ph = static_cast<long>(blockIdx.x) * ATTB_WARPS + wid;
    if (ph >= pose_heads) return;
    const long pose = ph / heads;
    const int head = static_cast<int>(ph - pose * heads);
    const int Hd = heads * DH, N = tokens, R = tokens + 1;
    const int MT = (N + 15) / 16, NT = (N + 7) / 8;
    const size_t per_warp = (static_cast<size_t>(4) * R * LDS + static_cast<size_t>(2) * R * ATTB_LDP) * 2;
    unsigned short* sq = reinterpret_cast<unsigned short*>(attb_smem + wid * per_warp);
    unsigned short* sP = sq + 4 * R * LDS;
    unsigned short* sS = sP + R * ATTB_LDP;
    const unsigned short* src = static_cast<const unsigned short*>(qkv) + pose * N * (3L * Hd) + head * DH;
    const unsigned short* srcdo = static_cast<const unsigned short*>(dO) + pose * N * static_cast<long>(Hd) + head * DH;
    const int vec = lane % VPR, rsub = lane / VPR;
#pragma unroll
    for (int slice = 0; slice < 4; ++slice) {
        const unsigned short* g0 = slice < 3 ? src + slice * Hd + vec * 8 : srcdo + vec * 8;
        const long gstride = slice < 3 ? 3L * Hd : static_cast<long>(Hd);
        for (int n = rsub; n < R; n += RPP) {
            unsigned short* dst = sq + (slice * R + n) * LDS + vec * 8;
            if (n < N) cp_async16(smem_u32(dst), g0 + n * gstride);
            else *reinterpret_cast<uint4*>(dst) = make_uint4(0u, 0u, 0u, 0u);
        }
    }
    if (lane < 2 * (ATTB_LDP / 8)) {                                  // zero rows of the P and dS tiles
        unsigned short* z = (lane < ATTB_LDP / 8 ? sP : sS) + N * ATTB_LDP + (lane % (ATTB_LDP / 8)) * 8;
        *reinterpret_cast<uint4*>(z) = make_uint4(0u, 0u, 0u, 0u);
    }
    cp_async_wait_all();
    __syncwarp();
    const uint32_t q_addr = smem_u32(sq), k_addr = q_addr + R * LDS * 2, v_addr = k_addr + R * LDS * 2, do_addr = v_addr + R * LDS * 2;
    const uint32_t p_addr = smem_u32(sP), s_addr = smem_u32(sS);
    const int g = lane >> 2, t4 = lane & 3;
    const float scale_log2e = scale * 1.4426950408889634f;
    // ldmatrix row selectors of this lane: "a" = A-operand / transposed-B pattern, "b" = B-operand pattern
    const int row_a = (lane & 7) + ((lane >> 3) & 1) * 8, col_a = (lane >> 4) * 8;
    const int row_b = (lane & 7) + (lane >> 4) * 8, col_b = ((lane >> 3) & 1) * 8;
    // ---- phase 1 ------------------------------------------------------------------------------------------------------
    for (int mt = 0; mt < MT; ++mt) {
        float sc[4][4], dp[4][4];
#pragma unroll
        for (int nt = 0; nt < 4; ++nt)
#pragma unroll
            for (int e = 0; e < 4; ++e) sc[nt][e] = dp[nt][e] = 0.0f;
        const int arow = min(mt * 16 + row_a, N);
#pragma unroll
        for (int ks = 0; ks < DH / 16; ++ks) {
            uint32_t aq[4], ad[4];
            attb_ldsm_x4(aq, q_addr + (arow * LDS + ks * 16 + col_a) * 2);
            attb_ldsm_x4(ad, do_addr + (arow * LDS + ks * 16 + col_a) * 2);
#pragma unroll
            for (int np = 0; np < 2; ++np) {
                if (np * 2 < NT) {
                    uint32_t bk[4], bv[4];
                    const int brow = min(np * 16 + row_b, N);
                    attb_ldsm_x4(bk, k_addr + (brow * LDS + ks * 16 + col_b) * 2);
                    attb_ldsm_x4(bv, v_addr + (brow * LDS + ks * 16 + col_b) * 2);
                    attb_mma<FMT>(sc[np * 2], aq, bk[0], bk[1]);
                    attb_mma<FMT>(dp[np * 2], ad, bv[0], bv[1]);
                    if (np * 2 + 1 < NT) {
                        attb_mma<FMT>(sc[np * 2 + 1], aq, bk[2], bk[3]);
                        attb_mma<FMT>(dp[np * 2 + 1], ad, bv[2], bv[3]);
                    }
                }
            }
        }
        // softmax rows g, g + 8
        float mx[2] = {-CUDART_INF_F, -CUDART_INF_F}, den[2] = {0.0f, 0.0f}, D[2] = {0.0f, 0.0f};
#pragma unroll
        for (int nt = 0; nt < 4; ++nt)
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                const int col = nt * 8 + t4 * 2 + (e & 1);
                sc[nt][e] = col < N ? sc[nt][e] * scale_log2e : -CUDART_INF_F;
                mx[e >> 1] = fmaxf(mx[e >> 1], sc[nt][e]);
            }
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            mx[h] = fmaxf(mx[h], __shfl_xor_sync(0xffffffffu, mx[h], 1));
            mx[h] = fmaxf(mx[h], __shfl_xor_sync(0xffffffffu, mx[h], 2));
        }
#pragma unroll
        for (int nt = 0; nt < 4; ++nt)
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                sc[nt][e] = exp2f(sc[nt][e] - mx[e >> 1]);
                den[e >> 1] += sc[nt][e];
            }
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            den[h] += __shfl_xor_sync(0xffffffffu, den[h], 1);
            den[h] += __shfl_xor_sync(0xffffffffu, den[h], 2);
            den[h] = 1.0f / den[h];
        }
#pragma unroll
        for (int nt = 0; nt < 4; ++nt)
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                sc[nt][e] *= den[e >> 1];                                         // P
                D[e >> 1] = fmaf(sc[nt][e], dp[nt][e], D[e >> 1]);
            }
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            D[h] += __shfl_xor_sync(0xffffffffu, D[h], 1);
            D[h] += __shfl_xor_sync(0xffffffffu, D[h], 2);
        }
        const int r0 = mt * 16 + g, r1 = r0 + 8;                                  // query rows past the token count stay the zero row
#pragma unroll
        for (int nt = 0; nt < 4; ++nt) {
#pragma unroll
            for (int e = 0; e < 4; ++e) dp[nt][e] = sc[nt][e] * (dp[nt][e] - D[e >> 1]) * scale;      // dS
            const int col = nt * 8 + t4 * 2;
            if (r0 < N) {
                *reinterpret_cast<uint32_t*>(sP + r0 * ATTB_LDP + col) = pack16(sc[nt][0], sc[nt][1], FMT);
                *reinterpret_cast<uint32_t*>(sS + r0 * ATTB_LDP + col) = pack16(dp[nt][0], dp[nt][1], FMT);
            }
            if (r1 < N) {
                *reinterpret_cast<uint32_t*>(sP + r1 * ATTB_LDP + col) = pack16(sc[nt][2], sc[nt][3], FMT);
                *reinterpret_cast<uint32_t*>(sS + r1 * ATTB_LDP + col) = pack16(dp[nt][2], dp[nt][3], FMT);
            }
        }
    }
    __syncwarp();
    auto stage = [&](unsigned short* region, int row0, const float (&oc)[DH / 8][4]) {     // accumulator tile -> staged rows
        const int r0 = row0 + g, r1 = r0 + 8;
#pragma unroll
        for (int dn = 0; dn < DH / 8; ++dn) {
            if (r0 < N) *reinterpret_cast<uint32_t*>(region + r0 * LDS + dn * 8 + t4 * 2) = pack16(oc[dn][0], oc[dn][1], FMT);
            if (r1 < N) *reinterpret_cast<uint32_t*>(region + r1 * LDS + dn * 8 + t4 * 2) = pack16(oc[dn][2], oc[dn][3], FMT);
        }
    };
    // ---- phases 2 and 3: dv = P^T dO over the v rows, then dk = dS^T q over the dO rows ---------------------------------------
#pragma unroll
    for (int which = 0; which < 2; ++which) {
        const uint32_t a_base = which == 0 ? p_addr : s_addr;
        const uint32_t b_base = which == 0 ? do_addr : q_addr;
        unsigned short* region = which == 0 ? sq + 2 * R * LDS : sq + 3 * R * LDS;
        float oc[2][DH / 8][4];                                                  // both key tiles at once: one load of the B
#pragma unroll                                                                   // fragments (q / dO columns) serves both
        for (int jt = 0; jt < 2; ++jt)
#pragma unroll
            for (int dn = 0; dn < DH / 8; ++dn)
#pragma unroll
                for (int e = 0; e < 4; ++e) oc[jt][dn][e] = 0.0f;
        for (int it = 0; it < MT; ++it) {                                        // queries it * 16 .. + 15
            const int qrow_a = min(it * 16 + (lane & 7) + ((lane >> 4) & 1) * 8, N);
            const int qrow_b = min(it * 16 + row_a, N);
            uint32_t b[DH / 16][4];
#pragma unroll
            for (int dq = 0; dq < DH / 16; ++dq) attb_ldsm_x4_trans(b[dq], b_base + (qrow_b * LDS + dq * 16 + col_a) * 2);
#pragma unroll
            for (int jt = 0; jt < 2; ++jt) {
                if (jt < MT) {
                    uint32_t a[4];                                               // A = (tile stored [query][key])^T
                    attb_ldsm_x4_trans(a, a_base + (qrow_a * ATTB_LDP + jt * 16 + ((lane >> 3) & 1) * 8) * 2);
#pragma unroll
                    for (int dq = 0; dq < DH / 16; ++dq) {
                        attb_mma<FMT>(oc[jt][dq * 2], a, b[dq][0], b[dq][1]);
                        attb_mma<FMT>(oc[jt][dq * 2 + 1], a, b[dq][2], b[dq][3]);
                    }
                }
            }
        }
        __syncwarp();                                                            // every lane is done reading the region's old rows
#pragma unroll
        for (int jt = 0; jt < 2; ++jt)
            if (jt < MT) stage(region, jt * 16, oc[jt]);
    }
    // ---- phase 4: dq = dS k over the q rows ----------------------------------------------------------------------------------
    for (int mt = 0; mt < MT; ++mt) {
        float oc[DH / 8][4];
#pragma unroll
        for (int dn = 0; dn < DH / 8; ++dn)
#pragma unroll
            for (int e = 0; e < 4; ++e) oc[dn][e] = 0.0f;
        const int arow = min(mt * 16 + row_a, N);
#pragma unroll
        for (int kt = 0; kt < 2; ++kt) {
            if (kt * 2 < NT) {
                uint32_t pa[4];
                attb_ldsm_x4(pa, s_addr + (arow * ATTB_LDP + kt * 16 + col_a) * 2);
                const int krow = min(kt * 16 + row_a, N);
#pragma unroll
                for (int dq = 0; dq < DH / 16; ++dq) {
                    uint32_t bk[4];
                    attb_ldsm_x4_trans(bk, k_addr + (krow * LDS + dq * 16 + col_a) * 2);
                    attb_mma<FMT>(oc[dq * 2], pa, bk[0], bk[1]);
                    attb_mma<FMT>(oc[dq * 2 + 1], pa, bk[2], bk[3]);
                }
            }
        }
        stage(sq, mt * 16, oc);                                                  // q: last read in phase 3
    }
    __syncwarp();
    unsigned short* out_base = static_cast<unsigned short*>(dqkv) + pose * N * (3L * Hd) + head * DH + vec * 8;
#pragma unroll
    for (int slice = 0; slice < 3; ++slice) {                                    // dq <- q rows, dk <- dO rows, dv <- v rows
        const unsigned short* region = sq + (slice == 0 ? 0 : slice == 1 ? 3 : 2) * R * LDS + vec * 8;
        for (int n = rsub; n < N; n += RPP)
            *reinterpret_cast<uint4*>(out_base + n * (3L * Hd) + slice * Hd) = *reinterpret_cast<const uint4*>(region + n * LDS);
    }
}

size_t attb_smem_bytes(int tokens, int dh) {
    return static_cast<size_t>(ATTB_WARPS) * (static_cast<size_t>(4) * (tokens + 1) * (dh + 8) + static_cast<size_t>(2) * (tokens + 1) * ATTB_LDP) * 2;
}

template <int DH, int FMT>
int attention_bwd_launch(const void* qkv, const void* dO, long pose_heads, int tokens, int heads, float scale, void* dqkv, size_t smem,
                         cudaStream_t st) {
    static PerDeviceOnce once;
    bool* configured = once.slot();
    if (!*configured) {
        LIFTER_CHECK_CUDA(cudaFuncSetAttribute(vit_attention_bwd_kernel<DH, FMT>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024));
        *configured = true;
    }
    LIFTER_CHECK_CUDA(launch_pdl("vit_attention_bwd_kernel", vit_attention_bwd_kernel<DH, FMT>, dim3(static_cast<unsigned>((pose_heads + ATTB_WARPS - 1) / ATTB_WARPS)),
                                 dim3(ATTB_WARPS * 32), smem, st, qkv, dO, pose_heads, tokens, heads, scale, dqkv));
    return LIFTER_OK;
}
int attention_bwd(const void* qkv, const void* dO, long pose_heads, int tokens, int heads, int dh, float scale, void* dqkv, int fmt,
                  size_t smem, cudaStream_t st) {
    if (dh == 64) return fmt == FMT_BF16 ? attention_bwd_launch<64, FMT_BF16>(qkv, dO, pose_heads, tokens, heads, scale, dqkv, smem, st)
                                         : attention_bwd_launch<64, FMT_F16>(qkv, dO, pose_heads, tokens, heads, scale, dqkv, smem, st);
    return fmt == FMT_BF16 ? attention_bwd_launch<32, FMT_BF16>(qkv, dO, pose_heads, tokens, heads, scale, dqkv, smem, st)
                           : attention_bwd_launch<32, FMT_F16>(qkv, dO, pose_heads, tokens, heads, scale, dqkv, smem, st);
}

}  // namespace
}  // namespace lifter

namespace lifter {

namespace {

struct SavedLayout {                       // 16-bit matrices kept by the train-mode forward, element offsets
    long T; int H, M, Hh, nb;
    long h0, blocks, block_stride, r, total;
    long u1(int b) const { return blocks + b * block_stride; }
    long qkv(int b) const { return u1(b) + T * H; }
    long o(int b) const { return qkv(b) + T * 3L * H; }
    long h1(int b) const { return o(b) + T * H; }
    long u2(int b) const { return h1(b) + T * H; }
    long z1(int b) const { return u2(b) + T * H; }
    long m(int b) const { return z1(b) + T * static_cast<long>(M); }
    long hout(int b) const { return m(b) + T * static_cast<long>(M); }
    long hin(int b) const { return b == 0 ? h0 : hout(b - 1); }
    long hlast() const { return nb == 0 ? h0 : hout(nb - 1); }
};

SavedLayout saved_layout(const VitDesc& d, int64_t poses) {
    SavedLayout s;
    s.T = poses * d.tokens; s.H = d.hidden; s.M = d.mlp_dim; s.Hh = d.head_dim; s.nb = d.blocks;
    s.h0 = 0;
    s.blocks = s.T * s.H;
    s.block_stride = s.T * (8L * s.H + 2L * s.M);
    s.r = s.blocks + s.nb * s.block_stride;
    s.total = s.r + s.T * s.Hh;
    return s;
}

long slab_floats(int M_out, int N_in, int K) {
    return static_cast<long>(gemm_effective_splits(K, wgrad_splits(M_out, N_in, K))) * gemm_slab_rows(M_out) * N_in;
}

struct BwdScratch {
    unsigned short *dout16, *dr, *dhA, *dhB, *du, *dO, *dm, *dqkv;
    float *slabs, *part, *part_side;      // part_side: column-sum partials of the weight-gradient stream
    size_t bytes;
};

int part_rows(long T) { return static_cast<int>((T + 8 * VIT_RPW - 1) / (8 * VIT_RPW)); }
int colsum_blocks(long T) { const long b = (T + 63) / 64; return static_cast<int>(b < 592 ? b : 592); }

BwdScratch bwd_scratch(const VitDesc& d, int64_t poses, void* base) {
    const long T = poses * d.tokens;
    const int H = d.hidden, M = d.mlp_dim, Hh = d.head_dim, K = static_cast<int>(T);
    BwdScratch s;
    char* p = static_cast<char*>(base);
    auto take = [&](size_t bytes) { char* q = p; p += (bytes + 255) / 256 * 256; return q; };
    s.dout16 = reinterpret_cast<unsigned short*>(take(T * 64 * 2));
    s.dr = reinterpret_cast<unsigned short*>(take(T * Hh * 2L));
    s.dhA = reinterpret_cast<unsigned short*>(take(T * H * 2L));
    s.dhB = reinterpret_cast<unsigned short*>(take(T * H * 2L));
    s.du = reinterpret_cast<unsigned short*>(take(T * H * 2L));
    s.dO = reinterpret_cast<unsigned short*>(take(T * H * 2L));
    s.dm = reinterpret_cast<unsigned short*>(take(T * M * 2L));
    s.dqkv = reinterpret_cast<unsigned short*>(take(T * 3L * H * 2));
    long slab = slab_floats(64, Hh, K);
    const long cands[5] = {slab_floats(Hh, H, K), slab_floats(3 * H, H, K), slab_floats(H, H, K), slab_floats(M, H, K), slab_floats(H, M, K)};
    for (long c : cands) slab = c > slab ? c : slab;
    s.slabs = reinterpret_cast<float*>(take(slab * 4));
    long part = static_cast<long>(part_rows(T)) * 5 * H;                          // LayerNorm (4 H) / embedding (<= 5 H) partial rows
    const long cs = static_cast<long>(colsum_blocks(T)) * (M > 3 * H ? M : 3 * H);
    part = cs > part ? cs : part;
    const long po = static_cast<long>(ew_partial_rows(T, 64)) * 64;
    part = po > part ? po : part;
    s.part = reinterpret_cast<float*>(take(part * 4));
    s.part_side = reinterpret_cast<float*>(take(cs * 4));
    s.bytes = static_cast<size_t>(p - static_cast<char*>(base));
    return s;
}

int pointwise(const void* a, const void* b, void* out, long n, int mode, int fmt, cudaStream_t st) {
    const long n8 = n / 8;
    long blocks = (n8 + 255) / 256;
    const long cap = static_cast<long>(num_sms()) * 8;
    if (blocks > cap) blocks = cap;
    LIFTER_CHECK_CUDA(launch_pdl("vit_pointwise_kernel", vit_pointwise_kernel, dim3(static_cast<unsigned>(blocks)), dim3(256), 0, st,
                                 static_cast<const uint4*>(a), static_cast<const uint4*>(b), static_cast<uint4*>(out), n8, mode, fmt));
    return LIFTER_OK;
}

// out[W] = column sums of the 16-bit matrix m [T][W]
int colsum16(const unsigned short* m, long T, int W, float* part, float* out, int fmt, cudaStream_t st) {
    const int blocks = colsum_blocks(T);
    const long rpb = (T + blocks - 1) / blocks;
    LIFTER_CHECK_CUDA(launch_pdl("vit_colsum16_kernel", vit_colsum16_kernel, dim3(blocks), dim3(CS_THREADS), 0, st, m, T, W, rpb, part, fmt));
    return reduce_rows_f32(part, blocks, W, W, out, st);
}

// dW [M_out][N_in] (dense fp32, rows_valid x N_in) = dZ^T X over the T rows
int wgrad16(const unsigned short* dZ, int M_out, const unsigned short* X, int N_in, int T, int rows_valid, float* slabs, float* grad,
            int fmt, cudaStream_t st) {
    GemmArgs g;
    g.A = dZ; g.lda = M_out; g.a_mn_major = 1;
    g.B = X; g.ldb = N_in; g.b_mn_major = 1;
    g.M = M_out; g.N = N_in; g.K = T; g.fmt = fmt; g.out_mode = OUT_32_TMA; g.D = slabs; g.ldd = N_in;
    g.splits = wgrad_splits(M_out, N_in, T);
    g.tag = "vit_wgrad";
    LIFTER_PROPAGATE(gemm_launch(g, st));
    return wgrad_reduce(slabs, gemm_effective_splits(T, g.splits), gemm_slab_rows(M_out), N_in, rows_valid, N_in, grad, st);
}

// dX [T][N_in] (16-bit) = dZ [T][M_out] W [M_out][N_in] (+ residual)
int dgrad16(const unsigned short* dZ, int M_out, const void* W, int N_in, int T, unsigned short* dX, const unsigned short* residual,
            int fmt, cudaStream_t st, int residual_op = 0) {
    GemmArgs g;
    g.A = dZ; g.lda = M_out;
    g.B = W; g.ldb = N_in; g.b_mn_major = 1;
    g.M = T; g.N = N_in; g.K = M_out; g.fmt = fmt; g.out_mode = OUT_16_TMA; g.D = dX; g.ldd = N_in;
    if (residual != nullptr) { g.residual = residual; g.ldr = N_in; g.residual_op = residual_op; }
    g.tag = "vit_dgrad";
    return gemm_launch(g, st);
}

// Nothing on the data-gradient chain reads a parameter gradient, so the weight-gradient GEMMs (tensor pipe) and the bias
// column sums run on a second stream beside the attention / LayerNorm / GELU backward kernels (FP32 pipe, HBM) of the chain.
// One library-owned stream and a few timing-disabled events per device, created on first use and kept until the process
// exits (the ownership exception include/lifter_b200.h documents); forks from and joins back into the caller's stream with
// events, so the call stays ordered on `st` and is legal under stream capture.  LIFTER_VIT_WGRAD_STREAM=0: one stream.
struct VitSide {
    cudaStream_t st = nullptr;
    cudaEvent_t fork = nullptr, dhA = nullptr, dm = nullptr, dhB = nullptr, dqkv = nullptr, join = nullptr;
};
VitSide* vit_side() {
    static VitSide per_device[64];
    const char* e = getenv("LIFTER_VIT_WGRAD_STREAM");
    if (e != nullptr && e[0] == '0') return nullptr;
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return nullptr;
    VitSide* s = &per_device[dev];
    if (s->st == nullptr) {
        if (cudaStreamCreateWithFlags(&s->st, cudaStreamNonBlocking) != cudaSuccess) { s->st = nullptr; return nullptr; }
        cudaEvent_t* ev[6] = {&s->fork, &s->dhA, &s->dm, &s->dhB, &s->dqkv, &s->join};
        for (cudaEvent_t* x : ev)
            if (cudaEventCreateWithFlags(x, cudaEventDisableTiming) != cudaSuccess) return nullptr;
    }
    return s;
}
// everything enqueued on `from` so far happens before whatever is enqueued on `to` next
int order_after(cudaEvent_t ev, cudaStream_t from, cudaStream_t to) {
    if (from == to) return LIFTER_OK;
    LIFTER_CHECK_CUDA(cudaEventRecord(ev, from));
    LIFTER_CHECK_CUDA(cudaStreamWaitEvent(to, ev, 0));
    return LIFTER_OK;
}

// LayerNorm backward over T rows; returns the number of partial rows it wrote to `part` (each [4][H]) in *blocks
template <int NV, bool TWO>
int ln_bwd_launch(const void* x, const void* dy, const void* dres, long T, const float* g1, const float* b1, const float* g2, const float* b2,
                  float eps, void* dx_out, float* part, int max_blocks, int fmt, int* blocks, cudaStream_t st) {
    static PerDeviceOnce once;
    bool* configured = once.slot();
    if (!*configured) {
        LIFTER_CHECK_CUDA(cudaFuncSetAttribute(vit_ln_bwd_kernel<NV, TWO>, cudaFuncAttributeMaxDynamicSharedMemorySize, LnbCfg<NV>::SMEM));
        *configured = true;
    }
    int grid = 2 * num_sms();
    if (grid > max_blocks) grid = max_blocks;
    const long rows_per_warp = (T + grid * 8L - 1) / (grid * 8L);
    grid = static_cast<int>((T + rows_per_warp * 8 - 1) / (rows_per_warp * 8));
    *blocks = grid;
    LIFTER_CHECK_CUDA(launch_pdl("vit_ln_bwd_kernel", vit_ln_bwd_kernel<NV, TWO>, dim3(grid), dim3(256), static_cast<size_t>(LnbCfg<NV>::SMEM), st,
                                 x, dy, dres, T, rows_per_warp, g1, b1, g2, b2, eps, dx_out, part, fmt));
    return LIFTER_OK;
}
int ln_bwd(int nv, const void* x, const void* dy, const void* dres, long T, const float* g1, const float* b1, const float* g2, const float* b2,
           float eps, void* dx_out, float* part, int max_blocks, int fmt, int* blocks, cudaStream_t st) {
    if (g2 != nullptr) { VIT_DISPATCH_NV(nv, return (ln_bwd_launch<NV, true>(x, dy, dres, T, g1, b1, g2, b2, eps, dx_out, part, max_blocks, fmt, blocks, st))); }
    else               { VIT_DISPATCH_NV(nv, return (ln_bwd_launch<NV, false>(x, dy, dres, T, g1, b1, g2, b2, eps, dx_out, part, max_blocks, fmt, blocks, st))); }
    return LIFTER_OK;
}

template <int NV>
int embed_bwd_launch(const void* dh, const float* x, long T, int in_d, float* part, int max_blocks, const float* w, float* dx, int fmt,
                     int* blocks, cudaStream_t st) {
    constexpr int SMEM = 8 * EMB_DEPTH * NV * 512;
    static PerDeviceOnce once;
    bool* configured = once.slot();
    if (!*configured) {
        LIFTER_CHECK_CUDA(cudaFuncSetAttribute(vit_embed_bwd_kernel<NV>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM));
        *configured = true;
    }
    int grid = 2 * num_sms();
    if (grid > max_blocks) grid = max_blocks;
    const long rows_per_warp = (T + grid * 8L - 1) / (grid * 8L);
    grid = static_cast<int>((T + rows_per_warp * 8 - 1) / (rows_per_warp * 8));
    *blocks = grid;
    LIFTER_CHECK_CUDA(launch_pdl("vit_embed_bwd_kernel", vit_embed_bwd_kernel<NV>, dim3(grid), dim3(256), static_cast<size_t>(SMEM), st,
                                 dh, x, T, rows_per_warp, in_d, part, w, dx, fmt));
    return LIFTER_OK;
}

int validate_train(const VitDesc& d) {
    LIFTER_PROPAGATE(vit_validate(d));
    LIFTER_REQUIRE(d.mlp_dim <= 2048 && 3 * d.hidden <= 2048, "vit training: hidden width %d unsupported (256 or 512)", d.hidden);
    return LIFTER_OK;
}

}  // namespace

int64_t vit_saved_bytes(const VitDesc& d, int64_t poses) {
    if (validate_train(d) != LIFTER_OK || poses < 0) return LIFTER_ERR_INVALID;
    return saved_layout(d, poses).total * 2 + 256;
}

int64_t vit_bwd_scratch_bytes(const VitDesc& d, int64_t poses) {
    if (validate_train(d) != LIFTER_OK || poses < 1) return LIFTER_ERR_INVALID;
    return static_cast<int64_t>(bwd_scratch(d, poses, nullptr).bytes) + 256;
}

}  // namespace lifter

namespace lifter {

int vit_train_forward(const VitDesc& d, const float* x, int64_t poses, void* saved, float* out, cudaStream_t st) {
    LIFTER_PROPAGATE(validate_train(d));
    if (poses <= 0) return LIFTER_OK;
    LIFTER_REQUIRE(x && saved && out, "vit_train_forward: null pointer");
    LIFTER_REQUIRE(reinterpret_cast<uintptr_t>(saved) % 16 == 0, "vit_train_forward: saved-activation buffer must be 16-byte aligned");
    const SavedLayout L = saved_layout(d, poses);
    LIFTER_REQUIRE(L.T <= 0x7fffffffL, "vit_train_forward: %ld tokens in one call", L.T);
    const int H = d.hidden, dh = H / d.heads, M = static_cast<int>(L.T);
    unsigned short* S = static_cast<unsigned short*>(saved);
    const long T = L.T;
    LIFTER_PROPAGATE(vit_embed(d, x, T, S + L.h0, d.blocks > 0 ? d.ln : nullptr, d.blocks > 0 ? static_cast<void*>(S + L.u1(0)) : nullptr, st));
    const long pose_heads = poses * d.heads;
    for (int b = 0; b < d.blocks; ++b) {
        const float* ln = d.ln + static_cast<long>(b) * 6 * H;
        const char* wqkv = static_cast<const char*>(d.w_qkv) + static_cast<long>(b) * 3 * H * H * 2;
        const char* wout = static_cast<const char*>(d.w_out) + static_cast<long>(b) * H * H * 2;
        const char* w1 = static_cast<const char*>(d.w_mlp1) + static_cast<long>(b) * d.mlp_dim * H * 2;
        const char* w2 = static_cast<const char*>(d.w_mlp2) + static_cast<long>(b) * H * d.mlp_dim * 2;
        unsigned short *hin = S + L.hin(b), *u1 = S + L.u1(b), *qkv = S + L.qkv(b), *o = S + L.o(b), *h1 = S + L.h1(b),
                       *u2 = S + L.u2(b), *z1 = S + L.z1(b), *m = S + L.m(b), *hout = S + L.hout(b);
        if (b > 0) LIFTER_PROPAGATE(vit_layernorm(hin, T, ln, ln + H, ln + 2 * H, ln + 3 * H, d.ln_eps, u1, H, d.fmt, st));
        GemmArgs g;
        g.A = u1; g.lda = H; g.B = wqkv; g.ldb = H; g.M = M; g.N = 3 * H; g.K = H; g.fmt = d.fmt; g.D = qkv; g.ldd = 3 * H; g.tag = "vit_qkv";
        LIFTER_PROPAGATE(gemm_launch(g, st));
        LIFTER_PROPAGATE(vit_attention(qkv, pose_heads, d.tokens, d.heads, dh, o, d.fmt, st));
        GemmArgs go;
        go.A = o; go.lda = H; go.B = wout; go.ldb = H; go.M = M; go.N = H; go.K = H; go.fmt = d.fmt; go.D = h1; go.ldd = H;
        go.residual = hin; go.ldr = H; go.tag = "vit_attn_out";
        LIFTER_PROPAGATE(gemm_launch(go, st));
        LIFTER_PROPAGATE(vit_layernorm(h1, T, ln + 4 * H, ln + 5 * H, nullptr, nullptr, d.ln_eps, u2, H, d.fmt, st));
        GemmArgs g1;                                              // z1 = u2 W1^T + b1 kept (GELU is not invertible), then m = GELU(z1)
        g1.A = u2; g1.lda = H; g1.B = w1; g1.ldb = H; g1.M = M; g1.N = d.mlp_dim; g1.K = H; g1.fmt = d.fmt; g1.D = z1; g1.ldd = d.mlp_dim;
        g1.shift = d.mlp_b1 + static_cast<long>(b) * d.mlp_dim; g1.tag = "vit_mlp1";
        LIFTER_PROPAGATE(gemm_launch(g1, st));
        LIFTER_PROPAGATE(pointwise(z1, nullptr, m, T * d.mlp_dim, 0, d.fmt, st));
        GemmArgs g2;
        g2.A = m; g2.lda = d.mlp_dim; g2.B = w2; g2.ldb = d.mlp_dim; g2.M = M; g2.N = H; g2.K = d.mlp_dim; g2.fmt = d.fmt; g2.D = hout; g2.ldd = H;
        g2.shift = d.mlp_b2 + static_cast<long>(b) * H; g2.residual = h1; g2.ldr = H; g2.tag = "vit_mlp2";
        LIFTER_PROPAGATE(gemm_launch(g2, st));
    }
    GemmArgs gh;
    gh.A = S + L.hlast(); gh.lda = H; gh.B = d.w_head1; gh.ldb = H; gh.M = M; gh.N = d.head_dim; gh.K = H; gh.fmt = d.fmt;
    gh.D = S + L.r; gh.ldd = d.head_dim; gh.shift = d.head_b1; gh.relu = 1; gh.tag = "vit_head1";
    LIFTER_PROPAGATE(gemm_launch(gh, st));
    GemmArgs gf;
    gf.A = S + L.r; gf.lda = d.head_dim; gf.B = d.w_head2; gf.ldb = d.head_dim; gf.M = M; gf.N = 64; gf.K = d.head_dim; gf.fmt = d.fmt;
    gf.out_mode = OUT_32_DIRECT; gf.D = out; gf.n_valid = d.out_dim; gf.shift = d.head_b2; gf.tag = "vit_head2";
    return gemm_launch(gf, st);
}

int vit_backward(const VitDesc& d, const float* x, int64_t poses, const void* saved, const float* dout, void* scratch,
                 const VitGrads& G, float* dx, cudaStream_t st) {
    LIFTER_PROPAGATE(validate_train(d));
    LIFTER_REQUIRE(poses >= 1, "vit_backward: empty batch");
    LIFTER_REQUIRE(x && saved && dout && scratch, "vit_backward: null pointer");
    LIFTER_REQUIRE(G.map && G.w_head1 && G.head_b1 && G.w_head2 && G.head_b2, "vit_backward: null gradient pointer");
    LIFTER_REQUIRE(d.blocks == 0 || (G.ln && G.w_qkv && G.w_out && G.w_mlp1 && G.mlp_b1 && G.w_mlp2 && G.mlp_b2), "vit_backward: null block gradient pointer");
    LIFTER_REQUIRE(reinterpret_cast<uintptr_t>(scratch) % 256 == 0, "vit_backward: scratch must be 256-byte aligned");
    const SavedLayout L = saved_layout(d, poses);
    const BwdScratch B = bwd_scratch(d, poses, scratch);
    const int H = d.hidden, nv = H / 256, dh = H / d.heads, Mm = d.mlp_dim, Hh = d.head_dim, fmt = d.fmt;
    const long T = L.T;
    const int Ti = static_cast<int>(T);
    const unsigned short* S = static_cast<const unsigned short*>(saved);
    const unsigned row_blocks = static_cast<unsigned>(part_rows(T));
    // LIFTER_VIT_FUSE_GELU_BWD=1: d z1 = (dhA W2) * GELU'(z1) in the epilogue of the data-gradient GEMM instead of a pointwise
    // pass.  Measured SLOWER on B200 (3.53 vs 3.50 ms per step): that GEMM has K = 256 and is epilogue-bound already, the
    // GELU' arithmetic costs it +190 us per launch against the 145 us pass it replaces.  Default off.
    const char* fg = getenv("LIFTER_VIT_FUSE_GELU_BWD");
    const bool fuse_gelu_bwd = fg != nullptr && fg[0] == '1' && Mm > 128;
    VitSide* side = vit_side();
    cudaStream_t ws = side != nullptr ? side->st : st;                 // weight-gradient stream
    auto fork = [&]() -> int { return side != nullptr ? order_after(side->fork, st, ws) : LIFTER_OK; };
    auto side_done = [&](cudaEvent_t VitSide::*ev) -> int {            // the side stream is done reading a dZ buffer
        if (side != nullptr) LIFTER_CHECK_CUDA(cudaEventRecord(side->*ev, ws));
        return LIFTER_OK;
    };
    auto wait_side = [&](cudaEvent_t VitSide::*ev) -> int {            // ... before the chain overwrites it
        if (side != nullptr) LIFTER_CHECK_CUDA(cudaStreamWaitEvent(st, side->*ev, 0));
        return LIFTER_OK;
    };
    // ---- head -------------------------------------------------------------------------------------------------------------
    LIFTER_PROPAGATE(prep_grad_out(dout, T, d.out_dim, 64, fmt, B.dout16, B.part, st));
    LIFTER_PROPAGATE(reduce_rows_f32(B.part, ew_partial_rows(T, 64), d.out_dim, 64, G.head_b2, st));
    LIFTER_PROPAGATE(fork());
    LIFTER_PROPAGATE(wgrad16(B.dout16, 64, S + L.r, Hh, Ti, d.out_dim, B.slabs, G.w_head2, fmt, ws));
    LIFTER_PROPAGATE(dgrad16(B.dout16, 64, d.w_head2, Hh, Ti, B.dr, nullptr, fmt, st));
    LIFTER_PROPAGATE(pointwise(S + L.r, B.dr, B.dr, T * Hh, 2, fmt, st));
    LIFTER_PROPAGATE(fork());
    LIFTER_PROPAGATE(colsum16(B.dr, T, Hh, B.part_side, G.head_b1, fmt, ws));
    LIFTER_PROPAGATE(wgrad16(B.dr, Hh, S + L.hlast(), H, Ti, Hh, B.slabs, G.w_head1, fmt, ws));
    LIFTER_PROPAGATE(dgrad16(B.dr, Hh, d.w_head1, H, Ti, B.dhA, nullptr, fmt, st));
    // ---- blocks, last to first; dhA holds d loss / d (block output) -------------------------------------------------------------
    const size_t smem = attb_smem_bytes(d.tokens, dh);
    LIFTER_REQUIRE(d.tokens <= 32, "vit_backward: at most 32 tokens per pose");
    const long pose_heads = poses * d.heads;
    const float att_scale = 1.0f / sqrtf(static_cast<float>(dh));
    for (int b = d.blocks - 1; b >= 0; --b) {
        const bool first = b == d.blocks - 1;
        const float* ln = d.ln + static_cast<long>(b) * 6 * H;
        float* gln = G.ln + static_cast<long>(b) * 6 * H;
        const char* wqkv = static_cast<const char*>(d.w_qkv) + static_cast<long>(b) * 3 * H * H * 2;
        const char* wout = static_cast<const char*>(d.w_out) + static_cast<long>(b) * H * H * 2;
        const char* w1 = static_cast<const char*>(d.w_mlp1) + static_cast<long>(b) * Mm * H * 2;
        const char* w2 = static_cast<const char*>(d.w_mlp2) + static_cast<long>(b) * H * Mm * 2;
        // MLP: hout = h1 + W2 GELU(W1 LN2(h1) + b1) + b2
        LIFTER_PROPAGATE(fork());
        LIFTER_PROPAGATE(colsum16(B.dhA, T, H, B.part_side, G.mlp_b2 + static_cast<long>(b) * H, fmt, ws));
        LIFTER_PROPAGATE(wgrad16(B.dhA, H, S + L.m(b), Mm, Ti, H, B.slabs, G.w_mlp2 + static_cast<long>(b) * H * Mm, fmt, ws));
        LIFTER_PROPAGATE(side_done(&VitSide::dhA));
        if (!first) LIFTER_PROPAGATE(wait_side(&VitSide::dm));
        if (fuse_gelu_bwd) {     // d z1 = (dhA W2) * GELU'(z1) in the epilogue of the data-gradient GEMM
            LIFTER_PROPAGATE(dgrad16(B.dhA, H, w2, Mm, Ti, B.dm, S + L.z1(b), fmt, st, 1));
        } else {
            LIFTER_PROPAGATE(dgrad16(B.dhA, H, w2, Mm, Ti, B.dm, nullptr, fmt, st));
            LIFTER_PROPAGATE(pointwise(S + L.z1(b), B.dm, B.dm, T * Mm, 1, fmt, st));
        }
        LIFTER_PROPAGATE(fork());
        LIFTER_PROPAGATE(colsum16(B.dm, T, Mm, B.part_side, G.mlp_b1 + static_cast<long>(b) * Mm, fmt, ws));
        LIFTER_PROPAGATE(wgrad16(B.dm, Mm, S + L.u2(b), H, Ti, Mm, B.slabs, G.w_mlp1 + static_cast<long>(b) * Mm * H, fmt, ws));
        LIFTER_PROPAGATE(side_done(&VitSide::dm));
        LIFTER_PROPAGATE(dgrad16(B.dm, Mm, w1, H, Ti, B.du, nullptr, fmt, st));
        if (!first) LIFTER_PROPAGATE(wait_side(&VitSide::dhB));
        int lnb = 0;
        LIFTER_PROPAGATE(ln_bwd(nv, S + L.h1(b), B.du, B.dhA, T, ln + 4 * H, ln + 5 * H, nullptr, nullptr, d.ln_eps, B.dhB, B.part,
                                static_cast<int>(row_blocks), fmt, &lnb, st));
        LIFTER_PROPAGATE(reduce_rows_f32(B.part, lnb, 2 * H, 4 * H, gln + 4 * H, st));
        // attention: h1 = hin + Wout attn(Wqkv LN(LN(hin)))
        LIFTER_PROPAGATE(fork());
        LIFTER_PROPAGATE(wgrad16(B.dhB, H, S + L.o(b), H, Ti, H, B.slabs, G.w_out + static_cast<long>(b) * H * H, fmt, ws));
        LIFTER_PROPAGATE(side_done(&VitSide::dhB));
        LIFTER_PROPAGATE(dgrad16(B.dhB, H, wout, H, Ti, B.dO, nullptr, fmt, st));
        if (!first) LIFTER_PROPAGATE(wait_side(&VitSide::dqkv));
        LIFTER_PROPAGATE(attention_bwd(S + L.qkv(b), B.dO, pose_heads, d.tokens, d.heads, dh, att_scale, B.dqkv, fmt, smem, st));
        LIFTER_PROPAGATE(fork());
        LIFTER_PROPAGATE(wgrad16(B.dqkv, 3 * H, S + L.u1(b), H, Ti, 3 * H, B.slabs, G.w_qkv + static_cast<long>(b) * 3 * H * H, fmt, ws));
        LIFTER_PROPAGATE(side_done(&VitSide::dqkv));
        LIFTER_PROPAGATE(dgrad16(B.dqkv, 3 * H, wqkv, H, Ti, B.du, nullptr, fmt, st));
        LIFTER_PROPAGATE(wait_side(&VitSide::dhA));                  // this block's mlp_b2 / w_mlp2 tasks have read dhA
        LIFTER_PROPAGATE(ln_bwd(nv, S + L.hin(b), B.du, B.dhB, T, ln, ln + H, ln + 2 * H, ln + 3 * H, d.ln_eps, B.dhA, B.part,
                                static_cast<int>(row_blocks), fmt, &lnb, st));
        LIFTER_PROPAGATE(reduce_rows_f32(B.part, lnb, 4 * H, 4 * H, gln, st));
    }
    // ---- embedding: d linear_mapper (positional table has requires_grad = False, baselineModel.py:330) -------------------------------
    int emb_blocks = 0;
    VIT_DISPATCH_NV(nv, LIFTER_PROPAGATE((embed_bwd_launch<NV>(B.dhA, x, T, d.in_dim, B.part, static_cast<int>(row_blocks), d.map_w, dx, fmt, &emb_blocks, st))));
    LIFTER_PROPAGATE(reduce_rows_f32(B.part, emb_blocks, (1 + d.in_dim) * H, (1 + d.in_dim) * H, G.map, st));
    return side != nullptr ? order_after(side->join, ws, st) : LIFTER_OK;       // the call is complete on `st`
}

}  // namespace lifter

// ---- C ABI (include/lifter_b200.h) ---------------------------------------------------------------------------------
#include "../../include/lifter_b200.h"

lifter::VitDesc lifter_vit_to_desc(const lifter_vit_t* v);        // vit_ops.cu

namespace lifter {
namespace {
__global__ void vit_map_grad_kernel(const float* __restrict__ gmap, int in_d, int H, float* __restrict__ dw, float* __restrict__ db) {
    griddep_launch();
    griddep_wait();
    const int h = blockIdx.x * blockDim.x + threadIdx.x;
    if (h >= H) return;
    db[h] = gmap[h];
    for (int i = 0; i < in_d; ++i) dw[h * in_d + i] = gmap[(1 + i) * H + h];
}
}  // namespace
}  // namespace lifter

extern "C" {

int64_t lifter_vit_saved_bytes(const lifter_vit_t* vit, int64_t poses) {
    if (vit == nullptr) return LIFTER_ERR_INVALID;
    return lifter::vit_saved_bytes(lifter_vit_to_desc(vit), poses);
}

int64_t lifter_vit_bwd_scratch_bytes(const lifter_vit_t* vit, int64_t poses) {
    if (vit == nullptr) return LIFTER_ERR_INVALID;
    return lifter::vit_bwd_scratch_bytes(lifter_vit_to_desc(vit), poses);
}

int lifter_vit_train_forward(const lifter_vit_t* vit, const float* x, int64_t poses, void* saved, float* out, void* stream) {
    LIFTER_REQUIRE(vit != nullptr, "vit_train_forward: null descriptor");
    return lifter::vit_train_forward(lifter_vit_to_desc(vit), x, poses, saved, out, static_cast<cudaStream_t>(stream));
}

int lifter_vit_backward(const lifter_vit_t* vit, const float* x, int64_t poses, const void* saved, const float* dout, void* scratch,
                        const lifter_vit_grads_t* g, float* dx, void* stream) {
    LIFTER_REQUIRE(vit != nullptr && g != nullptr, "vit_backward: null descriptor");
    lifter::VitGrads G;
    G.map = g->map; G.ln = g->ln; G.w_qkv = g->w_qkv; G.w_out = g->w_out; G.w_mlp1 = g->w_mlp1; G.mlp_b1 = g->mlp_b1;
    G.w_mlp2 = g->w_mlp2; G.mlp_b2 = g->mlp_b2; G.w_head1 = g->w_head1; G.head_b1 = g->head_b1; G.w_head2 = g->w_head2; G.head_b2 = g->head_b2;
    return lifter::vit_backward(lifter_vit_to_desc(vit), x, poses, saved, dout, scratch, G, dx, static_cast<cudaStream_t>(stream));
}

int lifter_vit_map_grad(const float* gmap, int in_dim, int hidden, float* d_weight, float* d_bias, void* stream) {
    LIFTER_REQUIRE(gmap && d_weight && d_bias && in_dim >= 1 && hidden >= 1, "vit_map_grad: bad arguments");
    LIFTER_CHECK_CUDA(lifter::launch_pdl("vit_map_grad_kernel", lifter::vit_map_grad_kernel, dim3((hidden + 255) / 256), dim3(256), 0,
                                         static_cast<cudaStream_t>(stream), gmap, in_dim, hidden, d_weight, d_bias));
    return LIFTER_OK;
}

}  // extern "C"
