// Persistent, warp-specialised tcgen05 GEMM for the lifter's Linear layers (forward, dgrad, wgrad).
//
// Replaces the aten::addmm / cuBLAS sgemm calls under torch.nn.Linear at
// phase1_lifting/baselineModel.py:24-28,69-78, with the elementwise work the
// reference runs as separate kernels (bias, eval-mode BatchNorm affine, ReLU, residual add,
// BatchNorm batch statistics) fused into the epilogue.
//
// Two variants of one kernel:
//   TWO_SM = true  : a CTA PAIR (cluster of 2, one TPC) computes a 256 x BLOCK_N tile with
//                    tcgen05.mma.cta_group::2 (M = 256).  Each CTA stages its own 128 rows of A and HALF of the
//                    B tile; the tensor cores of both SMs read both halves.  Per SM and per K block this moves
//                    16 KB + BLOCK_N/2 * 128 B through shared memory instead of 16 KB + BLOCK_N * 128 B, which
//                    is what bounds the 1-SM variant (measured: operand fill + MMA operand reads saturate the
//                    shared-memory port at ~63 % tensor utilisation; profiles/README.md).
//   TWO_SM = false : one CTA per 128 x BLOCK_N tile (cta_group::1): small / odd problems.
//
// CTA layout (128 + 32*EPI_WARPS threads, one CTA per SM, persistent over output tiles):
//   warp 0         : TMA producer   global -> smem ring (SWIZZLE_128B tiles, 64 K-elements per stage)
//   warp 1         : UMMA issuer    (pair leader only when TWO_SM) 128|256 x BLOCK_N x 16 per instruction,
//                                   fp32 accumulators in TMEM, two accumulator stages (MMA of tile i+1
//                                   overlaps the epilogue of tile i)
//   warp 2         : TMEM allocator
//   warps 4..      : epilogue       EPI_WARPS = 8: two warps per TMEM lane quarter, each owning half of the
//                                   columns.  tcgen05.ld (32 lanes x 32 columns, next chunk in flight while the
//                                   current one is processed) -> affine / ReLU / residual (TMA-prefetched into the
//                                   staging buffer) / column statistics -> swizzled smem staging -> TMA store
// Operands are fp16 or bf16 (runtime flag in the instruction descriptor); both operands may be K-major
// or MN-major (wgrad contracts over the batch, so both of its operands are MN-major).
#include "gemm_sm100.cuh"

#include "bn_device.cuh"

#include <stdlib.h>
#include <string.h>

namespace lifter {

namespace {

constexpr int BLOCK_M = 128;         // rows of A per CTA
constexpr int BLOCK_K = 64;          // 64 x 2 B = one 128 B swizzle atom
constexpr int UMMA_K = 16;
constexpr int A_STAGE_BYTES = BLOCK_M * BLOCK_K * 2;     // 16 KB
constexpr int STAGING_BYTES_PER_WARP = 8192;             // two 32-row x 128 B buffers
constexpr int SMEM_LIMIT = 232448;                       // 227 KB

template <int BLOCK_N, bool TWO_SM>
struct Cfg {
    static constexpr int B_ROWS = TWO_SM ? BLOCK_N / 2 : BLOCK_N;               // rows of the B tile one CTA stages
    static constexpr int B_STAGE_BYTES = B_ROWS * BLOCK_K * 2;
    static constexpr int STAGE_BYTES = A_STAGE_BYTES + B_STAGE_BYTES;
    static constexpr int EPI_WARPS = BLOCK_N >= 128 ? 8 : 4;
    static constexpr int COLS_PER_WARP = BLOCK_N / (EPI_WARPS / 4);
    static constexpr int NCH = COLS_PER_WARP / 32;                              // 32-column chunks per epilogue warp
    static constexpr int NUM_THREADS = 128 + 32 * EPI_WARPS;
    static constexpr int STG_BYTES = EPI_WARPS * STAGING_BYTES_PER_WARP;
    // shared memory: [staging | affine | barriers | TMEM pointer] then the operand ring, A stages followed by B stages.
    // The ring is last so that a launch asking for fewer stages also asks for less shared memory (room for a small
    // kernel of another stream on the same SM).
    static constexpr int MAX_STAGES = 8;
    static constexpr int OFF_STG = 0;
    static constexpr int OFF_AFFINE = OFF_STG + STG_BYTES;                      // scale[BLOCK_N], shift[BLOCK_N]
    static constexpr int OFF_BAR = OFF_AFFINE + 2 * BLOCK_N * 4;                // full[S], empty[S], tfull[2], tempty[2], resid[EPI_WARPS][2]
    static constexpr int NUM_BARS = 2 * MAX_STAGES + 4 + 2 * EPI_WARPS;
    static constexpr int OFF_TMEM_PTR = OFF_BAR + NUM_BARS * 8;
    static constexpr int OFF_RING = (OFF_TMEM_PTR + 16 + 1023) / 1024 * 1024;   // SWIZZLE_128B tiles: 1024 B aligned
    static constexpr int STAGES_FIT = (SMEM_LIMIT - OFF_RING) / STAGE_BYTES;
    static constexpr int STAGES = STAGES_FIT > MAX_STAGES ? MAX_STAGES : STAGES_FIT;
    static constexpr int smem_bytes(int stages) { return OFF_RING + stages * STAGE_BYTES; }
    static constexpr int SMEM_BYTES = OFF_RING + STAGES * STAGE_BYTES;
    static constexpr int TMEM_COLS = 2 * BLOCK_N;                               // 128 / 256 / 512: powers of two
    static_assert(STAGES >= 3, "operand ring too shallow");
    static_assert(SMEM_BYTES <= SMEM_LIMIT, "exceeds 227 KB of shared memory");
    static_assert(NCH >= 1 && NCH <= 4, "epilogue warp owns 1..4 chunks");
};

struct alignas(64) GemmParams {
    CUtensorMap tmA, tmB, tmD, tmR;
    int M, N, K;
    int num_m_blocks, num_n_blocks, splits, k_blocks_total, k_blocks_per_split, num_tiles;   // num_tiles counts M-groups
    int a_mn_major, b_mn_major, fmt, out_mode, relu, has_affine, has_residual, affine_n;
    int resid_mul_gelu_grad;           // GELU instantiation only: result *= GELU'(residual) instead of GELU(result) + residual
    const float* scale;
    const float* shift;
    float* out_direct;
    int n_valid;
    int slab_rows;
    float* stat_sum;
    float* stat_sq;
    const float* resid32;              // OUT_32_TMA: fp32 residual [M][ldr32] (nullptr: none)
    long ldr32;
    unsigned long long* timeline;      // debug hook: [gridDim.x][32] globaltimer stamps (nullptr in production)
    int stages;                        // ring depth in use (<= Cfg::STAGES; lowered only by the debug hook)
    // whole-layer mode (FUSED instantiation only; see FusedBnArgs in gemm_sm100.cuh)
    CUtensorMap tmOut;                 // the activation a [M][N], 64 x 32 boxes like tmD (tmR: the block residual)
    BnFwdFinalize fbn;
    double* fb_sums_out;
    uint8_t* fb_bits;
    unsigned* fb_barrier;
    int fb_stat_rows, fb_has_resid, fb_dropout;
    uint32_t fb_threshold;             // one-bit lanes: 0 keep all, 1 keep where the bit is set, 2 keep none
    float fb_inv_keep;
    PhiloxKeys fb_keys;
    uint32_t fb_site, fb_step, fb_step_mul, fb_step_add;
    const unsigned* fb_step_dev;
    long long fb_row_offset;
};

__device__ __forceinline__ unsigned long long globaltimer_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
#define TL_STAMP(slot)                                                                     \
    do {                                                                                   \
        if (p.timeline != nullptr) p.timeline[blockIdx.x * 32 + (slot)] = globaltimer_ns(); \
    } while (0)

// 64-bit shared-memory matrix descriptor, SWIZZLE_128B, sm_100 version bit set.
//   K-major : rows of 128 B (64 elements of K), 8-row groups 1024 B apart (SBO); LBO unused.
//   MN-major: 64-element MN atoms of [BLOCK_K rows of k][128 B]; 8-k groups 1024 B apart (SBO),
//             atoms along M/N 8192 B apart (LBO).
__device__ __forceinline__ uint64_t make_smem_desc(uint32_t saddr, uint32_t lbo_bytes) {
    uint64_t d = 0;
    d |= static_cast<uint64_t>((saddr & 0x3FFFFu) >> 4);
    d |= static_cast<uint64_t>(lbo_bytes >> 4) << 16;
    d |= static_cast<uint64_t>(1024u >> 4) << 32;
    d |= 1ull << 46;
    d |= 2ull << 61;
    return d;
}

// instruction descriptor, kind::f16: D = fp32, A/B = fmt (0 fp16, 1 bf16), M = m (128, or 256 for the pair), N = n
__device__ __forceinline__ uint32_t make_idesc(int fmt, int a_mn, int b_mn, int m, int n) {
    return (1u << 4) | (static_cast<uint32_t>(fmt) << 7) | (static_cast<uint32_t>(fmt) << 10) |
           (static_cast<uint32_t>(a_mn) << 15) | (static_cast<uint32_t>(b_mn) << 16) |
           (static_cast<uint32_t>(n >> 3) << 17) | (static_cast<uint32_t>(m >> 4) << 24);
}

__device__ __forceinline__ uint4 lds128(uint32_t addr) {
    uint4 r;
    asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "r"(addr));
    return r;
}
__device__ __forceinline__ uint32_t lds32(uint32_t addr) {
    uint32_t r;
    asm volatile("ld.shared.b32 %0, [%1];" : "=r"(r) : "r"(addr));
    return r;
}
__device__ __forceinline__ void sts128(uint32_t addr, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
    asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}

// 128 registers x 384 threads = 48 K of the SM's 64 K registers: the epilogue's two 32-column tcgen05.ld buffers fit
// without spills, and a small kernel on another stream (optimiser range, split-K reduction) can share the SM.
#ifndef LIFTER_GEMM_REGS
#define LIFTER_GEMM_REGS __maxnreg__(128)
#endif
// torch.nn.GELU() (exact erf form) for the epilogue: x * Phi(x) with Phi(-|x|) = erfc(|x| / sqrt 2) / 2 from Abramowitz &
// Stegun 7.1.26 (|error| <= 1.5e-7 on erf): one rcp.approx, one ex2.approx and ten FMA-class operations, against ~25
// for erff() -- with K = 256 the epilogue, not the MMA, bounds the MLP-up GEMM of the transformer lifter.
__device__ __forceinline__ float gelu_erf(float x) {
    const float ax = fabsf(x);
    float t, e;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(t) : "f"(fmaf(ax, 0.3275911f * 0.70710678118654752f, 1.0f)));
    float p = fmaf(0.5f * 1.061405429f, t, 0.5f * -1.453152027f);     // the 1/2 of Phi folded into the coefficients
    p = fmaf(p, t, 0.5f * 1.421413741f);
    p = fmaf(p, t, 0.5f * -0.284496736f);
    p = fmaf(p, t, 0.5f * 0.254829592f);
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(x * x * (-0.5f * 1.4426950408889634f)));
    return fmaf(-ax, p * t * e, fmaxf(x, 0.0f));                     // relu(x) - |x| Phi(-|x|)
}

// d GELU(x) / dx = Phi(x) + x phi(x), same polynomial
__device__ __forceinline__ float gelu_erf_grad(float x) {
    const float ax = fabsf(x);
    float t, e;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(t) : "f"(fmaf(ax, 0.3275911f * 0.70710678118654752f, 1.0f)));
    float p = fmaf(0.5f * 1.061405429f, t, 0.5f * -1.453152027f);
    p = fmaf(p, t, 0.5f * 1.421413741f);
    p = fmaf(p, t, 0.5f * -0.284496736f);
    p = fmaf(p, t, 0.5f * 0.254829592f);
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(x * x * (-0.5f * 1.4426950408889634f)));
    const float tail = p * t * e;                                     // Phi(-|x|)
    return fmaf(x, e * 0.3989422804014327f, x >= 0.0f ? 1.0f - tail : tail);
}

// arrive-and-wait of the epilogue warps of EVERY CTA of the grid (one thread per CTA talks to the counter; the grid is at
// most one CTA per SM, so all CTAs are resident).  `counter` starts at zero; a lost CTA traps after ~4 s instead of hanging.
template <int EPI_THREADS>
__device__ __forceinline__ void epilogue_grid_barrier(unsigned* counter, unsigned nblocks) {
    __threadfence();
    named_bar_sync(1, EPI_THREADS);
    if (threadIdx.x == 128) {
        unsigned seen;
        asm volatile("atom.add.release.gpu.global.u32 %0, [%1], 1;" : "=r"(seen) : "l"(counter) : "memory");
        long long t0 = 0;
        unsigned spins = 0;
        while (true) {
            asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(seen) : "l"(counter) : "memory");
            if (seen >= nblocks) break;
            if ((++spins & 0x3FF) == 0) {
                const long long now = clock64();
                if (t0 == 0) t0 = now;
                else if (now - t0 > 8000000000LL) __trap();
            }
        }
    }
    named_bar_sync(1, EPI_THREADS);
}

template <int BLOCK_N, bool TWO_SM, bool GELU, bool FUSED = false>
__global__ void LIFTER_GEMM_REGS
lifter_gemm_kernel(const __grid_constant__ GemmParams p) {
    using C = Cfg<BLOCK_N, TWO_SM>;
    constexpr int CL = TWO_SM ? 2 : 1;
    extern __shared__ __align__(1024) uint8_t smem[];
    const uint32_t smem_base = smem_u32(smem);
    const uint32_t warp = threadIdx.x >> 5;
    const uint32_t lane = threadIdx.x & 31;

    const uint32_t bar_full = smem_base + C::OFF_BAR;
    const uint32_t bar_empty = bar_full + C::MAX_STAGES * 8;
    const uint32_t bar_tfull = bar_empty + C::MAX_STAGES * 8;
    const uint32_t ring_a = smem_base + C::OFF_RING;                       // p.stages A tiles, then p.stages B tiles
    const uint32_t ring_b = ring_a + static_cast<uint32_t>(p.stages) * A_STAGE_BYTES;
    const uint32_t bar_tempty = bar_tfull + 2 * 8;
    const uint32_t bar_resid = bar_tempty + 2 * 8;                 // [EPI_WARPS][2]
    uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(smem + C::OFF_TMEM_PTR);

    if (threadIdx.x == 0 && (smem_base & 1023u) != 0) __trap();   // SWIZZLE_128B needs 1024 B alignment
    if (threadIdx.x == 0) TL_STAMP(0);

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&p.tmA);
        tma_prefetch_desc(&p.tmB);
        if (p.out_mode != OUT_32_DIRECT) tma_prefetch_desc(&p.tmD);
        if (p.has_residual || (FUSED && p.fb_has_resid)) tma_prefetch_desc(&p.tmR);
        if (FUSED) tma_prefetch_desc(&p.tmOut);
    } else if (warp == 1 && lane == 0) {
        for (int i = 0; i < p.stages; ++i) {
            mbar_init(bar_full + i * 8, 1);            // the (leader's) producer arms it; TMA bytes complete it
            mbar_init(bar_empty + i * 8, 1);           // one MMA-commit arrival
        }
        for (int i = 0; i < 2; ++i) {
            mbar_init(bar_tfull + i * 8, 1);
            mbar_init(bar_tempty + i * 8, C::EPI_WARPS * CL);     // one arrival per epilogue warp (of both CTAs)
        }
        for (int i = 0; i < 2 * C::EPI_WARPS; ++i) mbar_init(bar_resid + i * 8, 1);
        fence_mbar_init();
    } else if (warp == 2) {
        if (TWO_SM) tmem_alloc_2sm<C::TMEM_COLS>(smem_u32(tmem_ptr_smem));
        else tmem_alloc<C::TMEM_COLS>(smem_u32(tmem_ptr_smem));
    }
    griddep_launch();                                         // let the next kernel's prologue overlap our tail
    tcgen05_fence_before();
    if (TWO_SM) cluster_sync_all(); else __syncthreads();     // the peer's barriers are initialised before any remote arrive
    tcgen05_fence_after();
    griddep_wait();                                           // everything above touched no global data
    const uint32_t tmem_base = *tmem_ptr_smem;
    if (threadIdx.x == 0) TL_STAMP(1);

    const int tiles_per_m = p.num_n_blocks * p.splits;
    const uint32_t crank = TWO_SM ? cluster_ctarank() : 0;
    const int first_tile = blockIdx.x / CL, tile_step = gridDim.x / CL;

    if (warp == 0) {
        if (elect_one_sync()) {
            // ===================== TMA producer =====================
            uint32_t stage = 0, phase = 0;
            for (int tile = first_tile; tile < p.num_tiles; tile += tile_step) {
                const int m_grp = tile / tiles_per_m;
                const int m_blk = m_grp * CL + crank;
                const int rem = tile - m_grp * tiles_per_m;
                const int split = rem / p.num_n_blocks;
                const int n_blk = rem - split * p.num_n_blocks;
                const int n0 = n_blk * BLOCK_N + crank * C::B_ROWS;           // my part of the B tile
                const int kb0 = split * p.k_blocks_per_split;
                const int kb1 = min(kb0 + p.k_blocks_per_split, p.k_blocks_total);
                for (int kb = kb0; kb < kb1; ++kb) {
                    mbar_wait(bar_empty + stage * 8, phase ^ 1);
                    const uint32_t full = bar_full + stage * 8;
                    // pair: only the leader arms its barrier, with the bytes of both CTAs
                    if (!TWO_SM) mbar_arrive_expect_tx(full, C::STAGE_BYTES);
                    else if (crank == 0) mbar_arrive_expect_tx(full, 2 * C::STAGE_BYTES);
                    const uint32_t sa = ring_a + stage * A_STAGE_BYTES;
                    const uint32_t sb = ring_b + stage * C::B_STAGE_BYTES;
                    if (!p.a_mn_major) {
                        if (TWO_SM) tma_load_2d_2sm(&p.tmA, sa, full, kb * BLOCK_K, m_blk * BLOCK_M);
                        else tma_load_2d(&p.tmA, sa, full, kb * BLOCK_K, m_blk * BLOCK_M);
                    } else {
#pragma unroll
                        for (int j = 0; j < BLOCK_M / 64; ++j) {
                            if (TWO_SM) tma_load_2d_2sm(&p.tmA, sa + j * 8192, full, m_blk * BLOCK_M + j * 64, kb * BLOCK_K);
                            else tma_load_2d(&p.tmA, sa + j * 8192, full, m_blk * BLOCK_M + j * 64, kb * BLOCK_K);
                        }
                    }
                    if (!p.b_mn_major) {
                        if (TWO_SM) tma_load_2d_2sm(&p.tmB, sb, full, kb * BLOCK_K, n0);
                        else tma_load_2d(&p.tmB, sb, full, kb * BLOCK_K, n0);
                    } else {
#pragma unroll
                        for (int j = 0; j < C::B_ROWS / 64; ++j) {
                            if (TWO_SM) tma_load_2d_2sm(&p.tmB, sb + j * 8192, full, n0 + j * 64, kb * BLOCK_K);
                            else tma_load_2d(&p.tmB, sb + j * 8192, full, n0 + j * 64, kb * BLOCK_K);
                        }
                    }
                    if (++stage == static_cast<uint32_t>(p.stages)) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        // ===================== UMMA issuer (the pair's leader only) =====================
        // The whole warp walks the pipeline (waits are warp-wide); one elected lane issues the MMAs and commits.
        if (!TWO_SM || crank == 0) {
            const uint32_t idesc = make_idesc(p.fmt, p.a_mn_major, p.b_mn_major, BLOCK_M * CL, BLOCK_N);
            const uint32_t a_kstep = (p.a_mn_major ? 2048u : 32u) >> 4, b_kstep = (p.b_mn_major ? 2048u : 32u) >> 4;
            const uint64_t da0 = make_smem_desc(ring_a, p.a_mn_major ? 8192u : 0u);
            const uint64_t db0 = make_smem_desc(ring_b, p.b_mn_major ? 8192u : 0u);
            const uint32_t a_hi = static_cast<uint32_t>(da0 >> 32), b_hi = static_cast<uint32_t>(db0 >> 32);
            const uint32_t a_lo0 = static_cast<uint32_t>(da0), b_lo0 = static_cast<uint32_t>(db0);
            uint32_t stage = 0, phase = 0;
            int it = 0;
            for (int tile = first_tile; tile < p.num_tiles; tile += tile_step, ++it) {
                const int rem = tile % tiles_per_m;
                const int split = rem / p.num_n_blocks;
                const int kb0 = split * p.k_blocks_per_split;
                const int kb1 = min(kb0 + p.k_blocks_per_split, p.k_blocks_total);
                const uint32_t as = it & 1, aph = (it >> 1) & 1;
                mbar_wait(bar_tempty + as * 8, aph ^ 1);               // epilogue(s) drained this accumulator
                tcgen05_fence_after();
                const uint32_t tmem_d = tmem_base + as * BLOCK_N;
                for (int kb = kb0; kb < kb1; ++kb) {
                    mbar_wait(bar_full + stage * 8, phase);            // TMA bytes (of both CTAs) landed
                    tcgen05_fence_after();
                    if (elect_one_sync()) {
                        const uint32_t a_lo = a_lo0 + stage * (A_STAGE_BYTES >> 4);
                        const uint32_t b_lo = b_lo0 + stage * (C::B_STAGE_BYTES >> 4);
#pragma unroll
                        for (int k = 0; k < BLOCK_K / UMMA_K; ++k) {
                            const uint64_t da = (static_cast<uint64_t>(a_hi) << 32) | (a_lo + k * a_kstep);
                            const uint64_t db = (static_cast<uint64_t>(b_hi) << 32) | (b_lo + k * b_kstep);
                            const uint32_t acc = (kb > kb0 || k > 0) ? 1u : 0u;
                            if (TWO_SM) umma_f16_2sm(tmem_d, da, db, idesc, acc);
                            else umma_f16(tmem_d, da, db, idesc, acc);
                        }
                        // frees the smem slot (in both CTAs of a pair) when these MMAs retire
                        if (TWO_SM) umma_commit_2sm(bar_empty + stage * 8, 0x3);
                        else umma_commit(bar_empty + stage * 8);
                        if (kb == kb1 - 1) {                               // accumulator complete -> epilogue(s)
                            if (TWO_SM) umma_commit_2sm(bar_tfull + as * 8, 0x3);
                            else umma_commit(bar_tfull + as * 8);
                        }
                        if (kb == kb0 && it == 0) TL_STAMP(2);
                        if (kb == kb1 - 1 && it < 8) TL_STAMP(4 + 3 * it);
                    }
                    __syncwarp();
                    if (++stage == static_cast<uint32_t>(p.stages)) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp >= 4) {
        // ===================== epilogue =====================
        const uint32_t ew = warp - 4;
        const uint32_t q = warp & 3;                               // TMEM lane quarter: lanes [32 q, 32 q + 32)
        const uint32_t h = ew >> 2;                                // column half (always 0 with four epilogue warps)
        const int col_base = static_cast<int>(h) * C::COLS_PER_WARP;   // first tile column owned by this warp
        uint8_t* stg = smem + C::OFF_STG + ew * STAGING_BYTES_PER_WARP;
        const uint32_t stg_addr = smem_u32(stg);
        const uint32_t rbar = bar_resid + ew * 16;                 // residual barriers of my two staging buffers
        float* s_scale = reinterpret_cast<float*>(smem + C::OFF_AFFINE);
        float* s_shift = s_scale + BLOCK_N;
        const int fmt = p.fmt;
        uint32_t buf = 0;                                          // staging buffer the next pair / chunk uses
        uint32_t rph = 0;                                          // bit b: phase of residual barrier b
        int cached_n_blk = -1;
        int it = 0;
        for (int tile = first_tile; tile < p.num_tiles; tile += tile_step, ++it) {
            const int m_grp = tile / tiles_per_m;
            const int m_blk = m_grp * CL + crank;                  // may lie past the last M-block (odd count): all
            const int rem = tile - m_grp * tiles_per_m;            // rows are then >= M and every store is clipped
            const int split = rem / p.num_n_blocks;
            const int n_blk = rem - split * p.num_n_blocks;
            const uint32_t as = it & 1, aph = (it >> 1) & 1;

            if (p.has_affine && n_blk != cached_n_blk) {           // uniform over the epilogue warps
                named_bar_sync(1, 32 * C::EPI_WARPS);
                for (int i = ew * 32 + lane; i < BLOCK_N; i += 32 * C::EPI_WARPS) {
                    const int n = n_blk * BLOCK_N + i;
                    s_scale[i] = (p.scale != nullptr && n < p.affine_n) ? p.scale[n] : 1.0f;
                    s_shift[i] = (p.shift != nullptr && n < p.affine_n) ? p.shift[n] : 0.0f;
                }
                named_bar_sync(1, 32 * C::EPI_WARPS);
                cached_n_blk = n_blk;
            }

            const int row0 = m_blk * BLOCK_M + q * 32;             // first row of this warp's slab
            const bool tile_ok = m_blk < p.num_m_blocks;
            const int rows_valid = min(max(p.M - row0, 0), 32);
            const int ncol0 = n_blk * BLOCK_N + col_base;          // first global column of this warp

            // residual tiles of this warp's 64-column pairs: TMA into the staging buffers, ahead of the accumulator
            if (p.has_residual) {
                if (lane == 0) {
                    tma_store_wait_read<0>();                      // the buffers' previous stores have read them
#pragma unroll
                    for (int j = 0; j < C::NCH / 2; ++j) {
                        const uint32_t b = (buf + j) & 1u;
                        mbar_arrive_expect_tx(rbar + b * 8, 4096);
                        tma_load_2d(&p.tmR, stg_addr + b * 4096, rbar + b * 8, ncol0 + j * 64, row0);
                    }
                }
                __syncwarp();
            }

            mbar_wait(bar_tfull + as * 8, aph);
            tcgen05_fence_after();
            if (threadIdx.x == 128 && it < 8) TL_STAMP(5 + 3 * it);

            const uint32_t taddr = tmem_base + ((q * 32u) << 16) + as * BLOCK_N + col_base;
            float* stg_f = reinterpret_cast<float*>(stg);
            float v[2][32];
            tmem_ld_32x32(taddr, v[0]);
            tmem_ld_wait();

#pragma unroll
            for (int c = 0; c < C::NCH; ++c) {
                float (&x)[32] = v[c & 1];
                if (c + 1 < C::NCH) {
                    tmem_ld_32x32(taddr + (c + 1) * 32, v[(c + 1) & 1]);      // in flight while chunk c is processed
                } else if (!FUSED) {                                        // accumulator fully read: hand TMEM back
                    tcgen05_fence_before();
                    __syncwarp();
                    if (lane == 0) {
                        if (!TWO_SM || crank == 0) mbar_arrive(bar_tempty + as * 8);
                        else mbar_arrive_remote(bar_tempty + as * 8, 0);
                    }
                }                                                           // (FUSED: the tile stays in TMEM for pass 2)
                const int n0 = ncol0 + c * 32;
                if (p.has_affine) {
#pragma unroll
                    for (int j = 0; j < 32; ++j)
                        x[j] = fmaf(x[j], s_scale[col_base + c * 32 + j], s_shift[col_base + c * 32 + j]);
                }
                if constexpr (GELU) {          // torch.nn.GELU(); its own instantiation: the lifter's kernels carry none of it
                    if (p.relu == 2) {
#pragma unroll
                        for (int j = 0; j < 32; ++j) x[j] = gelu_erf(x[j]);
                    }
                } else if (p.relu) {
#pragma unroll
                    for (int j = 0; j < 32; ++j) x[j] = fmaxf(x[j], 0.0f);
                }

                if (p.out_mode == OUT_16_TMA) {
                    if constexpr (GELU) {      // fp32 residual stream folded into a 16-bit output (last block of the MyViT eval path)
                        if (p.resid32 != nullptr && tile_ok && static_cast<int>(lane) < rows_valid && n0 + 32 <= p.N) {
                            const float* rr = p.resid32 + static_cast<long>(row0 + lane) * p.ldr32 + n0;
#pragma unroll
                            for (int qq = 0; qq < 8; ++qq) {
                                const float4 r4 = *reinterpret_cast<const float4*>(rr + 4 * qq);
                                x[4 * qq] += r4.x; x[4 * qq + 1] += r4.y; x[4 * qq + 2] += r4.z; x[4 * qq + 3] += r4.w;
                            }
                        }
                    }
                    const uint32_t half = c & 1;
                    const uint32_t dst = stg_addr + buf * 4096 + lane * 128;
                    if (half == 0) {
                        if (p.has_residual) {                      // residual tile of this pair has landed
                            mbar_wait(rbar + buf * 8, (rph >> buf) & 1u);
                            rph ^= 1u << buf;
                        } else {                                   // buffer reuse: its previous TMA store has read it
                            if (lane == 0) tma_store_wait_read<1>();
                            __syncwarp();
                        }
                    }
                    if (p.has_residual) {
#pragma unroll
                        for (int qq = 0; qq < 4; ++qq) {
                            const uint4 r = lds128(dst + (((half * 4 + qq) ^ (lane & 7)) * 16));
                            const uint32_t w[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
                            for (int e = 0; e < 4; ++e) {
                                const float2 f = unpack16(w[e], fmt);
                                if (GELU && p.resid_mul_gelu_grad) {
                                    x[qq * 8 + e * 2] *= gelu_erf_grad(f.x);
                                    x[qq * 8 + e * 2 + 1] *= gelu_erf_grad(f.y);
                                } else {
                                    x[qq * 8 + e * 2] += f.x;
                                    x[qq * 8 + e * 2 + 1] += f.y;
                                }
                            }
                        }
                    }
                    uint32_t pk[16];
#pragma unroll
                    for (int j = 0; j < 16; ++j) pk[j] = pack16(x[2 * j], x[2 * j + 1], fmt);
#pragma unroll
                    for (int qq = 0; qq < 4; ++qq)                  // SWIZZLE_128B: 16-byte chunk index ^ (row & 7)
                        sts128(dst + (((half * 4 + qq) ^ (lane & 7)) * 16), pk[4 * qq], pk[4 * qq + 1], pk[4 * qq + 2],
                               pk[4 * qq + 3]);
                    if (half == 1) {
                        if (p.stat_sum != nullptr) {               // column sums of the values actually stored
                            __syncwarp();
                            const uint32_t src = stg_addr + buf * 4096 + (lane & 3) * 4;
                            const uint32_t ch = lane >> 2;
                            float s0 = 0.0f, s1 = 0.0f, q0 = 0.0f, q1 = 0.0f;
                            for (int r = 0; r < rows_valid; ++r) {
                                const float2 f = unpack16(lds32(src + r * 128 + ((ch ^ (r & 7)) * 16)), fmt);
                                s0 += f.x; s1 += f.y;
                                q0 = fmaf(f.x, f.x, q0); q1 = fmaf(f.y, f.y, q1);
                            }
                            const int n = ncol0 + (c >> 1) * 64 + 2 * lane;
                            if (tile_ok) {
                                const long o = static_cast<long>(m_blk * 4 + q) * p.N + n;
                                if (n < p.N) { p.stat_sum[o] = s0; p.stat_sq[o] = q0; }
                                if (n + 1 < p.N) { p.stat_sum[o + 1] = s1; p.stat_sq[o + 1] = q1; }
                            }
                        }
                        fence_proxy_async_smem();
                        __syncwarp();
                        if (lane == 0 && tile_ok) {
                            tma_store_2d(&p.tmD, stg_addr + buf * 4096, ncol0 + (c >> 1) * 64, row0);
                            tma_store_commit();
                        }
                        buf ^= 1;
                    }
                } else if (p.out_mode == OUT_32_TMA) {
                    if (p.resid32 != nullptr && tile_ok && static_cast<int>(lane) < rows_valid) {
                        // fp32 residual stream (MyViT eval path): this lane's row, 32 consecutive columns = one 128-byte line
                        const float* rr = p.resid32 + static_cast<long>(row0 + lane) * p.ldr32 + n0;
                        if (n0 + 32 <= p.N) {
#pragma unroll
                            for (int qq = 0; qq < 8; ++qq) {
                                const float4 r4 = *reinterpret_cast<const float4*>(rr + 4 * qq);
                                x[4 * qq] += r4.x; x[4 * qq + 1] += r4.y; x[4 * qq + 2] += r4.z; x[4 * qq + 3] += r4.w;
                            }
                        } else {
#pragma unroll
                            for (int j = 0; j < 32; ++j) if (n0 + j < p.N) x[j] += rr[j];
                        }
                    }
                    if (lane == 0) tma_store_wait_read<1>();
                    __syncwarp();
                    const uint32_t dst = stg_addr + buf * 4096 + lane * 128;
#pragma unroll
                    for (int qq = 0; qq < 8; ++qq)
                        sts128(dst + ((qq ^ (lane & 7)) * 16), __float_as_uint(x[4 * qq]), __float_as_uint(x[4 * qq + 1]),
                               __float_as_uint(x[4 * qq + 2]), __float_as_uint(x[4 * qq + 3]));
                    fence_proxy_async_smem();
                    __syncwarp();
                    if (lane == 0 && tile_ok) {
                        tma_store_2d(&p.tmD, stg_addr + buf * 4096, n0, split * p.slab_rows + row0);
                        tma_store_commit();
                    }
                    buf ^= 1;
                } else {                                           // OUT_32_DIRECT: dense [32][n_valid] slab in smem
#pragma unroll
                    for (int j = 0; j < 32; ++j) {
                        const int n = col_base + c * 32 + j;
                        if (n < p.n_valid) stg_f[lane * p.n_valid + n] = x[j];
                    }
                }
                if (c + 1 < C::NCH) tmem_ld_wait();                // chunk c + 1 is in registers
            }

            if (p.out_mode == OUT_32_DIRECT) {                     // rows are contiguous in global: linear copy-out
                __syncwarp();
                const int limit = rows_valid * p.n_valid;
                float* gout = p.out_direct + static_cast<long>(row0) * p.n_valid;
                for (int i = lane; i < limit; i += 32) gout[i] = stg_f[i];
                __syncwarp();
            }
            if (threadIdx.x == 128 && it < 8) TL_STAMP(6 + 3 * it);
        }
        if constexpr (FUSED) {
            constexpr int EPI_THREADS = 32 * C::EPI_WARPS;
            // ================= column finalize, shared out over the grid =================
            if (lane == 0) tma_store_wait_read<0>();               // pass 1's z stores have read the staging buffers
            if (threadIdx.x == 128) TL_STAMP(27);                  // pass 1 done in this CTA
            epilogue_grid_barrier<EPI_THREADS>(p.fb_barrier, gridDim.x);       // every partial row of the layer is written
            if (threadIdx.x == 128) TL_STAMP(28);
            {
                double (*sh)[32][8] = reinterpret_cast<double (*)[32][8]>(smem + C::OFF_STG);   // [2][32][8], 4 KB
                const int tid = static_cast<int>(threadIdx.x) - 128;
                const int cc = tid & 7, lane_r = tid >> 3;         // 8 columns x 32 row lanes
                const BnFwdFinalize& f = p.fbn;
                for (int cb = blockIdx.x; cb * 8 < p.N; cb += gridDim.x) {
                    const int col = cb * 8 + cc;
                    BnColumnInputs ci = {0.0f, 1.0f, 0.0f, 0.0f, 1.0f};
                    if (lane_r == 0 && col < p.N) ci = bn_column_inputs(f, col);
                    double sa = 0.0, sb = 0.0;
                    if (col < p.N) {
                        for (int r = lane_r; r < p.fb_stat_rows; r += 32) {
                            sa += static_cast<double>(ldg_f32(p.stat_sum + static_cast<long>(r) * p.N + col));
                            sb += static_cast<double>(ldg_f32(p.stat_sq + static_cast<long>(r) * p.N + col));
                        }
                    }
                    sh[0][lane_r][cc] = sa;
                    sh[1][lane_r][cc] = sb;
                    named_bar_sync(1, EPI_THREADS);
#pragma unroll
                    for (int st = 16; st >= 1; st >>= 1) {         // fixed pairing: deterministic
                        if (lane_r < st) {
                            sh[0][lane_r][cc] += sh[0][lane_r + st][cc];
                            sh[1][lane_r][cc] += sh[1][lane_r + st][cc];
                        }
                        named_bar_sync(1, EPI_THREADS);
                    }
                    double ta = sh[0][0][cc], tb = sh[1][0][cc];
                    named_bar_sync(1, EPI_THREADS);                // sh is reused by the next column group
                    if (f.ex.world > 1 && lane_r < 4) exchange_column_sums_warp(f.ex, p.N, col, col < p.N, lane_r, ta, tb);
                    if (lane_r == 0 && col < p.N) {
                        p.fb_sums_out[col] = ta;
                        p.fb_sums_out[p.N + col] = tb;
                        if (col == 0 && f.num_batches_tracked != nullptr) *f.num_batches_tracked += 1;
                        bn_finalize_fwd_col(f, col, ta, tb, ci);
                    }
                }
            }
            if (threadIdx.x == 128) TL_STAMP(29);                  // this CTA's share of the column finalize done
            epilogue_grid_barrier<EPI_THREADS>(p.fb_barrier + 1, gridDim.x);   // scale / shift of every column exist
            if (threadIdx.x == 128) TL_STAMP(30);

            // ================= pass 2: BatchNorm + ReLU + dropout (+ residual) straight from TMEM =================
            const uint32_t dstep = (p.fb_step_dev != nullptr ? *p.fb_step_dev : p.fb_step) * p.fb_step_mul + p.fb_step_add;
            const int bits_ld = p.N >> 3;
            it = 0;
            for (int tile = first_tile; tile < p.num_tiles; tile += tile_step, ++it) {
                const int m_grp = tile / tiles_per_m;
                const int m_blk = m_grp * CL + crank;
                const int n_blk = tile - m_grp * tiles_per_m;      // splits == 1
                const uint32_t as = it & 1;
                named_bar_sync(1, EPI_THREADS);                    // everyone is done with the previous tile's coefficients
                for (int i = ew * 32 + lane; i < BLOCK_N; i += EPI_THREADS) {
                    const int n = n_blk * BLOCK_N + i;             // 1 / (1 - p) folded in (positive: ReLU unchanged)
                    s_scale[i] = n < p.N ? __ldcg(p.fbn.scale + n) * p.fb_inv_keep : 0.0f;
                    s_shift[i] = n < p.N ? __ldcg(p.fbn.shift + n) * p.fb_inv_keep : 0.0f;
                }
                named_bar_sync(1, EPI_THREADS);
                const int row0 = m_blk * BLOCK_M + q * 32;
                const bool tile_ok = m_blk < p.num_m_blocks;
                const int rows_valid = min(max(p.M - row0, 0), 32);
                const int ncol0 = n_blk * BLOCK_N + col_base;
                if (p.fb_has_resid) {                              // block input tiles ahead of the arithmetic
                    if (lane == 0) {
                        tma_store_wait_read<0>();
#pragma unroll
                        for (int j = 0; j < C::NCH / 2; ++j) {
                            const uint32_t b = (buf + j) & 1u;
                            mbar_arrive_expect_tx(rbar + b * 8, 4096);
                            tma_load_2d(&p.tmR, stg_addr + b * 4096, rbar + b * 8, ncol0 + j * 64, row0);
                        }
                    }
                    __syncwarp();
                }
                // dropout stream (philox.cuh, one-bit lanes): a call covers 16 rows x 8 columns; the warp's 32 rows are two
                // row groups, a 32-column chunk four column groups.  Lane L computes the call of (row group L / 16, column
                // group (L / 4) % 4) and keeps word L % 4 of its 128 bits; row r's byte for column group k is byte r % 4 of
                // word (r / 4) % 4 of call (r / 16, k), i.e. it sits on lane (L & 16) | (k << 2) | ((L >> 2) & 3).
                const uint32_t grow = static_cast<uint32_t>(row0 + static_cast<int>(lane) + p.fb_row_offset);
                const uint32_t taddr = tmem_base + ((q * 32u) << 16) + as * BLOCK_N + col_base;
                float v[2][32];
                tmem_ld_32x32(taddr, v[0]);
                tmem_ld_wait();
#pragma unroll
                for (int c = 0; c < C::NCH; ++c) {
                    float (&x)[32] = v[c & 1];
                    if (c + 1 < C::NCH) tmem_ld_32x32(taddr + (c + 1) * 32, v[(c + 1) & 1]);
                    const int n0 = ncol0 + c * 32;
                    uint32_t keep = 0xFFFFFFFFu;
                    if (p.fb_dropout) {
                        if (p.fb_threshold == 1u) {
                            const Philox4 r4 = philox4x32_10(grow >> 4, static_cast<uint32_t>(n0 >> 3) + ((lane >> 2) & 3u),
                                                             p.fb_site, dstep, p.fb_keys);
                            const uint32_t sel = lane & 3u;
                            const uint32_t mine = sel == 0 ? r4.x : (sel == 1 ? r4.y : (sel == 2 ? r4.z : r4.w));
                            keep = 0u;
#pragma unroll
                            for (uint32_t k = 0; k < 4; ++k) {
                                const uint32_t word = __shfl_sync(0xffffffffu, mine, (lane & 16u) | (k << 2) | ((lane >> 2) & 3u));
                                keep |= ((word >> ((lane & 3u) * 8u)) & 0xFFu) << (8u * k);
                            }
                        } else if (p.fb_threshold != 0u) {
                            keep = 0u;
                        }
                    }
                    uint32_t active = 0u;
#pragma unroll
                    for (int j = 0; j < 32; j += 2) {              // z as the 16-bit value pass 1 stored (backward reads that)
                        const float2 zr = unpack16(pack16(x[j], x[j + 1], fmt), fmt);
                        const float y0 = fmaf(zr.x, s_scale[col_base + c * 32 + j], s_shift[col_base + c * 32 + j]);
                        const float y1 = fmaf(zr.y, s_scale[col_base + c * 32 + j + 1], s_shift[col_base + c * 32 + j + 1]);
                        const bool on0 = ((keep >> j) & 1u) && (y0 > 0.0f), on1 = ((keep >> (j + 1)) & 1u) && (y1 > 0.0f);
                        active |= (on0 ? 1u : 0u) << j;
                        active |= (on1 ? 1u : 0u) << (j + 1);
                        x[j] = on0 ? y0 : 0.0f;
                        x[j + 1] = on1 ? y1 : 0.0f;
                    }
                    if (tile_ok && static_cast<int>(lane) < rows_valid && n0 < p.N)
                        *reinterpret_cast<uint32_t*>(p.fb_bits + static_cast<long>(row0 + lane) * bits_ld + (n0 >> 3)) = active;
                    const uint32_t half = c & 1;
                    const uint32_t dst = stg_addr + buf * 4096 + lane * 128;
                    if (half == 0) {
                        if (p.fb_has_resid) {
                            mbar_wait(rbar + buf * 8, (rph >> buf) & 1u);
                            rph ^= 1u << buf;
                        } else {
                            if (lane == 0) tma_store_wait_read<1>();
                            __syncwarp();
                        }
                    }
                    if (p.fb_has_resid) {
#pragma unroll
                        for (int qq = 0; qq < 4; ++qq) {
                            const uint4 r = lds128(dst + (((half * 4 + qq) ^ (lane & 7)) * 16));
                            const uint32_t w[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
                            for (int e = 0; e < 4; ++e) {
                                const float2 f2 = unpack16(w[e], fmt);
                                x[qq * 8 + e * 2] += f2.x;
                                x[qq * 8 + e * 2 + 1] += f2.y;
                            }
                        }
                    }
                    uint32_t pk[16];
#pragma unroll
                    for (int j = 0; j < 16; ++j) pk[j] = pack16(x[2 * j], x[2 * j + 1], fmt);
#pragma unroll
                    for (int qq = 0; qq < 4; ++qq)
                        sts128(dst + (((half * 4 + qq) ^ (lane & 7)) * 16), pk[4 * qq], pk[4 * qq + 1], pk[4 * qq + 2],
                               pk[4 * qq + 3]);
                    if (half == 1) {
                        fence_proxy_async_smem();
                        __syncwarp();
                        if (lane == 0 && tile_ok) {
                            tma_store_2d(&p.tmOut, stg_addr + buf * 4096, ncol0 + (c >> 1) * 64, row0);
                            tma_store_commit();
                        }
                        buf ^= 1;
                    }
                    if (c + 1 < C::NCH) tmem_ld_wait();
                }
            }
            if (threadIdx.x == 128) TL_STAMP(31);                  // pass 2 issued (its bulk stores may still be reading smem)
        }
        // smem must outlive the READS of the last bulk stores; their global writes complete with the grid (what the
        // next kernel's griddepcontrol.wait / stream order waits for), so the CTA need not sit through them
        if (lane == 0) tma_store_wait_read<0>();
    }

    tcgen05_fence_before();
    if (TWO_SM) cluster_sync_all(); else __syncthreads();     // no CTA exits while its peer may still signal it
    if (warp == 2) {
        tcgen05_fence_after();
        if (TWO_SM) tmem_dealloc_2sm<C::TMEM_COLS>(tmem_base);
        else tmem_dealloc<C::TMEM_COLS>(tmem_base);
    }
}

// ------------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn get_encode_fn() {
    static EncodeTiledFn fn = nullptr;
    if (fn == nullptr) {
        void* ptr = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &qres) == cudaSuccess &&
            qres == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(ptr);
    }
    return fn;
}

// 2D row-major tensor [rows][cols] (cols contiguous, row stride ld elements), box = box_cols x box_rows
int encode_2d(CUtensorMap* map, CUtensorMapDataType dt, int elem_bytes, const void* base, long rows, long cols,
              long ld, int box_cols, int box_rows) {
    EncodeTiledFn fn = get_encode_fn();
    if (fn == nullptr) {
        set_error("cuTensorMapEncodeTiled is not available from the driver");
        return LIFTER_ERR_CUDA;
    }
    LIFTER_REQUIRE((reinterpret_cast<uintptr_t>(base) & 15) == 0, "TMA base pointer must be 16-byte aligned");
    LIFTER_REQUIRE((ld * elem_bytes) % 16 == 0, "TMA row stride (%ld bytes) must be a multiple of 16", ld * elem_bytes);
    LIFTER_REQUIRE(box_cols * elem_bytes == 128, "TMA box must span one 128-byte swizzle atom");
    cuuint64_t gdim[2] = {static_cast<cuuint64_t>(cols), static_cast<cuuint64_t>(rows)};
    cuuint64_t gstr[1] = {static_cast<cuuint64_t>(ld) * elem_bytes};
    cuuint32_t box[2] = {static_cast<cuuint32_t>(box_cols), static_cast<cuuint32_t>(box_rows)};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = fn(map, dt, 2, const_cast<void*>(base), gdim, gstr, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                    CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        set_error("cuTensorMapEncodeTiled failed with CUresult %d (rows=%ld cols=%ld ld=%ld box=%dx%d)",
                  static_cast<int>(r), rows, cols, ld, box_cols, box_rows);
        return LIFTER_ERR_CUDA;
    }
    return LIFTER_OK;
}

template <int BLOCK_N, bool TWO_SM, bool GELU = false, bool FUSED = false>
int launch_impl(GemmParams& p, const GemmArgs& a, cudaStream_t stream) {
    using C = Cfg<BLOCK_N, TWO_SM>;
    constexpr int CL = TWO_SM ? 2 : 1;
    static PerDeviceOnce once;
    bool& configured = *once.slot();
    if (!configured) {
        LIFTER_CHECK_CUDA(cudaFuncSetAttribute(lifter_gemm_kernel<BLOCK_N, TWO_SM, GELU, FUSED>,
                                               cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM_BYTES));
        configured = true;
    }
    const CUtensorMapDataType dt16 =
        a.fmt == FMT_BF16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16;
    // operand maps (a pair's CTA loads half of the B tile)
    if (!a.a_mn_major) LIFTER_PROPAGATE(encode_2d(&p.tmA, dt16, 2, a.A, a.M, a.K, a.lda, BLOCK_K, BLOCK_M));
    else               LIFTER_PROPAGATE(encode_2d(&p.tmA, dt16, 2, a.A, a.K, a.M, a.lda, 64, BLOCK_K));
    if (!a.b_mn_major) LIFTER_PROPAGATE(encode_2d(&p.tmB, dt16, 2, a.B, a.N, a.K, a.ldb, BLOCK_K, C::B_ROWS));
    else               LIFTER_PROPAGATE(encode_2d(&p.tmB, dt16, 2, a.B, a.K, a.N, a.ldb, 64, BLOCK_K));
    if (a.out_mode == OUT_16_TMA) {
        LIFTER_PROPAGATE(encode_2d(&p.tmD, dt16, 2, a.D, a.M, a.N, a.ldd, 64, 32));
        if (a.residual != nullptr) LIFTER_PROPAGATE(encode_2d(&p.tmR, dt16, 2, a.residual, a.M, a.N, a.ldr, 64, 32));
        if (FUSED) {
            const FusedBnArgs& fb = *a.fused_bn;
            LIFTER_PROPAGATE(encode_2d(&p.tmOut, dt16, 2, fb.a_out, a.M, a.N, a.N, 64, 32));
            if (fb.resid != nullptr) LIFTER_PROPAGATE(encode_2d(&p.tmR, dt16, 2, fb.resid, a.M, a.N, fb.ldr, 64, 32));
        }
    } else if (a.out_mode == OUT_32_TMA) {
        // rows of the last slab are clipped to M, so a single-slab output may be exactly [M][ldd]
        LIFTER_PROPAGATE(encode_2d(&p.tmD, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, a.D,
                                   static_cast<long>(p.splits - 1) * p.slab_rows + a.M, a.N, a.ldd, 32, 32));
    }
    p.num_n_blocks = ceil_div(a.N, BLOCK_N);
    p.stages = (a.max_stages > 0 && a.max_stages < C::STAGES) ? a.max_stages : C::STAGES;
    p.num_tiles = ceil_div(p.num_m_blocks, CL) * p.num_n_blocks * p.splits;
    cudaLaunchConfig_t cfg = {};
    cfg.blockDim = dim3(C::NUM_THREADS);
    cfg.dynamicSmemBytes = C::smem_bytes(p.stages);
    cfg.stream = stream;
    cudaLaunchAttribute attr[2];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = CL; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[1].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    const int sms = (a.max_ctas > 0 && a.max_ctas < num_sms()) ? a.max_ctas : num_sms();
    const int max_clusters = sms / CL < 1 ? 1 : sms / CL;                       // a pair shares a TPC: 74 pairs on 148 SMs
    const int clusters = p.num_tiles < max_clusters ? p.num_tiles : max_clusters;
    cfg.gridDim = dim3(clusters * CL);
    cfg.numAttrs = pdl_enabled() ? 2 : 1;
    if (FUSED)      // every tile must stay in TMEM across the grid barriers: at most two per pair, all pairs resident
        LIFTER_REQUIRE(p.num_tiles <= 2 * clusters && a.max_ctas == 0, "gemm: %d tiles do not fit the fused layer mode", p.num_tiles);
    LIFTER_CHECK_CUDA(cudaLaunchKernelEx(&cfg, lifter_gemm_kernel<BLOCK_N, TWO_SM, GELU, FUSED>, p));
    if (a.tag != nullptr) trace_mark(a.tag, stream); else trace_mark("gemm", stream);
    return LIFTER_OK;
}

}  // namespace

int gemm_slab_rows(int M) { return ceil_div(M, BLOCK_M) * BLOCK_M; }
int gemm_effective_splits(int K, int splits) {
    const int kb = ceil_div(K, BLOCK_K);
    int s = splits < kb ? splits : kb;
    if (s < 1) s = 1;
    const int per = ceil_div(kb, s);
    return ceil_div(kb, per);
}
int gemm_stat_rows(int M) { return ceil_div(M, BLOCK_M) * 4; }

// whole-layer mode: CTA pairs, 256-wide column blocks, at most two tiles per pair, one-bit dropout lanes on 16-row groups
bool gemm_fused_bn_fits(int M, int N, float p, int apply_dropout, long long row_offset) {
    if (M < 2 * BLOCK_M || N < 256 || N % 256 != 0) return false;
    const int pairs = num_sms() / 2;
    const long tiles = static_cast<long>(ceil_div(ceil_div(M, BLOCK_M), 2)) * (N / 256);
    if (tiles > 2l * pairs) return false;
    if (apply_dropout && p > 0.0f) {
        if (dropout_lanes(p).bits != 1 || row_offset % 16 != 0) return false;
    }
    return true;
}

int gemm_launch(const GemmArgs& a, cudaStream_t stream) {
    LIFTER_REQUIRE(a.M > 0 && a.N > 0 && a.K > 0, "gemm: empty problem (M=%d N=%d K=%d)", a.M, a.N, a.K);
    LIFTER_REQUIRE(a.fmt == FMT_F16 || a.fmt == FMT_BF16, "gemm: unknown operand format %d", a.fmt);
    LIFTER_REQUIRE(a.A && a.B, "gemm: null operand");
    LIFTER_REQUIRE(a.splits >= 1, "gemm: splits must be >= 1");
    LIFTER_REQUIRE(a.splits == 1 || a.out_mode == OUT_32_TMA, "gemm: split-K needs the fp32 slab output");
    LIFTER_REQUIRE(a.D != nullptr, "gemm: null output");
    if (a.out_mode == OUT_32_DIRECT)
        LIFTER_REQUIRE(a.n_valid > 0 && a.n_valid <= 64 && a.N <= 64, "gemm: direct fp32 output holds at most 64 columns");
    LIFTER_REQUIRE(a.stat_sum == nullptr || (a.out_mode == OUT_16_TMA && a.stat_sq != nullptr),
                   "gemm: column statistics ride on the 16-bit output");
    LIFTER_REQUIRE(a.residual32 == nullptr || (((a.out_mode == OUT_32_TMA && a.splits == 1) ||
                                                (a.out_mode == OUT_16_TMA && a.N > 128 && a.N % 32 == 0 && a.residual == nullptr)) &&
                                               a.ldr32 % 4 == 0 && (reinterpret_cast<uintptr_t>(a.residual32) & 15) == 0),
                   "gemm: the fp32 residual rides on the single-slab fp32 output, or on a 16-bit output with N > 128 and N % 32 == 0, "
                   "and needs 16-byte aligned rows");
    LIFTER_REQUIRE(a.residual == nullptr || (a.out_mode == OUT_16_TMA && a.ldr % 8 == 0),
                   "gemm: the residual rides on the 16-bit output and needs a 16-byte row stride");

    GemmParams p;
    memset(&p, 0, sizeof(p));
    p.M = a.M; p.N = a.N; p.K = a.K;
    p.num_m_blocks = ceil_div(a.M, BLOCK_M);
    p.k_blocks_total = ceil_div(a.K, BLOCK_K);
    p.splits = gemm_effective_splits(a.K, a.splits);               // no empty split
    p.k_blocks_per_split = ceil_div(p.k_blocks_total, p.splits);
    p.a_mn_major = a.a_mn_major; p.b_mn_major = a.b_mn_major;
    p.fmt = a.fmt; p.out_mode = a.out_mode; p.relu = a.relu;
    p.has_affine = (a.scale != nullptr || a.shift != nullptr) ? 1 : 0;
    p.has_residual = a.residual != nullptr ? 1 : 0;
    LIFTER_REQUIRE(a.residual_op == 0 || (a.residual_op == 1 && a.residual != nullptr && a.relu == 0 && a.N > 128),
                   "gemm: residual_op 1 (multiply by GELU'(residual)) needs a residual, no activation and N > 128");
    p.resid_mul_gelu_grad = a.residual_op == 1 ? 1 : 0;
    p.scale = a.scale; p.shift = a.shift;
    p.affine_n = (a.shift_valid > 0 && a.shift_valid < a.N) ? a.shift_valid : a.N;
    p.out_direct = static_cast<float*>(a.D); p.n_valid = a.n_valid;
    p.slab_rows = gemm_slab_rows(a.M);
    p.stat_sum = a.stat_sum; p.stat_sq = a.stat_sq;
    p.resid32 = a.residual32; p.ldr32 = a.ldr32;
    p.timeline = a.timeline;

    if (a.fused_bn != nullptr) {
        const FusedBnArgs& fb = *a.fused_bn;
        LIFTER_REQUIRE(a.out_mode == OUT_16_TMA && a.stat_sum != nullptr && a.splits == 1 && a.residual == nullptr &&
                       a.scale == nullptr && a.shift == nullptr && a.relu == 0 && a.cluster != 1,
                       "gemm: the fused layer mode takes the plain z = A W^T launch with column statistics");
        LIFTER_REQUIRE(gemm_fused_bn_fits(a.M, a.N, fb.p, fb.apply_dropout, fb.row_offset), "gemm: shape does not fit the fused layer mode");
        LIFTER_REQUIRE(fb.a_out && fb.bits && fb.barrier && fb.sums_out && fb.f.scale && fb.f.shift && fb.f.mean_z && fb.f.rstd,
                       "gemm: fused layer mode: null pointer");
        LIFTER_REQUIRE(fb.f.use_bn && fb.f.bn_mode == 0 && fb.f.H == a.N, "gemm: the fused layer mode is batch-statistics BatchNorm");
        p.fbn = fb.f; p.fb_sums_out = fb.sums_out; p.fb_bits = fb.bits; p.fb_barrier = fb.barrier;
        p.fb_stat_rows = gemm_stat_rows(a.M); p.fb_has_resid = fb.resid != nullptr ? 1 : 0;
        const bool drop = fb.apply_dropout && fb.p > 0.0f;
        p.fb_dropout = drop ? 1 : 0;
        p.fb_threshold = drop ? dropout_lanes(fb.p).threshold : 0u;
        p.fb_inv_keep = drop ? (fb.p < 1.0f ? 1.0f / (1.0f - fb.p) : 0.0f) : 1.0f;
        p.fb_keys = philox_round_keys(fb.seed);
        p.fb_site = fb.site; p.fb_step = fb.step; p.fb_step_mul = fb.step_mul; p.fb_step_add = fb.step_add;
        p.fb_step_dev = fb.step_dev; p.fb_row_offset = fb.row_offset;
        return launch_impl<256, true, false, true>(p, a, stream);
    }
    // CTA pairs (cta_group::2, 256-row tiles) whenever there are two M-blocks to pair up.
    // a.cluster: 0 = auto, 1 = single-CTA tiles, 2 = pairs (when the shape allows).
    const bool pair = a.cluster != 1 && p.num_m_blocks >= 2 && a.N > 64;
    LIFTER_REQUIRE(a.relu >= 0 && a.relu <= 2, "gemm: activation %d (0 none, 1 ReLU, 2 GELU)", a.relu);
    if (a.relu == 2 || a.residual_op == 1 || (a.residual32 != nullptr && a.out_mode == OUT_16_TMA)) {
        // the transformer lifter's instantiation: GELU epilogue of the MLP-up projection (forward), GELU' of its backward, fp32 residual on a 16-bit output
        LIFTER_REQUIRE(a.N > 128, "gemm: the GELU epilogue is built for N > 128 (got %d)", a.N);      // or its backward
        return pair ? launch_impl<256, true, true>(p, a, stream) : launch_impl<256, false, true>(p, a, stream);
    }
    if (a.N <= 64) return launch_impl<64, false>(p, a, stream);
    if (a.N <= 128) return pair ? launch_impl<128, true>(p, a, stream) : launch_impl<128, false>(p, a, stream);
    // 256-wide column blocks halve the B-operand fill per FLOP, but the LAST tile's epilogue is exposed (4-5 us of a
    // 17.7 us launch at M = 8,192).  With few tiles per CTA pair, 128-wide blocks quantise into waves the same way
    // (1.73 -> 2 waves of 256 = 3.46 -> 4 waves of 128) and leave half the tail.  LIFTER_GEMM_NARROW: 0 never, 1 always
    // (pairs), default = when a pair gets at most `narrow_below` 256-wide tiles.
    static const int narrow = [] { const char* e = getenv("LIFTER_GEMM_NARROW"); return e != nullptr ? atoi(e) : -1; }();
    if (pair && a.out_mode == OUT_16_TMA && narrow == 1) return launch_impl<128, true>(p, a, stream);
    return pair ? launch_impl<256, true>(p, a, stream) : launch_impl<256, false>(p, a, stream);
}

}  // namespace lifter
