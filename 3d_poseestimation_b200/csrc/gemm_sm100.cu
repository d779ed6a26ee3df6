// Persistent, warp-specialised tcgen05 GEMM for the lifter's Linear layers (forward, dgrad, wgrad).
//
// Replaces the aten::addmm / cuBLAS sgemm calls under torch.nn.Linear at
// phase1_lifting/baselineModel.py:24-28,69-78, with the elementwise work the
// reference runs as separate kernels (bias, eval-mode BatchNorm affine, ReLU, residual add,
// BatchNorm batch statistics) fused into the epilogue.
//
// CTA layout (256 threads, one CTA per SM, persistent over output tiles):
//   warp 0  lane 0 : TMA producer   global -> smem ring (SWIZZLE_128B tiles, 64 K-elements per stage)
//   warp 1  lane 0 : UMMA issuer    tcgen05.mma.cta_group::1.kind::f16, 128 x BLOCK_N x 16 per instruction,
//                                   fp32 accumulators in TMEM, two accumulator stages (MMA of tile i+1
//                                   overlaps the epilogue of tile i)
//   warp 2         : TMEM allocator
//   warps 4..7     : epilogue       tcgen05.ld (32 lanes x 32 columns per warp) -> affine / ReLU / residual /
//                                   column statistics -> swizzled smem staging -> TMA store
// Operands are fp16 or bf16 (runtime flag in the instruction descriptor); both operands may be K-major
// or MN-major (wgrad contracts over the batch, so both of its operands are MN-major).
#include "gemm_sm100.cuh"

#include <stdlib.h>
#include <string.h>

namespace lifter {

namespace {

constexpr int BLOCK_M = 128;
constexpr int BLOCK_K = 64;          // 64 x 2 B = one 128 B swizzle atom
constexpr int UMMA_K = 16;
constexpr int A_STAGE_BYTES = BLOCK_M * BLOCK_K * 2;     // 16 KB
constexpr int STAGING_BYTES_PER_WARP = 8192;             // two 32-row x 128 B buffers
constexpr int NUM_THREADS = 256;

template <int BLOCK_N>
struct Cfg {
    static constexpr int B_STAGE_BYTES = BLOCK_N * BLOCK_K * 2;
    static constexpr int STAGE_BYTES = A_STAGE_BYTES + B_STAGE_BYTES;
    static constexpr int STAGES = (196608 / STAGE_BYTES) > 8 ? 8 : (196608 / STAGE_BYTES);
    static constexpr int OFF_A = 0;
    static constexpr int OFF_B = OFF_A + STAGES * A_STAGE_BYTES;
    static constexpr int OFF_STG = OFF_B + STAGES * B_STAGE_BYTES;
    static constexpr int OFF_AFFINE = OFF_STG + 4 * STAGING_BYTES_PER_WARP;     // scale[BLOCK_N], shift[BLOCK_N]
    static constexpr int OFF_BAR = OFF_AFFINE + 2 * BLOCK_N * 4;                // full[S], empty[S], tfull[2], tempty[2]
    static constexpr int OFF_TMEM_PTR = OFF_BAR + (2 * STAGES + 4) * 8;
    static constexpr int SMEM_BYTES = OFF_TMEM_PTR + 16;
    static constexpr int TMEM_COLS = 2 * BLOCK_N;                               // 128 / 256 / 512: powers of two
    static_assert(SMEM_BYTES <= 232448, "exceeds 227 KB of shared memory");
};

struct alignas(64) GemmParams {
    CUtensorMap tmA, tmB, tmD;
    int M, N, K;
    int num_m_blocks, num_n_blocks, splits, k_blocks_total, k_blocks_per_split, num_tiles;   // num_tiles counts M-groups of CL blocks
    int a_mn_major, b_mn_major, fmt, out_mode, relu, has_affine;
    const float* scale;
    const float* shift;
    const uint16_t* residual;
    long ldr;
    float* out_direct;
    int n_valid;
    int slab_rows;
    float* stat_sum;
    float* stat_sq;
    unsigned long long* timeline;      // debug hook: [gridDim.x][32] globaltimer stamps (nullptr in production)
    int dbg;                           // debug hook: 1 = MMAs without operand loads, 2 = operand loads without MMAs
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

// instruction descriptor, kind::f16: D = fp32, A/B = fmt (0 fp16, 1 bf16), M = 128, N = BLOCK_N
__device__ __forceinline__ uint32_t make_idesc(int fmt, int a_mn, int b_mn, int n) {
    return (1u << 4) | (static_cast<uint32_t>(fmt) << 7) | (static_cast<uint32_t>(fmt) << 10) |
           (static_cast<uint32_t>(a_mn) << 15) | (static_cast<uint32_t>(b_mn) << 16) |
           (static_cast<uint32_t>(n >> 3) << 17) | (static_cast<uint32_t>(BLOCK_M >> 4) << 24);
}

// Sum over the 32 lanes of v[j] for every j; lane l returns the total of column l.
__device__ __forceinline__ float warp_transpose_sum(float (&v)[32], uint32_t lane) {
#pragma unroll
    for (int s = 16; s >= 1; s >>= 1) {
        const bool upper = (lane & s) != 0;
#pragma unroll
        for (int j = 0; j < s; ++j) {
            const float keep = upper ? v[j + s] : v[j];
            const float send = upper ? v[j] : v[j + s];
            v[j] = keep + __shfl_xor_sync(0xffffffffu, send, s);
        }
    }
    return v[0];
}

// CL = CTAs per cluster (1, 2, 4 or 8).  The CTAs of a cluster work on vertically adjacent M-blocks of the same
// N-block and K-range: each loads its own A tile and 1/CL of the shared B tile, multicast into every CTA's smem, so
// the L2-side operand traffic per stage drops from 16 KB + B to 16 KB + B/CL per SM (the L2 -> SM fabric, ~9.5 TB/s
// measured, is what bounds these GEMMs: profiles/README.md).  The MMAs stay per-CTA (cta_group::1); only the
// smem-slot release is cluster-wide (each CTA's commit arrives on the empty barrier of every CTA).
template <int BLOCK_N, int CL>
__global__ void __launch_bounds__(NUM_THREADS, 1) lifter_gemm_kernel(const __grid_constant__ GemmParams p) {
    using C = Cfg<BLOCK_N>;
    static_assert(CL == 1 || ((CL == 2 || CL == 4 || CL == 8) && BLOCK_N / CL >= 32),
                  "cluster multicast splits the B tile in CL parts of >= 32 rows");
    constexpr uint16_t CL_MASK = static_cast<uint16_t>((1u << CL) - 1u);
    extern __shared__ __align__(1024) uint8_t smem[];
    const uint32_t smem_base = smem_u32(smem);
    const uint32_t warp = threadIdx.x >> 5;
    const uint32_t lane = threadIdx.x & 31;

    const uint32_t bar_full = smem_base + C::OFF_BAR;
    const uint32_t bar_empty = bar_full + C::STAGES * 8;
    const uint32_t bar_tfull = bar_empty + C::STAGES * 8;
    const uint32_t bar_tempty = bar_tfull + 2 * 8;
    uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(smem + C::OFF_TMEM_PTR);

    if (threadIdx.x == 0 && (smem_base & 1023u) != 0) __trap();   // SWIZZLE_128B needs 1024 B alignment
    if (threadIdx.x == 0) TL_STAMP(0);

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&p.tmA);
        tma_prefetch_desc(&p.tmB);
        if (p.out_mode != OUT_32_DIRECT) tma_prefetch_desc(&p.tmD);
    } else if (warp == 1 && lane == 0) {
        for (int i = 0; i < C::STAGES; ++i) {
            mbar_init(bar_full + i * 8, 1);
            mbar_init(bar_empty + i * 8, CL);          // one MMA-commit arrival per CTA of the cluster
        }
        for (int i = 0; i < 2; ++i) {
            mbar_init(bar_tfull + i * 8, 1);
            mbar_init(bar_tempty + i * 8, 4);      // one arrival per epilogue warp
        }
        fence_mbar_init();
    } else if (warp == 2) {
        tmem_alloc<C::TMEM_COLS>(smem_u32(tmem_ptr_smem));
    }
    griddep_launch();                                         // let the next kernel's prologue overlap our tail
    tcgen05_fence_before();
    if (CL > 1) cluster_sync_all(); else __syncthreads();     // peers' barriers are initialised before any remote arrive
    tcgen05_fence_after();
    griddep_wait();                                           // everything above touched no global data
    const uint32_t tmem_base = *tmem_ptr_smem;
    if (threadIdx.x == 0) TL_STAMP(1);

    const int tiles_per_m = p.num_n_blocks * p.splits;
    const uint32_t crank = CL > 1 ? cluster_ctarank() : 0;
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
                const int kb0 = split * p.k_blocks_per_split;
                const int kb1 = min(kb0 + p.k_blocks_per_split, p.k_blocks_total);
                for (int kb = kb0; kb < kb1; ++kb) {
                    if (p.dbg == 1) continue;
                    mbar_wait(bar_empty + stage * 8, phase ^ 1);
                    const uint32_t full = bar_full + stage * 8;
                    mbar_arrive_expect_tx(full, C::STAGE_BYTES);
                    const uint32_t sa = smem_base + C::OFF_A + stage * A_STAGE_BYTES;
                    const uint32_t sb = smem_base + C::OFF_B + stage * C::B_STAGE_BYTES;
                    if (!p.a_mn_major) {
                        tma_load_2d(&p.tmA, sa, full, kb * BLOCK_K, m_blk * BLOCK_M);
                    } else {
#pragma unroll
                        for (int j = 0; j < BLOCK_M / 64; ++j)
                            tma_load_2d(&p.tmA, sa + j * 8192, full, m_blk * BLOCK_M + j * 64, kb * BLOCK_K);
                    }
                    if (CL == 1) {
                        if (!p.b_mn_major) {
                            tma_load_2d(&p.tmB, sb, full, kb * BLOCK_K, n_blk * BLOCK_N);
                        } else {
#pragma unroll
                            for (int j = 0; j < BLOCK_N / 64; ++j)
                                tma_load_2d(&p.tmB, sb + j * 8192, full, n_blk * BLOCK_N + j * 64, kb * BLOCK_K);
                        }
                    } else {                                       // my part of B, delivered to every CTA
                        constexpr int PART = BLOCK_N / CL;
                        if (!p.b_mn_major) {
                            tma_load_2d_multicast(&p.tmB, sb + crank * (PART * 128), full, kb * BLOCK_K,
                                                  n_blk * BLOCK_N + crank * PART, CL_MASK);
                        } else if (PART >= 64) {                   // (MN-major B moves in 64-column atoms)
#pragma unroll
                            for (int j = 0; j < PART / 64; ++j) {
                                const int atom = crank * (PART / 64) + j;
                                tma_load_2d_multicast(&p.tmB, sb + atom * 8192, full, n_blk * BLOCK_N + atom * 64,
                                                      kb * BLOCK_K, CL_MASK);
                            }
                        }
                    }
                    if (++stage == C::STAGES) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        // ===================== UMMA issuer =====================
        // The whole warp walks the pipeline (waits are warp-wide); one elected lane issues the MMAs and commits.
        const uint32_t idesc = make_idesc(p.fmt, p.a_mn_major, p.b_mn_major, BLOCK_N);
        const uint32_t a_kstep = (p.a_mn_major ? 2048u : 32u) >> 4, b_kstep = (p.b_mn_major ? 2048u : 32u) >> 4;
        const uint64_t da0 = make_smem_desc(smem_base + C::OFF_A, p.a_mn_major ? 8192u : 0u);
        const uint64_t db0 = make_smem_desc(smem_base + C::OFF_B, p.b_mn_major ? 8192u : 0u);
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
            mbar_wait(bar_tempty + as * 8, aph ^ 1);               // epilogue drained this accumulator
            tcgen05_fence_after();
            const uint32_t tmem_d = tmem_base + as * BLOCK_N;
            for (int kb = kb0; kb < kb1; ++kb) {
                if (p.dbg != 1) mbar_wait(bar_full + stage * 8, phase);            // TMA bytes landed
                tcgen05_fence_after();
                if (p.dbg == 2) {
                    if (elect_one_sync()) {
                        if (CL == 1) mbar_arrive(bar_empty + stage * 8);
                        else for (uint32_t r = 0; r < CL; ++r) mbar_arrive_remote(bar_empty + stage * 8, r);
                        if (kb == kb1 - 1) umma_commit(bar_tfull + as * 8);
                        if (kb == kb0 && it == 0) TL_STAMP(2);
                        if (kb == kb1 - 1 && it < 8) TL_STAMP(4 + 3 * it);
                    }
                } else if (elect_one_sync()) {
                    const uint32_t a_lo = a_lo0 + stage * (A_STAGE_BYTES >> 4);
                    const uint32_t b_lo = b_lo0 + stage * (C::B_STAGE_BYTES >> 4);
#pragma unroll
                    for (int k = 0; k < BLOCK_K / UMMA_K; ++k) {
                        const uint64_t da = (static_cast<uint64_t>(a_hi) << 32) | (a_lo + k * a_kstep);
                        const uint64_t db = (static_cast<uint64_t>(b_hi) << 32) | (b_lo + k * b_kstep);
                        umma_f16(tmem_d, da, db, idesc, (kb > kb0 || k > 0) ? 1u : 0u);
                    }
                    // frees the smem slot (in every CTA of the cluster) when these MMAs retire
                    if (CL == 1) umma_commit(bar_empty + stage * 8);
                    else umma_commit_multicast(bar_empty + stage * 8, CL_MASK);
                    if (kb == kb1 - 1) umma_commit(bar_tfull + as * 8);    // accumulator complete -> epilogue
                    if (kb == kb0 && it == 0) TL_STAMP(2);
                    if (kb == kb1 - 1 && it < 8) TL_STAMP(4 + 3 * it);
                }
                __syncwarp();
                if (++stage == C::STAGES) { stage = 0; phase ^= 1; }
            }
        }
    } else if (warp >= 4) {
        // ===================== epilogue =====================
        const uint32_t ew = warp - 4;                              // == warp % 4: TMEM lanes [32 ew, 32 ew + 32)
        uint8_t* stg = smem + C::OFF_STG + ew * STAGING_BYTES_PER_WARP;
        const uint32_t stg_addr = smem_u32(stg);
        float* s_scale = reinterpret_cast<float*>(smem + C::OFF_AFFINE);
        float* s_shift = s_scale + BLOCK_N;
        const int fmt = p.fmt;
        uint32_t buf = 0;
        int cached_n_blk = -1;
        int it = 0;
        for (int tile = first_tile; tile < p.num_tiles; tile += tile_step, ++it) {
            const int m_grp = tile / tiles_per_m;
            const int m_blk = m_grp * CL + crank;                  // may lie past the last M-block (odd count): all
            const int rem = tile - m_grp * tiles_per_m;            // rows are then >= M and every store is clipped
            const int split = rem / p.num_n_blocks;
            const int n_blk = rem - split * p.num_n_blocks;
            const uint32_t as = it & 1, aph = (it >> 1) & 1;

            if (p.has_affine && n_blk != cached_n_blk) {           // uniform over the four epilogue warps
                named_bar_sync(1, 128);
                for (int i = ew * 32 + lane; i < BLOCK_N; i += 128) {
                    const int n = n_blk * BLOCK_N + i;
                    s_scale[i] = (p.scale != nullptr && n < p.N) ? p.scale[n] : 1.0f;
                    s_shift[i] = (p.shift != nullptr && n < p.N) ? p.shift[n] : 0.0f;
                }
                named_bar_sync(1, 128);
                cached_n_blk = n_blk;
            }

            mbar_wait(bar_tfull + as * 8, aph);
            tcgen05_fence_after();
            if (threadIdx.x == 128 && it < 8) TL_STAMP(5 + 3 * it);

            const int row0 = m_blk * BLOCK_M + ew * 32;            // first row of this warp's slab
            const int row = row0 + lane;
            const bool row_ok = row < p.M;
            const bool tile_ok = m_blk < p.num_m_blocks;
            const uint32_t taddr = tmem_base + ((ew * 32u) << 16) + as * BLOCK_N;
            float* stg_f = reinterpret_cast<float*>(stg);

#pragma unroll 1
            for (int c = 0; c < BLOCK_N / 32; ++c) {
                float v[32];
                tmem_ld_32x32(taddr + c * 32, v);
                tmem_ld_wait();
                if (c == BLOCK_N / 32 - 1) {                       // accumulator fully read: hand TMEM back
                    tcgen05_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(bar_tempty + as * 8);
                }
                const int n0 = n_blk * BLOCK_N + c * 32;
                if (p.has_affine) {
#pragma unroll
                    for (int j = 0; j < 32; ++j) v[j] = fmaf(v[j], s_scale[c * 32 + j], s_shift[c * 32 + j]);
                }
                if (p.relu) {
#pragma unroll
                    for (int j = 0; j < 32; ++j) v[j] = fmaxf(v[j], 0.0f);
                }
                if (p.residual != nullptr && row_ok && n0 < p.N) {
                    const uint4* rp = reinterpret_cast<const uint4*>(p.residual + static_cast<long>(row) * p.ldr + n0);
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        const uint4 r = __ldg(rp + q);
                        const uint32_t w[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
                        for (int e = 0; e < 4; ++e) {
                            const float2 f = unpack16(w[e], fmt);
                            v[q * 8 + e * 2] += f.x;
                            v[q * 8 + e * 2 + 1] += f.y;
                        }
                    }
                }

                if (p.out_mode == OUT_16_TMA) {
                    uint32_t pk[16];
#pragma unroll
                    for (int j = 0; j < 16; ++j) pk[j] = pack16(v[2 * j], v[2 * j + 1], fmt);
                    const uint32_t half = c & 1;
                    if (half == 0) {                               // buffer reuse: its previous TMA store has read it
                        if (lane == 0) tma_store_wait_read<1>();
                        __syncwarp();
                    }
                    const uint32_t dst = stg_addr + buf * 4096 + lane * 128;
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        const uint32_t chunk = (half * 4 + q) ^ (lane & 7);     // SWIZZLE_128B
                        asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(dst + chunk * 16),
                                     "r"(pk[4 * q]), "r"(pk[4 * q + 1]), "r"(pk[4 * q + 2]), "r"(pk[4 * q + 3])
                                     : "memory");
                    }
                    if (p.stat_sum != nullptr) {                   // statistics of the values actually stored
                        float sq[32];
#pragma unroll
                        for (int j = 0; j < 16; ++j) {
                            float2 f = unpack16(pk[j], fmt);
                            if (!row_ok) { f.x = 0.0f; f.y = 0.0f; }
                            v[2 * j] = f.x; v[2 * j + 1] = f.y;
                            sq[2 * j] = f.x * f.x; sq[2 * j + 1] = f.y * f.y;
                        }
                        const float cs = warp_transpose_sum(v, lane);
                        const float cq = warp_transpose_sum(sq, lane);
                        const int n = n0 + lane;
                        if (n < p.N && tile_ok) {
                            const long o = static_cast<long>(m_blk * 4 + ew) * p.N + n;
                            p.stat_sum[o] = cs;
                            p.stat_sq[o] = cq;
                        }
                    }
                    if (half == 1) {
                        fence_proxy_async_smem();
                        __syncwarp();
                        if (lane == 0 && tile_ok) {
                            tma_store_2d(&p.tmD, stg_addr + buf * 4096, n_blk * BLOCK_N + (c >> 1) * 64, row0);
                            tma_store_commit();
                        }
                        buf ^= 1;
                    }
                } else if (p.out_mode == OUT_32_TMA) {
                    if (lane == 0) tma_store_wait_read<1>();
                    __syncwarp();
                    const uint32_t dst = stg_addr + buf * 4096 + lane * 128;
#pragma unroll
                    for (int q = 0; q < 8; ++q) {
                        const uint32_t chunk = q ^ (lane & 7);
                        asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(dst + chunk * 16),
                                     "r"(__float_as_uint(v[4 * q])), "r"(__float_as_uint(v[4 * q + 1])),
                                     "r"(__float_as_uint(v[4 * q + 2])), "r"(__float_as_uint(v[4 * q + 3]))
                                     : "memory");
                    }
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
                        const int n = c * 32 + j;
                        if (n < p.n_valid) stg_f[lane * p.n_valid + n] = v[j];
                    }
                }
            }

            if (p.out_mode == OUT_32_DIRECT) {                     // rows are contiguous in global: linear copy-out
                __syncwarp();
                const int rows_valid = min(max(p.M - row0, 0), 32);
                const int limit = rows_valid * p.n_valid;
                float* gout = p.out_direct + static_cast<long>(row0) * p.n_valid;
                for (int i = lane; i < limit; i += 32) gout[i] = stg_f[i];
                __syncwarp();
            }
            if (threadIdx.x == 128 && it < 8) TL_STAMP(6 + 3 * it);
        }
        if (lane == 0) tma_store_wait_all<0>();                    // smem must outlive the last bulk store
    }

    tcgen05_fence_before();
    if (CL > 1) cluster_sync_all(); else __syncthreads();     // no CTA exits while its peer may still signal it
    if (warp == 2) {
        tcgen05_fence_after();
        tmem_dealloc<C::TMEM_COLS>(tmem_base);
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

template <int BLOCK_N, int CL>
int launch_impl(GemmParams& p, const GemmArgs& a, cudaStream_t stream) {
    using C = Cfg<BLOCK_N>;
    static bool configured = false;
    if (!configured) {
        LIFTER_CHECK_CUDA(cudaFuncSetAttribute(lifter_gemm_kernel<BLOCK_N, CL>,
                                               cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM_BYTES));
        configured = true;
    }
    const CUtensorMapDataType dt16 =
        a.fmt == FMT_BF16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16;
    // operand maps (with CL = 2 every CTA loads half of the K-major B tile)
    if (!a.a_mn_major) LIFTER_PROPAGATE(encode_2d(&p.tmA, dt16, 2, a.A, a.M, a.K, a.lda, BLOCK_K, BLOCK_M));
    else               LIFTER_PROPAGATE(encode_2d(&p.tmA, dt16, 2, a.A, a.K, a.M, a.lda, 64, BLOCK_K));
    if (!a.b_mn_major) LIFTER_PROPAGATE(encode_2d(&p.tmB, dt16, 2, a.B, a.N, a.K, a.ldb, BLOCK_K, BLOCK_N / CL));
    else               LIFTER_PROPAGATE(encode_2d(&p.tmB, dt16, 2, a.B, a.K, a.N, a.ldb, 64, BLOCK_K));
    if (a.out_mode == OUT_16_TMA) {
        LIFTER_PROPAGATE(encode_2d(&p.tmD, dt16, 2, a.D, a.M, a.N, a.ldd, 64, 32));
    } else if (a.out_mode == OUT_32_TMA) {
        // rows of the last slab are clipped to M, so a single-slab output may be exactly [M][ldd]
        LIFTER_PROPAGATE(encode_2d(&p.tmD, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, a.D,
                                   static_cast<long>(p.splits - 1) * p.slab_rows + a.M, a.N, a.ldd, 32, 32));
    }
    p.num_n_blocks = ceil_div(a.N, BLOCK_N);
    p.num_tiles = ceil_div(p.num_m_blocks, CL) * p.num_n_blocks * p.splits;
    cudaLaunchConfig_t cfg = {};
    cfg.blockDim = dim3(NUM_THREADS);
    cfg.dynamicSmemBytes = C::SMEM_BYTES;
    cfg.stream = stream;
    cudaLaunchAttribute attr[2];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = CL; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[1].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    // clusters that can be co-resident (GPC boundaries may leave a few SMs out for CL > 1); queried once per variant
    static int max_clusters = 0;
    if (max_clusters == 0) {
        int n = 0;
        cfg.gridDim = dim3(num_sms() / CL * CL);
        if (CL == 1 || cudaOccupancyMaxActiveClusters(&n, lifter_gemm_kernel<BLOCK_N, CL>, &cfg) != cudaSuccess || n <= 0)
            n = num_sms() / CL;
        (void)cudaGetLastError();
        max_clusters = n < num_sms() / CL ? n : num_sms() / CL;
    }
    const int clusters = p.num_tiles < max_clusters ? p.num_tiles : max_clusters;
    cfg.gridDim = dim3(clusters * CL);
    cfg.numAttrs = pdl_enabled() ? 2 : 1;
    LIFTER_CHECK_CUDA(cudaLaunchKernelEx(&cfg, lifter_gemm_kernel<BLOCK_N, CL>, p));
    return LIFTER_OK;
}

}  // namespace

// default cluster size of long launches (LIFTER_GEMM_CLUSTER=1/2/4/8 overrides it for experiments)
static int gemm_auto_cluster() {
    static int cached = 0;
    if (cached == 0) {
        const char* e = getenv("LIFTER_GEMM_CLUSTER");
        const int v = e != nullptr ? atoi(e) : 0;
        cached = (v == 1 || v == 2 || v == 4 || v == 8) ? v : 2;
    }
    return cached;
}

int gemm_slab_rows(int M) { return ceil_div(M, BLOCK_M) * BLOCK_M; }
int gemm_effective_splits(int K, int splits) {
    const int kb = ceil_div(K, BLOCK_K);
    int s = splits < kb ? splits : kb;
    if (s < 1) s = 1;
    const int per = ceil_div(kb, s);
    return ceil_div(kb, per);
}
int gemm_stat_rows(int M) { return ceil_div(M, BLOCK_M) * 4; }

int gemm_launch(const GemmArgs& a, cudaStream_t stream) {
    LIFTER_REQUIRE(a.M > 0 && a.N > 0 && a.K > 0, "gemm: empty problem (M=%d N=%d K=%d)", a.M, a.N, a.K);
    LIFTER_REQUIRE(a.fmt == FMT_F16 || a.fmt == FMT_BF16, "gemm: unknown operand format %d", a.fmt);
    LIFTER_REQUIRE(a.A && a.B, "gemm: null operand");
    LIFTER_REQUIRE(a.splits >= 1, "gemm: splits must be >= 1");
    LIFTER_REQUIRE(a.splits == 1 || a.out_mode == OUT_32_TMA, "gemm: split-K needs the fp32 slab output");
    if (a.out_mode == OUT_32_DIRECT) {
        LIFTER_REQUIRE(a.n_valid > 0 && a.n_valid <= 64 && a.N <= 64, "gemm: direct fp32 output holds at most 64 columns");
        LIFTER_REQUIRE(a.D != nullptr, "gemm: null output");
    } else {
        LIFTER_REQUIRE(a.D != nullptr, "gemm: null output");
    }
    LIFTER_REQUIRE(a.stat_sum == nullptr || (a.out_mode == OUT_16_TMA && a.stat_sq != nullptr),
                   "gemm: column statistics ride on the 16-bit output");
    LIFTER_REQUIRE(a.residual == nullptr || (a.N % 32 == 0 && a.ldr % 8 == 0), "gemm: residual needs N %% 32 == 0");

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
    p.scale = a.scale; p.shift = a.shift;
    p.residual = static_cast<const uint16_t*>(a.residual); p.ldr = a.ldr;
    p.out_direct = static_cast<float*>(a.D); p.n_valid = a.n_valid;
    p.slab_rows = gemm_slab_rows(a.M);
    p.stat_sum = a.stat_sum; p.stat_sq = a.stat_sq;
    p.timeline = a.timeline;
    p.dbg = a.dbg;

    // Cluster size (B-tile multicast across CL vertically adjacent M-blocks).  a.cluster: 0 = auto, 1 = none,
    // 2 / 4 / 8 = forced (clamped to what the tile shape allows).  Auto: measured on B200, the lock-step of a
    // cluster only pays once every CTA works through several tiles.
    const int bn = a.N <= 64 ? 64 : (a.N <= 128 ? 128 : 256);
    const long tiles = static_cast<long>(p.num_m_blocks) * ceil_div(a.N, bn) * p.splits;
    int cl = a.cluster;
    if (cl == 0) cl = tiles >= 4L * num_sms() ? gemm_auto_cluster() : 1;
    int cl_max = a.b_mn_major ? bn / 64 : bn / 32;                 // MN-major B moves in 64-column atoms
    if (cl_max > 8) cl_max = 8;
    while (cl > 1 && (cl > cl_max || cl > p.num_m_blocks)) cl >>= 1;
    if (bn == 64) return launch_impl<64, 1>(p, a, stream);
    if (bn == 128) {
        if (cl >= 4) return launch_impl<128, 4>(p, a, stream);
        if (cl == 2) return launch_impl<128, 2>(p, a, stream);
        return launch_impl<128, 1>(p, a, stream);
    }
    if (cl == 8) return launch_impl<256, 8>(p, a, stream);
    if (cl == 4) return launch_impl<256, 4>(p, a, stream);
    if (cl == 2) return launch_impl<256, 2>(p, a, stream);
    return launch_impl<256, 1>(p, a, stream);
}

}  // namespace lifter
