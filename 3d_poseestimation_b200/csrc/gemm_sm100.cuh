// Host-side descriptor of one tcgen05 GEMM launch (see gemm_sm100.cu).
#pragma once
#include "common.cuh"
#include "philox.cuh"
#include "train_kernels.cuh"

namespace lifter {

enum : int {
    OUT_16_TMA = 0,     // 16-bit (fp16/bf16) output [M][N] through smem staging + TMA store
    OUT_32_TMA = 1,     // fp32 output (split-K partial slabs) through smem staging + TMA store
    OUT_32_DIRECT = 2,  // fp32 output with row stride n_valid (<= 64 columns, rows contiguous)
};

// A whole hidden layer of the training forward in ONE launch (GemmArgs::fused_bn): z = A W^T as usual (16-bit z stored,
// column partial sums written by the epilogue), but the fp32 accumulators STAY in TMEM while the grid meets at a barrier,
// shares out the column finalize (fp64 sums over the partial rows, SyncBN exchange, scale / shift, running statistics),
// meets again, and then every CTA applies BatchNorm + ReLU + Philox dropout (+ block residual) to its own tiles straight
// from TMEM and stores the activation and the "active" bits.  Replaces GEMM + bn_reduce_finalize_fwd + bn_apply_fwd
// (three launches, z written and read back).  Needs: CTA pairs with at most TWO 256 x 256 tiles per pair (the 512 TMEM
// columns hold both; 8,192 x 1,024 on 74 pairs = 128 tiles fits), N % 256 == 0, one-bit dropout lanes (p = 0.5) or no
// dropout, row_offset % 16 == 0.  gemm_fused_bn_fits() tells; the caller falls back to the three launches otherwise.
struct FusedBnArgs {
    BnFwdFinalize f;                  // `sums` unused (the kernel reduces the epilogue's partial rows itself); ex = SyncBN
    double* sums_out = nullptr;       // [2][N] global column sums (kept for tests / the NCCL path)
    void* a_out = nullptr;            // 16-bit [M][N] activation
    const void* resid = nullptr; long ldr = 0;   // 16-bit block input added after dropout (nullable)
    uint8_t* bits = nullptr;          // [M][N/8]: ReLU-active AND kept
    int apply_dropout = 0; float p = 0.0f; unsigned long long seed = 0;
    unsigned step = 0, site = 0, step_mul = 1, step_add = 0; const unsigned* step_dev = nullptr; long long row_offset = 0;
    unsigned* barrier = nullptr;      // two words, zero when the kernel starts
};
bool gemm_fused_bn_fits(int M, int N, float p, int apply_dropout, long long row_offset);

// D[M][N] = sum_k A(m,k) * B(n,k), 16-bit operands, fp32 accumulation in TMEM.
//   A "K-major":  global A  is [M rows][K cols]   (row-major, K contiguous)
//   A "MN-major": global A^T is [K rows][M cols]  (row-major, M contiguous)      -- likewise for B with N.
struct GemmArgs {
    const void* A = nullptr;  long lda = 0;   // leading dimension in elements (row stride of the global matrix)
    const void* B = nullptr;  long ldb = 0;
    int M = 0, N = 0, K = 0;
    int a_mn_major = 0, b_mn_major = 0;
    int fmt = FMT_BF16;                       // operand (and 16-bit output) format
    int out_mode = OUT_16_TMA;
    void* D = nullptr;        long ldd = 0;   // OUT_16_TMA / OUT_32_TMA: [rows][ldd]; OUT_32_DIRECT: [M][n_valid]
    int n_valid = 0;                          // OUT_32_DIRECT only
    int splits = 1;                           // split-K (OUT_32_TMA): slab s starts at row s * slab_rows(M)
    int relu = 0;
    int cluster = 0;                          // 0 auto, 1 single-CTA tiles (cta_group::1), 2 CTA pairs (cta_group::2)
    const float* scale = nullptr;             // per-column multiply (nullptr = 1)
    const float* shift = nullptr;             // per-column add      (nullptr = 0)
    int shift_valid = 0;                      // entries of scale/shift that exist (0 = N); columns past it use 1 / 0
    const void* residual = nullptr; long ldr = 0;   // 16-bit [M][ldr], added after the activation
    int residual_op = 0;                            // 1: MULTIPLY the result by GELU'(residual) instead (backward of the
                                                    // transformer lifter's GELU: d z1 = (d m) * GELU'(z1) in the dgrad epilogue)
    const float* residual32 = nullptr; long ldr32 = 0;   // OUT_32_TMA, one split: fp32 [M][ldr32] added after the activation (fp32 residual stream)
    float* stat_sum = nullptr;                // [4*ceil(M/128)][N] per-32-row-slab column sums of the stored values
    float* stat_sq = nullptr;                 // same, squares
    unsigned long long* timeline = nullptr;   // debug hook: [grid][32] globaltimer stamps
    const char* tag = nullptr;                // name in the debug trace
    int max_stages = 0;                       // debug hook: cap the operand ring depth (0 = all that fit)
    int max_ctas = 0;                         // cap the persistent grid (0 = every SM), for kernels meant to run side by side
    const FusedBnArgs* fused_bn = nullptr;    // whole-layer mode (see FusedBnArgs); needs OUT_16_TMA + stat_sum / stat_sq
};

int gemm_slab_rows(int M);                    // rows of one split-K slab (M rounded up to 128)
int gemm_effective_splits(int K, int splits);   // splits actually used (every split owns >= 1 K block)
int gemm_stat_rows(int M);                    // 4 * ceil(M / 128)
int gemm_launch(const GemmArgs& a, cudaStream_t stream);

}  // namespace lifter
