// Host-side descriptor of one tcgen05 GEMM launch (see gemm_sm100.cu).
#pragma once
#include "common.cuh"

namespace lifter {

enum : int {
    OUT_16_TMA = 0,     // 16-bit (fp16/bf16) output [M][N] through smem staging + TMA store
    OUT_32_TMA = 1,     // fp32 output (split-K partial slabs) through smem staging + TMA store
    OUT_32_DIRECT = 2,  // fp32 output with row stride n_valid (<= 64 columns, rows contiguous)
};

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
    float* stat_sum = nullptr;                // [4*ceil(M/128)][N] per-32-row-slab column sums of the stored values
    float* stat_sq = nullptr;                 // same, squares
    unsigned long long* timeline = nullptr;   // debug hook: [grid][32] globaltimer stamps
    const char* tag = nullptr;                // name in the debug trace
    int max_stages = 0;                       // debug hook: cap the operand ring depth (0 = all that fit)
    int max_ctas = 0;                         // cap the persistent grid (0 = every SM), for kernels meant to run side by side
};

int gemm_slab_rows(int M);                    // rows of one split-K slab (M rounded up to 128)
int gemm_effective_splits(int K, int splits);   // splits actually used (every split owns >= 1 K block)
int gemm_stat_rows(int M);                    // 4 * ceil(M / 128)
int gemm_launch(const GemmArgs& a, cudaStream_t stream);

}  // namespace lifter
