// Elementwise / column-reduction kernels around the GEMMs (HBM- or L2-bound; vectorised 16-byte accesses).
#pragma once
#include "common.cuh"

namespace lifter {

// fp32 [rows][cols] (dense) -> 16-bit [rows_p][cols_p], zero padded.  Weight operand copies.
int convert_pad(const float* src, int rows, int cols, void* dst, int rows_p, int cols_p, int fmt, cudaStream_t st);

// Eval-mode fold of Linear bias + BatchNorm1d into y = z * scale + shift:
//   use_bn: scale = gamma / sqrt(rv + eps), shift = (bias - rm) * scale + beta;   else scale = 1, shift = bias
int fold_bn(const float* bias, const float* gamma, const float* beta, const float* rm, const float* rv, int H,
            int use_bn, float eps, float* scale, float* shift, cudaStream_t st);

// x fp32 [rows][I] (row stride I) -> 16-bit [rows][Ip] zero padded; optional (x - mean) / std per column
// (main.py:334, the 2D keypoint standardisation).
// zero_words (nullable): n_zero <= 256 words that the kernel also zeroes -- the grid-barrier words of the pass it opens.
int prep_input(const float* x, int64_t rows, int I, int Ip, const float* mean, const float* stdv, void* dst, int fmt,
               cudaStream_t st, unsigned* zero_words = nullptr, int n_zero = 0);

// scale[j] = std[j] (or 1), shift[j] = bias[j] * std[j] + mean[j]  for j < O, zero beyond: the final layer's
// bias with the de-standardisation y * std + mean (main.py:473-479) folded in.
int make_out_affine(const float* bias, const float* mean, const float* stdv, int O, int Op, float* scale, float* shift,
                    cudaStream_t st);

}  // namespace lifter
