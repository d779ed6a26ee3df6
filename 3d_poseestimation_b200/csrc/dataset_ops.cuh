// Dataset-side keypoint processing: phase1_lifting/main.py:271-355 (process_data) on the device.
#pragma once
#include <math_constants.h>

#include "common.cuh"
#include "pose_ops.cuh"

namespace lifter {

enum : int { ROOT_KEEP = 0, ROOT_ZERO = 1, ROOT_DROP = 2 };

// Number of partial rows the statistics pass writes: scratch = blocks * 4 * J * D doubles.
int keypoint_stats_blocks(int64_t frames, int cols);
// data [frames][J][D] fp32 -> out [4][J*D] fp64: mean, population std, min, max of the (optionally root-centred,
// main.py:275-277) values; main.py:280-289.  Deterministic: fixed-order partial sums, no atomics.
int keypoint_stats(const float* data, int64_t frames, int J, int D, int zero_centre, double* scratch, double* out,
                   cudaStream_t st);
// in [frames][J][D] -> out [frames][J or J-1][D]: root-centring, then mode COLS_NORMALISE ((v - mean) / std with the
// first two rows of `stats`), COLS_MINMAX_NORM (2v - 1) or -1 (none); root row kept / zeroed / dropped (main.py:340-344)
int process_keypoints(const float* in, float* out, int64_t frames, int J, int D, int zero_centre, const double* stats,
                      int mode, int root_mode, cudaStream_t st);

}  // namespace lifter
