// Keypoint-side kernels around the lifter (north_star (b)): flip_pose, mean/std (de)normalisation, and ONE fused pass
// computing MPJPE + camera reprojection + bone-length losses, their gradient and the per-joint MPJPE metric.
#pragma once
#include "common.cuh"

namespace lifter {

// phase3_direct/my_HybrIK/utils.py:372-396: mirror x (2D: 1 - x, 3D: -x) and swap the left/right joints of the
// H36M-17 skeleton.  grad = 1: the vector-Jacobian product of the same map (no "+1").  [poses][17][D], D in {2, 3}.
int flip_pose(const float* in, float* out, int64_t poses, int J, int D, int grad, cudaStream_t st);

enum : int { COLS_NORMALISE = 0, COLS_DENORMALISE = 1, COLS_MINMAX_NORM = 2, COLS_MINMAX_DENORM = 3 };
// per-column statistics broadcast over the rows: main.py:334,337 (x - a) / b ; main.py:473-479 y * b + a ;
// main.py:331 2x - 1 ; train_5.py:35-40 y * (a - b) + b with a = max, b = min.   [rows][cols], a/b [cols]
int affine_cols(const float* in, float* out, int64_t rows, int cols, const float* a, const float* b, int mode,
                cudaStream_t st);

// Keypoint ingest (video2keypoints.py:14-57 coco2h36m + the /1000 pixel scaling and the normalisation of main.py:331-337
// in the same pass): in [poses][17][in_channels] COCO order (x, y[, confidence]) -> out [poses][17][2] H36M order,
// out = norm(scale * joint);  mode -1: none, COLS_NORMALISE: (v - mean) / std with [34] statistics, COLS_MINMAX_NORM: 2v - 1
int coco2h36m(const float* in, int in_channels, float* out, int64_t poses, float scale, const float* mean,
              const float* stdv, int mode, cudaStream_t st);

struct PoseLossArgs {
    const float* pred; const float* target;     // [rows][J][3] (standardised when mean/std are given)
    const float* mean; const float* stdv;       // [J*3] nullable: joints = v * std + mean on both tensors
    const float* root;                          // [rows][3] nullable: added to pred and target before projecting
    const float* intrinsics;                    // [rows][4] (fx, fy, cx, cy) px, nullable: no reprojection term
    const float* target2d;                      // [rows][J][2] pixel/1000 (with intrinsics)
    int64_t rows, global_rows; int J;
    float w_mpjpe, w_reproj, w_bone;            // total = sum of w * term; the bone term needs J == 17
    float* dpred;                               // [rows][J][3] d total / d pred, nullable
    float* partials;                            // scratch: pose_loss_partial_rows(rows) * (4 + J) floats
    float* out;                                 // [4 + J]: total, mpjpe mean, reprojection L1, bone L1, per-joint sums
};
int pose_loss_partial_rows(int64_t rows);
int pose_loss(const PoseLossArgs& a, cudaStream_t st);

}  // namespace lifter
