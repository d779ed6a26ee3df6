// The loss of the phase-5 loop around the lifter (phase5_loop/backup_june8.py:137-161, train_5.py:145-181) in ONE
// kernel + one finalize: flip averaging of the four lifter outputs, the L1 terms of TriangleLoss / TriangleLoss_sep
// (losses.py:24-53, 81-110, Project = False), optional pinhole-reprojection and bone-length terms on the lifted pose
// (PARITY UNPINNED) and every gradient the loop back-propagates.
#pragma once
#include "common.cuh"

namespace lifter {

struct Phase5LossArgs {
    // lifter outputs [rows][J*3]: lift(pred 2D), lift(flip(pred 2D)), lift(gt 2D), lift(flip(gt 2D)).
    // *_flip nullable: no flip averaging (Flip = False).  out_gt nullable when mode == 0 (TriangleLoss never reads it).
    const float* out_pred; const float* out_pred_flip; const float* out_gt; const float* out_gt_flip;
    const float* pred2d; const float* gt2d;       // [rows][J*2]
    const float* pred3d; const float* gt3d;       // [rows][J*3]
    int mode;                                     // 0: TriangleLoss, 1: TriangleLoss_sep
    // unpinned terms on lift_pred, both nullable / zero-weighted
    const float* mean3d; const float* std3d;      // [J*3]: joints = v * std + mean (lift_pred and gt3d)
    const float* mean2d; const float* std2d;      // [J*2] nullable: gt2d * std + mean = key points in pixel / 1000
    const float* root;                            // [rows][3] nullable
    const float* intrinsics;                      // [rows][4] nullable: no reprojection term
    float w_reproj, w_bone;
    int64_t rows, global_rows; int J;
    float* d_out_pred; float* d_out_pred_flip;    // gradients w.r.t. the lifter outputs (flip ones nullable with their inputs)
    float* d_out_gt; float* d_out_gt_flip;        // mode 1 only (nullable in mode 0)
    float* d_pred2d; float* d_pred3d;             // the loss's own share of the gradient of the 2D / 3D predictions
    float* partials;                              // phase5_partial_rows(rows) * 8 floats
    float* out;                                   // [8]: total, loss_2d, loss_3d, loss_lift, domain gap, reprojection, bone, 0
};
int phase5_partial_rows(int64_t rows);
int phase5_loss(const Phase5LossArgs& a, cudaStream_t st);

// d(loss) / d(predicted 2D key points) = the loss's own share + the lifter's input gradient of the plain forward +
// the flipped input gradient of the flipped forward (flip_pose's vector-Jacobian product); dx_flip nullable.  [rows][J*2]
int phase5_input_grad(const float* d_own, const float* dx, const float* dx_flip, int64_t rows, int J, float* out,
                      cudaStream_t st);

// dst[i] = a[i] + b[i] (+ c[i] + d[i]); c, d nullable.  Sums the parameter gradients of the step's backward passes.
int sum_buffers(float* dst, const float* a, const float* b, const float* c, const float* d, int64_t n, cudaStream_t st);

}  // namespace lifter
