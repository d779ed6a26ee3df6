// Fused loss / metric kernel (b) and fused flat-buffer Adam (c) of BASELINE.json's north_star.
#pragma once
#include "common.cuh"

namespace lifter {

enum : int { LOSS_MSE = 0, LOSS_L1 = 1, LOSS_MPJPE = 2 };

// pred/target [rows][J*3] fp32.  One pass computes
//   loss   = mean over the GLOBAL batch (count rows) of  (d^2 | |d| | ||d_joint||)          -> out[0]
//   dpred  = d loss / d pred  (nullable)                                                      [rows][J*3]
//   metric = per-joint sum over the local rows of ||(pred - target) * std|| (loss_MPJPE)      -> out[1 .. J]
// `stdv` (nullable, [J*3]) de-standardises the error for the metric (y*std+mean on both sides cancels the mean).
// partials: [loss_partial_rows(rows)][1 + J] floats of scratch.  accumulate_metric: out[1..J] += (epoch accumulation).
int loss_partial_rows(int64_t rows);
int fused_loss(const float* pred, const float* target, int64_t rows, int J, int kind, double count,
               const float* stdv, float* dpred, float* partials, float* out, int accumulate_metric, cudaStream_t st);

// sum |a - b| over n elements -> out[0] = sum * scale (L1Loss building block of TriangleLoss, losses.py:21-53);
// centre: subtract row 0 (row_len elements) from rows 1.. of both tensors first (the reference's batch-dim quirk).
int l1_sum(const float* a, const float* b, int64_t n, int row_len, int centre, float scale, float* partials,
           float* out, cudaStream_t st);
// da = sign(a' - b') * gscale[0] * scale (and db = -da when db != nullptr); with centre, row 0 collects the
// negated sum of the other rows' gradients.
int l1_grad(const float* a, const float* b, int64_t n, int row_len, int centre, const float* gscale, float scale,
            float* da, float* db, cudaStream_t st);

constexpr int ADAM_MAX_WEIGHTS = 16;
struct AdamArgs {
    float* p; const float* g; float* m; float* v; int64_t n;
    long long base = 0;                                           // first element of the range (multiple of 4); n = its length
    float lr, beta1, beta2, eps, weight_decay; int decoupled;     // AdamW: decoupled = 1
    int step;                                                     // 1-based
    float grad_scale;                                             // multiplies g first (1/world for DP means)
    const float* dyn = nullptr;     // device lifter_step_state_t (nullable): step_size / inv_sqrt_bc2 / decay read from it
    // optional: refresh the 16-bit zero-padded GEMM operand copies of the Linear weights in the same pass
    void* operands = nullptr; int op_fmt = FMT_BF16; int nw = 0;
    long long w_p_off[ADAM_MAX_WEIGHTS]; long long w_numel[ADAM_MAX_WEIGHTS]; long long w_off[ADAM_MAX_WEIGHTS];
    int w_in[ADAM_MAX_WEIGHTS]; int w_pad[ADAM_MAX_WEIGHTS];
};
int adam_step(const AdamArgs& a, cudaStream_t st);
// Data-parallel optimiser step as ONE kernel over NVLink peer memory (no NCCL call): rank r owns slice r of the flat
// buffer.  After a flag barrier ("every rank's gradient is complete") it reads that slice of every rank's gradient
// with P2P loads, adds them in rank order, runs Adam on the slice and stores the new parameters (fp32 + 16-bit
// operand copies) into EVERY rank's buffers with P2P stores; a second flag barrier ends the kernel once all peers
// have finished writing here.  Reduce-scatter + Adam + all-gather fused: each gradient crosses NVLink once, each
// parameter once, and the Adam state traffic is 1/world of the replicated optimiser's.
constexpr int ADAM_DP_MAX_WORLD = 8;
struct AdamDpArgs {
    AdamArgs a;                                  // p/g = THIS rank's buffers; operands = this rank's operand copy
    int world, rank;
    const float* g_peer[ADAM_DP_MAX_WORLD];      // every rank's gradient buffer (peer-mapped)
    float* p_peer[ADAM_DP_MAX_WORLD];            // every rank's parameter buffer
    void* op_peer[ADAM_DP_MAX_WORLD];            // every rank's operand buffer (nullable when a.operands == nullptr)
    unsigned* flag_peer[ADAM_DP_MAX_WORLD];      // every rank's flag block: [2][world] epochs + [1] block counter
    const unsigned* epoch;                       // device step counter
    // a step may run several launches (one per parameter range): launch `phase` of the step uses the barrier epoch
    // step * ADAM_DP_PHASES + phase; only the `last` launch runs the closing "all stores have landed" barrier
    int phase = 0, last = 1;
    long long spin_limit = 0;                    // SM clocks a peer is waited for (0 = forever); set by adam_step_dp
    const float* mc_g = nullptr; float* mc_p = nullptr; void* mc_op = nullptr;   // NVSwitch multicast views (nullable)
    int external_flags = 0;                      // the flag handshakes run as their own one-block launches (set by adam_step_dp)
};
constexpr int ADAM_DP_PHASES = 64;
int adam_step_dp(const AdamDpArgs& a, cudaStream_t st);

// device step state (include/lifter_b200.h: lifter_step_state_t): dropout_step = adam_step; adam_step += 1; the Adam
// bias corrections of the new step from the lr stored in the state
int step_advance(void* state, float beta1, float beta2, float weight_decay, cudaStream_t st);

}  // namespace lifter
