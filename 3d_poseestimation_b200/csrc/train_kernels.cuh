// Elementwise + column-reduction kernels of the training path (BatchNorm statistics, BN/ReLU/dropout apply,
// their backward, gradient reductions).  All are L2/HBM-bound: 16-byte vector accesses, deterministic
// two-stage column reductions (per-block partial rows, then a fixed-order reduce) -- no float atomics.
#pragma once
#include "common.cuh"

namespace lifter {

constexpr int EW_THREADS = 256;
constexpr int EW_UNROLL = 4;                // row-iterations whose loads are issued together

// rows handled by one block of the column-reducing elementwise kernels, and the partial-row count for `rows`
int ew_rows_per_block(int64_t rows, int H);
int ew_partial_rows(int64_t rows, int H);

// out_a[j] = sum_p a[p * ld + j] for j < n (fp64 accumulate), same for b (nullable)
int reduce_rows_f64(const float* a, const float* b, int P, int n, int ld, double* out_a, double* out_b, cudaStream_t st);
// out[j] = sum_p a[p * ld + j] written as fp32 (bias gradients)
int reduce_rows_f32(const float* a, int P, int n, int ld, float* out, cudaStream_t st);

// Synchronised BatchNorm over the GPUs of one box WITHOUT a collective call: every rank stores its local column sums
// into its slot of every peer's symmetric buffer (NVLink P2P stores, each value tagged with the step's epoch), polls
// its own buffer for the peers' values and adds the slots in rank order -- identical, deterministic sums on every
// rank after one NVLink store latency, inside the BatchNorm finalize kernel (include/lifter_b200.h: lifter_dp_t).
constexpr int DP_MAX_WORLD = 8;
struct SyncExchange {
    int world = 1, rank = 0;           // world == 1: no exchange
    int site = 0, nsites = 1;          // which of the step's exchange points this is (forward l, backward l)
    void* data[DP_MAX_WORLD];          // peer p's buffer: [2 parities][nsites][world slots][2][H] 16-byte LL words
    const unsigned* epoch;             // device: the step counter (lifter_step_state_t.adam_step), >= 1
    long long spin_limit = 0;          // SM clocks a peer is waited for before the kernel traps (0 = forever)
    int serial = 1;                    // 1 (default): one thread per column polls slot after slot; 0: warp-cooperative polling
};

struct BnFwdFinalize {
    SyncExchange ex;
    const double* sums;        // [2][H] (sum z, sum z^2) over the GLOBAL batch; unused when bn_mode != 0
    double count;              // global batch rows
    const float* bias; const float* gamma; const float* beta;
    float* running_mean; float* running_var; long long* num_batches_tracked;   // nullable: no update
    float momentum, eps;
    int H, use_bn, bn_mode;    // bn_mode 0: batch statistics (train), 1: running statistics (eval with autograd)
    float* mean_z; float* rstd; float* scale; float* shift;                     // [H] each
};
int bn_finalize_fwd(const BnFwdFinalize& a, cudaStream_t st);
// single-GPU shortcut: second reduction stage over the GEMM's partial rows + finalize in one launch (batch statistics)
int bn_reduce_finalize_fwd(const float* psum, const float* psq, int P, const BnFwdFinalize& a, double* sums_out,
                           cudaStream_t st);

struct BnApplyFwd {
    const void* z; const float* scale; const float* shift; const void* resid;   // resid nullable
    int64_t rows; int H; int fmt;
    int apply_dropout; float p; unsigned long long seed; unsigned step, site; long long row_offset;
    const unsigned* step_dev;   // nullable: the dropout step is read from device memory (graph replay)
    unsigned step_mul = 1, step_add = 0;   // Philox step = (device or host step) * step_mul + step_add: a training step
                                           // that runs several lifter forwards (phase-5 loop) gives each its own masks
    void* a; uint8_t* bits;     // a [rows][H] 16-bit, bits [rows][H/8]: relu-active AND kept
};
int bn_apply_fwd(const BnApplyFwd& a, cudaStream_t st);
// bn_reduce_finalize_fwd + bn_apply_fwd as ONE launch with a grid barrier between the two (batch statistics only).
// `barrier`: one word, zero when the kernel starts (common.cuh: grid_barrier_once).
int bn_fwd_fused(const float* psum, const float* psq, int P, const BnFwdFinalize& f, double* sums_out,
                 const BnApplyFwd& a, unsigned* barrier, cudaStream_t st);

struct BwdCommon {
    const void* g; const void* z; const uint8_t* bits;
    int64_t rows; int H; int fmt;
};
// pass 1: partial column sums of gm = (g where the saved bit is set, else 0) and of gm * z: [ew_partial_rows][H] each
int bwd_stats(const BwdCommon& c, float* psum_g, float* psum_gz, cudaStream_t st);

struct BnBwdFinalize {
    SyncExchange ex;
    const double* sums;        // [2][H] (sum dy, sum dy*zhat) over the GLOBAL batch (bn_finalize_bwd only)
    double count;
    const float* gamma; const float* rstd; const float* mean_z;
    float inv_keep;            // dy = inv_keep * gm
    int H, use_bn, bn_mode;
    float* coef;               // [3][H]: dz = coef0 * gm + coef1 * z + coef2
    float* dbias;              // nullable: zeroed (the Linear bias in front of batch-statistics BN has no gradient)
};
int bn_finalize_bwd(const BnBwdFinalize& a, cudaStream_t st);
// second stage of pass 1: local sums (sum dy, sum dy*zhat) in fp64 -> sums_out [2][H], dgamma/dbeta from them;
// fused = also the pass-2 coefficients (single GPU, where the local sums are the global ones)
int bn_bwd_reduce(const float* pg, const float* pgz, int P, const BnBwdFinalize& a, double* sums_out, float* dgamma,
                  float* dbeta, bool fused, cudaStream_t st);

// pass 2: dz = coef0 * gm + coef1 * z + coef2 as 16-bit; psum_dz (nullable) receives partial column sums of dz
int bwd_dz(const BwdCommon& c, const float* coef, void* dz, float* psum_dz, cudaStream_t st);

// bwd_stats + bn_bwd_reduce<fused> + bwd_dz (no bias sums) as ONE launch with two grid barriers.  psum_g / psum_gz hold
// bn_bwd_fused_partial_rows(rows, H) rows of H floats; `barrier`: two words, zero when the kernel starts.
int bn_bwd_fused_partial_rows(int64_t rows, int H);
int bn_bwd_fused(const BwdCommon& c, float* psum_g, float* psum_gz, const BnBwdFinalize& f, double* sums_out,
                 float* dgamma, float* dbeta, void* dz, unsigned* barrier, cudaStream_t st);

// sum of split-K slabs -> dense fp32 gradient [rows][cols]
int wgrad_reduce(const float* slabs, int splits, int slab_rows, int ld, int rows, int cols, float* grad,
                 cudaStream_t st);

// grad_out fp32 [rows][O] -> 16-bit [rows][Op] zero padded, plus partial column sums (output-layer bias gradient)
// zero_words (nullable): n_zero <= 256 words the kernel also zeroes -- the grid-barrier words of the backward pass it opens
int prep_grad_out(const float* dout, int64_t rows, int O, int Op, int fmt, void* d16, float* psum, cudaStream_t st,
                  unsigned* zero_words = nullptr, int n_zero = 0);

}  // namespace lifter
