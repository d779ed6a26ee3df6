/* lifter_b200.h -- C ABI of the B200-native phase1_lifting LinearModel hot path.
 *
 * The reference (RHnejad/3D_PoseEstimation) has no FFI/plugin layer: its boundary is the Python
 * nn.Module contract of phase1_lifting/baselineModel.py:50-102.  This header is
 * what the Python mirror (3d_poseestimation_b200/baselineModel.py, losses.py, trainer.py) binds with
 * ctypes; every entry point names the reference code it replaces.
 *
 * Conventions
 *   - every pointer is a DEVICE pointer owned by the caller (PyTorch's caching allocator) unless it is
 *     a struct passed by address; the library allocates nothing and never synchronises the host;
 *   - every call is asynchronous on `stream` (a cudaStream_t passed as void*);
 *   - return value: 0 = OK, negative = error (LIFTER_ERR_*); text via lifter_last_error();
 *   - "fmt" is the 16-bit operand format of the tensor-core GEMMs: 0 = fp16, 1 = bf16 (fp32 accumulate);
 *   - flat parameter layout = reference model.parameters() order (baselineModel.py:68-85, 24-28), each
 *     tensor starting at a multiple of 32 floats: use lifter_param_offset().
 */
#ifndef LIFTER_B200_H
#define LIFTER_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define LIFTER_OK 0
#define LIFTER_ERR_INVALID (-1)
#define LIFTER_ERR_CUDA (-2)
#define LIFTER_ERR_BATCH1 (-3)   /* train-mode BatchNorm with one row: the reference raises ValueError */
#define LIFTER_ERR_NODEVICE (-4)

#define LIFTER_FMT_F16 0
#define LIFTER_FMT_BF16 1

/* constructor arguments of LinearModel (baselineModel.py:51-56) */
typedef struct lifter_config {
    int32_t i_dim;        /* input features  (17*2 = 34)            */
    int32_t o_dim;        /* output features (17*3 = 51), <= 64     */
    int32_t linear_size;  /* hidden width H: 64..2048, power of two */
    int32_t num_stage;    /* residual blocks                         */
    int32_t use_bn;       /* BN flag (baselineModel.py:85)           */
    float p_dropout;      /* baselineModel.py:81                     */
    float bn_eps;         /* 1e-5  (torch.nn.BatchNorm1d default)    */
    float bn_momentum;    /* 0.1                                      */
} lifter_config_t;

/* ---- library ------------------------------------------------------------------------------- */
int lifter_version(void);
const char* lifter_last_error(void);
int lifter_device_check(void);                       /* 0 if the current device is sm_100 */
const char* lifter_source_digest(void);              /* sha256 of the sources this binary was built from (build.py) */
/* Kernel launches issued by the library so far, host-side count (a launch captured into a CUDA graph counts once, at
 * capture): what bench.py reports as gpu_launches per step.                                                        */
long long lifter_launch_count(void);

/* ---- layout queries (host only) -------------------------------------------------------------- */
/* Flat fp32 parameter buffer.  tensor index t: per hidden layer l (0 .. 2*num_stage):
 *   4l+0 Linear.weight [H][in], 4l+1 Linear.bias [H], 4l+2 BatchNorm.weight [H], 4l+3 BatchNorm.bias [H];
 *   then w2.weight [o_dim][H], w2.bias [o_dim].                                                      */
int lifter_num_param_tensors(const lifter_config_t* cfg);
int64_t lifter_param_offset(const lifter_config_t* cfg, int tensor);      /* in floats            */
int64_t lifter_param_numel(const lifter_config_t* cfg, int tensor);
int64_t lifter_param_total(const lifter_config_t* cfg);                   /* padded flat length   */
/* BatchNorm buffers: running_mean | running_var, each [num_layers][H] fp32; num_batches_tracked int64[num_layers] */
int64_t lifter_operand_bytes(const lifter_config_t* cfg);                 /* 16-bit padded GEMM operand copy of the weights */
int64_t lifter_eval_workspace_bytes(const lifter_config_t* cfg, int64_t batch);

/* ---- weights -------------------------------------------------------------------------------- */
/* fp32 flat params -> zero-padded 16-bit operand copies (private to the kernels; refreshed after every update) */
int lifter_pack_weights(const lifter_config_t* cfg, const float* params, void* operands, int fmt, void* stream);
/* eval-mode BatchNorm (+ Linear bias) folded to per-column scale/shift [num_layers][2][H]
 * (torch BatchNorm1d eval: (z - running_mean) / sqrt(running_var + eps) * weight + bias)          */
int lifter_fold_bn(const lifter_config_t* cfg, const float* params, const float* running_mean,
                   const float* running_var, float* scale_shift, void* stream);

/* ---- inference: LinearModel.forward under model.eval() + torch.no_grad() (baselineModel.py:87-102) --- */
/* x [batch][i_dim] fp32 -> out [batch][o_dim] fp32.  in_mean/in_std (nullable, [i_dim]) fuse the input
 * normalisation (x - mean) / std of main.py:334; out_mean/out_std (nullable, [o_dim]) fuse the
 * de-standardisation y * std + mean of main.py:473-479.  chunk_rows = 0 picks an L2-resident chunk.   */
int lifter_forward_eval(const lifter_config_t* cfg, const void* operands, const float* scale_shift,
                        const float* w2_bias, const float* x, float* out, int64_t batch,
                        const float* in_mean, const float* in_std, const float* out_mean, const float* out_std,
                        void* workspace, int64_t workspace_bytes, int fmt, int64_t chunk_rows, void* stream);

/* ---- training: forward with saved activations, backward, loss, optimiser ----------------------------
 * Replaces the autograd graph of LinearModel.forward in train mode (baselineModel.py:87-102, 32-47),
 * loss.backward() and optimizer.step() (train_1.py:94-96).  One context describes one forward/backward
 * pair; `saved` is written by forward and read by backward.                                            */
/* Per-step scalars that live in DEVICE memory so that one captured CUDA graph can be replayed for every step:
 * lifter_step_advance() (first node of the step) moves the counters on and derives Adam's bias corrections from
 * `lr`, which the host may overwrite between steps (ReduceLROnPlateau, train_1.py:41,106).                   */
typedef struct lifter_step_state {
    uint32_t dropout_step;   /* Philox step of this training step's masks (= adam_step - 1)        */
    int32_t adam_step;       /* optimiser steps taken, 1-based after the first lifter_step_advance */
    float lr;                /* host-written                                                       */
    float step_size;         /* lr / (1 - beta1^t)                                                 */
    float inv_sqrt_bc2;      /* 1 / sqrt(1 - beta2^t)                                              */
    float decay;             /* 1 - lr * weight_decay (AdamW)                                      */
    float reserved[2];
} lifter_step_state_t;

/* Data parallelism inside one NVSwitch box without collective calls on the critical path: `sync` is a symmetric
 * buffer (same size on every rank, peer-mapped over NVLink, e.g. torch.distributed._symmetric_memory) of
 * lifter_dp_sync_bytes() bytes, zeroed once; sync[p] is rank p's copy as seen from this rank.  The BatchNorm
 * finalize kernels exchange their column sums through it (forward and backward of every layer).  Needs
 * step_state (its adam_step is the exchange epoch) and the single-call lifter_train_forward / _backward.      */
#define LIFTER_DP_MAX_WORLD 8
typedef struct lifter_dp {
    int32_t world, rank;
    void* sync[LIFTER_DP_MAX_WORLD];       /* BatchNorm exchange region, lifter_dp_sync_bytes()                      */
    /* lifter_dp_adam_step only (else NULL): every rank's flat gradient, flat parameters, 16-bit operand copies and a
     * zeroed flag block of LIFTER_DP_FLAG_BYTES, all inside symmetric (peer-mapped) memory                           */
    void* grad[LIFTER_DP_MAX_WORLD];
    void* params[LIFTER_DP_MAX_WORLD];
    void* operands[LIFTER_DP_MAX_WORLD];
    void* flags[LIFTER_DP_MAX_WORLD];
    /* Optional NVSwitch multicast mappings of the SAME three regions (one address that reaches every rank's copy, e.g.
     * torch symmetric memory's multicast_ptr + the region offset); all three set or all NULL.  With them the optimiser
     * kernel reduces its gradient slice in the switch (multimem.ld_reduce) and broadcasts the new parameters and dense
     * operand copies with one store each (multimem.st) instead of `world` unicast loads / stores per element.          */
    void* mc_grad; void* mc_params; void* mc_operands;
} lifter_dp_t;
#define LIFTER_DP_FLAG_BYTES 256
int64_t lifter_dp_sync_bytes(const struct lifter_config* cfg, int world);

typedef struct lifter_train_ctx {
    const float* params;            /* flat fp32 parameters                                              */
    const void* operands;           /* 16-bit operand copies in `fmt` (lifter_pack_weights)              */
    float* running_mean;            /* [layers][H]  nullable: no running-stat update                      */
    float* running_var;             /* [layers][H]                                                        */
    int64_t* num_batches_tracked;   /* [layers]     nullable                                              */
    const float* x;                 /* [batch][i_dim] fp32                                                */
    float* out;                     /* [batch][o_dim] fp32                                                */
    void* saved; int64_t saved_bytes;          /* lifter_train_saved_bytes()                              */
    int64_t batch;                  /* local rows                                                         */
    int64_t global_batch;           /* rows of the whole data-parallel step (BatchNorm statistics count)  */
    int64_t row_offset;             /* global index of local row 0 (dropout stream)                       */
    int32_t fmt;
    int32_t bn_mode;                /* 0: batch statistics (model.train()), 1: running statistics (eval) */
    int32_t apply_dropout;          /* 1 in train mode                                                    */
    uint32_t step;                  /* dropout stream: optimiser step                                     */
    uint64_t seed;                  /* dropout stream: seed                                               */
    const float* grad_out;          /* backward: [batch][o_dim] fp32                                      */
    float* grad_params;             /* backward: flat fp32, parameter layout; every tensor is overwritten */
    float* grad_x;                  /* backward: [batch][i_dim] fp32, nullable                            */
    void* scratch; int64_t scratch_bytes;      /* backward: lifter_train_scratch_bytes()                  */
    const struct lifter_step_state* step_state; /* nullable, DEVICE memory: the dropout step is read from it at run
                                                  time instead of `step` (a captured CUDA graph replays every step) */
    const lifter_dp_t* dp;          /* nullable: in-kernel synchronised BatchNorm over dp->world GPUs (host struct)   */
    /* A training step that runs SEVERAL lifter forwards (the phase-5 loop: predicted / ground-truth key points and their
     * flipped twins, backup_june8.py:137-154) gives each call its own dropout masks and its own slots of the SyncBN
     * exchange buffer: Philox step = step * calls_per_step + call_index; dp->sync must then hold calls_per_step times
     * lifter_dp_sync_bytes().  0 / 0 = one call per step.                                                           */
    int32_t calls_per_step;
    int32_t call_index;
} lifter_train_ctx_t;

int64_t lifter_train_saved_bytes(const lifter_config_t* cfg, int64_t batch);
int64_t lifter_train_scratch_bytes(const lifter_config_t* cfg, int64_t batch);
/* byte offset inside `saved` / `scratch` of layer l's double[2][H] column sums: the quantity a data-parallel
 * run all-reduces between ..._gemm/_stats and ..._finish (synchronised BatchNorm)                          */
int64_t lifter_train_fwd_sums_offset(const lifter_config_t* cfg, int64_t batch, int layer);
int64_t lifter_train_bwd_sums_offset(const lifter_config_t* cfg, int64_t batch, int layer);

int lifter_train_fwd_layer_gemm(const lifter_config_t* cfg, const lifter_train_ctx_t* ctx, int layer, void* stream);
int lifter_train_fwd_layer_finish(const lifter_config_t* cfg, const lifter_train_ctx_t* ctx, int layer, void* stream);
int lifter_train_fwd_output(const lifter_config_t* cfg, const lifter_train_ctx_t* ctx, void* stream);
int lifter_train_forward(const lifter_config_t* cfg, const lifter_train_ctx_t* ctx, void* stream);   /* all layers, one GPU */

int lifter_train_bwd_begin(const lifter_config_t* cfg, const lifter_train_ctx_t* ctx, void* stream);
int lifter_train_bwd_layer_stats(const lifter_config_t* cfg, const lifter_train_ctx_t* ctx, int layer, void* stream);
int lifter_train_bwd_layer_finish(const lifter_config_t* cfg, const lifter_train_ctx_t* ctx, int layer, void* stream);
int lifter_train_backward(const lifter_config_t* cfg, const lifter_train_ctx_t* ctx, void* stream);  /* all layers, one GPU */
/* Ownership note (an exception to "the library owns only plan handles"): lifter_train_backward and
 * lifter_train_backward_step run the weight-gradient GEMMs (and the optimiser ranges) on side streams so that they overlap
 * the BatchNorm-backward chain.  The library creates, on first use and per device, two non-blocking streams and a few
 * timing-disabled events for this and keeps them until the process exits; they fork from and join back into `stream`
 * with events, so the call is still ordered on `stream` and legal under stream capture.  LIFTER_WGRAD_STREAM=0 keeps
 * everything on `stream` (the per-layer entry points above never use the side streams).  lifter_vit_backward does the
 * same with one more stream and six events of its own (LIFTER_VIT_WGRAD_STREAM=0: one stream).                         */

/* Fused loss + gradient + MPJPE metric (losses.py:3-7, train_1.py:37,94,100; two2three.py:387-390).
 * kind 0: MSELoss(mean), 1: L1Loss(mean), 2: mean MPJPE.  out[0] = loss over `global_rows`, out[1..joints] =
 * per-joint sum over the local rows of ||(pred - target) * std|| (std nullable).  dpred nullable.        */
#define LIFTER_LOSS_MSE 0
#define LIFTER_LOSS_L1 1
#define LIFTER_LOSS_MPJPE 2
int64_t lifter_loss_scratch_bytes(int64_t rows, int joints);
int lifter_loss(const float* pred, const float* target, int64_t rows, int joints, int kind, int64_t global_rows,
                const float* stdv, float* dpred, float* scratch, float* out, int accumulate_metric, void* stream);
/* lifter_loss WITHOUT caller-provided scratch and without any allocation.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA-graph capture, tight
 * serving loops, harnesses that forbid it).  Same arguments, arithmetic and results as lifter_loss; the per-block partial
 * sums live in a file-scope `static __device__` array of the library (88 KB, held by every GPU that loads the library
 * until the process exits), so rows <= LIFTER_LOSS_STATIC_MAX_ROWS and joints <= 21 -- larger calls are rejected with
 * LIFTER_ERR_INVALID.  Calls no cudaMalloc / cudaFree / allocating routine; the first call on a device looks the
 * array's address up once (cudaGetSymbolAddress, with that device current -- make it before a graph capture).  EVERY
 * call on a device shares the one array: calls that overlap on the GPU (two streams, two threads) race -- serialise them
 * on one stream.  Nothing in the array needs clearing between calls (each call overwrites what it reads).          */
#define LIFTER_LOSS_STATIC_MAX_ROWS 65536
int lifter_loss_static(const float* pred, const float* target, int64_t rows, int joints, int kind, int64_t global_rows,
                       const float* stdv, float* dpred, float* out, int accumulate_metric, void* stream);
/* L1 building blocks of TriangleLoss / TriangleLoss_sep (losses.py:10-119); centre = the reference's
 * "t[1:] -= t[0]" on both tensors.  scratch: ceil(n / 2048) floats.                                      */
int lifter_l1_sum(const float* a, const float* b, int64_t n, int row_len, int centre, float scale, float* scratch,
                  float* out, void* stream);
int lifter_l1_grad(const float* a, const float* b, int64_t n, int row_len, int centre, const float* gscale,
                   float scale, float* da, float* db, void* stream);

/* flip_pose (phase3_direct/my_HybrIK/utils.py:372-396): x -> 1 - x (2D) or -x (3D) and left/right joints swapped;
 * in/out [poses][17][dims], dims 2 or 3, out != in.  grad = 1: the backward map (no "+1").                    */
int lifter_flip_pose(const float* in, float* out, int64_t poses, int joints, int dims, int grad, void* stream);
/* Keypoint ingest: coco2h36m (phase1_lifting/video2keypoints.py:14-57) fused with the pixel scaling (/1000) and the
 * input normalisation of main.py:331-337.  in [poses][17][in_channels] (x, y[, confidence]) in COCO order, out
 * [poses][17][2] in H36M order = norm(scale * joint); mode -1 none, 0 (v - mean) / std (34 statistics), 2: 2v - 1. */
int lifter_coco2h36m(const float* in, int in_channels, float* out, int64_t poses, float scale, const float* mean,
                     const float* stdv, int mode, void* stream);
/* Dataset-side keypoint processing (phase1_lifting/main.py:271-355 process_data) without Python loops over frames.
 * lifter_keypoint_stats: data [frames][joints][dims] fp32 -> out [4][joints*dims] fp64 = mean, population standard
 * deviation (divide by frames, main.py:280-289), min, max (train_5.py:35-40) of the values after the optional
 * root-centring "frame[1:] -= frame[0]" (main.py:275-277).  One read of the data; deterministic (no atomics).
 * lifter_process_keypoints: root-centring, then mode 0 (v - mean) / std with `stats` = the array above (main.py:334,
 * 337), 2: 2v - 1 (main.py:331), -1 none; root_mode 0 keep, 1 zero the root row (main.py:343-344), 2 drop it (the
 * 16-joint variant, main.py:340-342: out is [frames][joints-1][dims]).  joints * dims <= 64.                     */
#define LIFTER_ROOT_KEEP 0
#define LIFTER_ROOT_ZERO 1
#define LIFTER_ROOT_DROP 2
int64_t lifter_keypoint_stats_scratch_bytes(int64_t frames, int joints, int dims);
int lifter_keypoint_stats(const float* data, int64_t frames, int joints, int dims, int zero_centre, void* scratch,
                          double* out, void* stream);
int lifter_process_keypoints(const float* in, float* out, int64_t frames, int joints, int dims, int zero_centre,
                             const double* stats, int mode, int root_mode, void* stream);
/* Volumetric Gaussian heat-map targets of the dataset (phase3_direct/my_HybrIK/H36_dataset.py:143-202:
 * xyz_to_uvw, _keypoint_to_heatmap_3D, keypoints_to_heatmap_3D).  keypoints [n][3] (x, y, z) in [-1, 1], fp32 or fp64
 * (keypoints_are_f64), n = frames * joints -> out [n][64][64][64] fp32: (u, v, w) = (-y, -z, x), k = 31.5 (1 + uvw),
 * c = rint(k); the cube of ceil(6 sigma) (made odd) voxels around c holds exp(-|voxel - k|^2 / (2 sigma^2)), clipped
 * at the borders, zero elsewhere (the reference's sigma is 0.5: 3 x 3 x 3 voxels).  fp64 arithmetic like the numpy
 * original, fp32 on the store; one streaming pass, 1 MB per key point.  status: nullable device int, bit 0 set when a
 * coordinate is outside [-1, 1] (the reference asserts, :156-157).                                                  */
int lifter_keypoint_heatmaps(const void* keypoints, int keypoints_are_f64, int64_t n, double sigma, float* out,
                             int* status, void* stream);
/* per-column statistics broadcast over rows: mode 0 (x - a) / b (main.py:334,337), 1 x * b + a (main.py:473-479,
 * train_5.py:42-47), 2 2x - 1 (main.py:331), 3 x * (a - b) + b with a = max, b = min (train_5.py:35-40)        */
int lifter_affine_cols(const float* in, float* out, int64_t rows, int cols, const float* a, const float* b, int mode,
                       void* stream);
/* ONE pass over the predicted and target joints: de-standardisation, mean MPJPE (two2three.py:389-392), per-joint
 * MPJPE sums (losses.py:3-7), pinhole reprojection L1 against 2D keypoints in pixel/1000 and bone-length L1 over the
 * 16 edges of the H36M-17 skeleton (utils.py:37,131-172; these two terms have no reference implementation: PARITY
 * UNPINNED), and d total / d pred.  out[0] = w_mpjpe * out[1] + w_reproj * out[2] +
 * w_bone * out[3]; out[4 .. 4 + joints) = per-joint sums over the local rows.                                  */
typedef struct lifter_pose_loss {
    const float* pred; const float* target;   /* [rows][joints][3]                                          */
    const float* mean; const float* stdv;     /* [joints*3], nullable: joints = v * std + mean               */
    const float* root;                        /* [rows][3], nullable: camera-frame root added before projecting */
    const float* intrinsics;                  /* [rows][4] fx fy cx cy (pixels), nullable: no reprojection    */
    const float* target2d;                    /* [rows][joints][2] pixel/1000                                */
    int64_t rows, global_rows; int32_t joints;
    float w_mpjpe, w_reproj, w_bone;
    float* dpred;                             /* [rows][joints][3], nullable                                 */
    float* scratch;                           /* lifter_pose_loss_scratch_bytes()                            */
    float* out;                               /* [4 + joints]                                                */
} lifter_pose_loss_t;
int64_t lifter_pose_loss_scratch_bytes(int64_t rows, int joints);
int lifter_pose_loss(const lifter_pose_loss_t* desc, void* stream);

/* The loss of the phase-5 loop around the lifter (phase5_loop/backup_june8.py:137-161, train_5.py:145-181) as ONE
 * kernel + one finalize: flip averaging lift = (flip_pose(lifter(flip_pose(x))) + lifter(x)) / 2 of both lifts, the L1
 * terms of TriangleLoss (mode 0, losses.py:24-53: L1(pred2d, gt2d) + L1(pred3d, gt3d) + L1(lift_pred, pred3d)) or
 * TriangleLoss_sep (mode 1, losses.py:81-110: L1(pred2d, gt2d) + L1(pred3d, gt3d) + L1(lift_gt, gt3d) +
 * L1(lift_pred, lift_gt)), Project = False, and every gradient the loop back-propagates: w.r.t. the (up to) four lifter
 * outputs, the predicted 2D key points (the loss's own share; the lifter's input gradient is added by
 * lifter_phase5_input_grad) and the predicted 3D pose.  Optional terms on lift_pred with NO reference implementation
 * (PARITY UNPINNED): pinhole reprojection L1 and bone-length L1, as in lifter_pose_loss.
 * out[8] = total, loss_2d, loss_3d, loss_lift, domain gap (mode 1), reprojection, bone, 0 -- the reference returns
 * (total, loss_2d, loss_3d, loss_lift, loss_proj) and reads each with .cpu().item() every step (losses.py:48-51).   */
typedef struct lifter_phase5_loss {
    const float* out_pred;        /* [rows][joints*3] lifter(pred 2D)                                              */
    const float* out_pred_flip;   /* lifter(flip_pose(pred 2D)); nullable: Flip = False                            */
    const float* out_gt;          /* lifter(gt 2D); nullable in mode 0 (TriangleLoss does not read it)             */
    const float* out_gt_flip;
    const float* pred2d; const float* gt2d;      /* [rows][joints*2]                                               */
    const float* pred3d; const float* gt3d;      /* [rows][joints*3]                                               */
    int32_t mode;
    const float* mean3d; const float* std3d;     /* [joints*3] nullable: joints = v * std + mean (unpinned terms)  */
    const float* mean2d; const float* std2d;     /* [joints*2] nullable: gt2d * std + mean = key points, pixel/1000 */
    const float* root;                           /* [rows][3] nullable                                             */
    const float* intrinsics;                     /* [rows][4] fx fy cx cy (pixels), nullable: no reprojection      */
    float w_reproj, w_bone;
    int64_t rows, global_rows; int32_t joints;
    float* d_out_pred; float* d_out_pred_flip; float* d_out_gt; float* d_out_gt_flip;
    float* d_pred2d; float* d_pred3d;
    float* scratch;                              /* lifter_phase5_loss_scratch_bytes()                             */
    float* out;                                  /* [8]                                                            */
} lifter_phase5_loss_t;
int64_t lifter_phase5_loss_scratch_bytes(int64_t rows);
int lifter_phase5_loss(const lifter_phase5_loss_t* desc, void* stream);
/* out = d_own + dx + flip_pose-VJP(dx_flip): gradient w.r.t. the predicted 2D key points from the loss itself, the plain
 * lifter forward and the flipped one (dx_flip nullable).  [rows][joints*2]                                          */
int lifter_phase5_input_grad(const float* d_own, const float* dx, const float* dx_flip, int64_t rows, int joints,
                             float* out, void* stream);
/* dst = a + b (+ c + d): sums the flat parameter gradients of the backward passes of one step; n %% 4 == 0.        */
int lifter_sum_buffers(float* dst, const float* a, const float* b, const float* c, const float* d, int64_t n,
                       void* stream);

/* torch.optim.Adam / AdamW (train_1.py:39) over the flat buffer; step is 1-based; n %% 4 == 0.           */
int lifter_adam_step(float* p, const float* g, float* m, float* v, int64_t n, float lr, float beta1, float beta2,
                     float eps, float weight_decay, int decoupled, int step, float grad_scale, void* stream);

/* Graph-replayable optimiser step: advances the device step state (call once at the top of every step) ...      */
int lifter_step_advance(lifter_step_state_t* state, float beta1, float beta2, float weight_decay, void* stream);
/* ... and Adam/AdamW over the flat buffer of `cfg` that ALSO refreshes the 16-bit GEMM operand copies of the Linear
 * weights (operands nullable) in the same pass.  state (device, nullable) overrides lr/step with the step state.   */
int lifter_adam_step_packed(const lifter_config_t* cfg, float* p, const float* g, float* m, float* v, float lr,
                            float beta1, float beta2, float eps, float weight_decay, int decoupled, int step,
                            float grad_scale, const lifter_step_state_t* state, void* operands, int fmt, void* stream);

/* The same fused pass for ANY flat fp32 buffer (the MyViT trainer uses it): Adam / AdamW over p[0 .. n) and, in the same
 * pass, the 16-bit GEMM operand copies of up to 16 DENSE ranges of it: range t covers p[p_off[t] .. p_off[t] + numel[t])
 * and is written to operands[op_off[t] ..] (elements; every offset and length a multiple of 4).                    */
typedef struct lifter_adam_table {
    int32_t n;                                   /* ranges in use, <= 16                                          */
    int64_t p_off[16]; int64_t numel[16]; int64_t op_off[16];
} lifter_adam_table_t;
int lifter_adam_step_table(float* p, const float* g, float* m, float* v, int64_t n, float lr, float beta1, float beta2,
                           float eps, float weight_decay, int decoupled, int step, float grad_scale,
                           const lifter_step_state_t* state, void* operands, int fmt, const lifter_adam_table_t* table,
                           void* stream);

/* Data-parallel optimiser step as ONE kernel over NVLink peer memory (replaces gradient all-reduce + Adam): rank r
 * sums slice r of every rank's gradient with P2P loads (rank order: deterministic), runs Adam/AdamW on that slice
 * (m, v: only slice r is touched on rank r) and stores the new parameters and operand copies into every rank's
 * buffers; flag barriers inside the kernel order it against the peers' backward and next forward.                 */
int lifter_dp_adam_step(const lifter_config_t* cfg, const lifter_dp_t* dp, float* m, float* v, float beta1, float beta2,
                        float eps, float weight_decay, int decoupled, float grad_scale,
                        const lifter_step_state_t* state, int fmt, void* stream);

/* The same kernel for ANY flat fp32 buffer of n parameters (n % 4 == 0) with a table of dense 16-bit operand ranges (the
 * data-parallel MyViT trainer): dp->grad / params / operands / flags must be set, dp->sync is not read.             */
int lifter_dp_adam_step_table(const lifter_dp_t* dp, int64_t n, float* m, float* v, float beta1, float beta2, float eps,
                              float weight_decay, int decoupled, float grad_scale, const lifter_step_state_t* state,
                              int fmt, const lifter_adam_table_t* table, void* stream);

/* Backward pass WITH the optimiser folded in (needs ctx->step_state): as soon as a layer's gradients are complete —
 * weight gradient reduced, its data-gradient GEMM done with the old weights — Adam/AdamW runs on that layer's range of
 * the flat buffer on a side stream while the layers below are still in backward.  With ctx->dp (world > 1, peer
 * gradient / parameter / operand / flag buffers set) each range goes through the fused NVLink reduce-scatter + Adam +
 * all-gather kernel instead, so the gradient exchange overlaps backward too.  Replaces lifter_train_backward followed
 * by lifter_adam_step_packed / lifter_dp_adam_step; the caller still advances the step state once per step.        */
typedef struct lifter_optim {
    float* params;               /* writable flat fp32 parameters (the buffer ctx->params points at)              */
    void* operands;              /* writable 16-bit operand copies (ctx->operands), refreshed in the same pass     */
    float* m; float* v;          /* flat Adam state                                                                */
    float beta1, beta2, eps, weight_decay;
    int32_t decoupled;           /* 1: AdamW                                                                       */
    float grad_scale;
} lifter_optim_t;
int lifter_train_backward_step(const lifter_config_t* cfg, const lifter_train_ctx_t* ctx, const lifter_optim_t* opt,
                               void* stream);

/* ---- MyViT lifter, eval-mode forward (phase1_lifting/baselineModel.py:307-362: the model train_1.py:35 and
 * train_5.py:58,63 instantiate).  Joints are tokens: Linear(in_dim -> hidden) + positional embedding, `blocks` x
 * [x + to_out(softmax(q k^T / sqrt(d)) v) with q, k, v = to_qkv(mhsa.norm(norm1(x))); x + mlp(norm2(x)), exact GELU],
 * then Linear(hidden -> hidden / 2) -> ReLU -> Linear(-> out_dim) per token.  Every Linear over the [poses * tokens][.]
 * token matrix is the tcgen05 GEMM with bias / ReLU / GELU / residual in its epilogue; LayerNorm, attention and the
 * embedding are separate kernels.  16-bit activations live in `workspace` (lifter_vit_workspace_bytes).  Limits:
 * tokens <= 32, in_dim <= 4, hidden in {256, 512, 768, 1024}, head width 32 or 64, out_dim <= 64.               */
typedef struct lifter_vit {
    int32_t tokens, in_dim, hidden, heads, blocks, mlp_dim, head_dim, out_dim;
    int32_t fmt;                  /* LIFTER_FMT_F16 / LIFTER_FMT_BF16: GEMM operands and stored activations       */
    float ln_eps;                 /* torch.nn.LayerNorm default 1e-5                                              */
    const float* map_w; const float* map_b; const float* pos;   /* [hidden][in_dim], [hidden], [tokens][hidden]     */
    const float* ln;              /* [blocks][6][hidden]: norm1 w, b; mhsa.norm w, b; norm2 w, b                   */
    const float* mlp_b1;          /* [blocks][mlp_dim]                                                             */
    const float* mlp_b2;          /* [blocks][hidden]                                                              */
    const float* head_b1;         /* [head_dim]                                                                    */
    const float* head_b2;         /* [64], zero past out_dim                                                       */
    const void* w_qkv;            /* 16-bit [blocks][3 * hidden][hidden]   (Linear.weight layout: [out][in])       */
    const void* w_out;            /* 16-bit [blocks][hidden][hidden]                                               */
    const void* w_mlp1;           /* 16-bit [blocks][mlp_dim][hidden]                                              */
    const void* w_mlp2;           /* 16-bit [blocks][hidden][mlp_dim]                                              */
    const void* w_head1;          /* 16-bit [head_dim][hidden]                                                     */
    const void* w_head2;          /* 16-bit [64][head_dim], rows past out_dim zero                                 */
} lifter_vit_t;
int64_t lifter_vit_workspace_bytes(const lifter_vit_t* vit, int64_t poses);
/* x [poses][tokens][in_dim] fp32 -> out [poses][tokens][out_dim] fp32; asynchronous on `stream`.                  */
int lifter_vit_forward(const lifter_vit_t* vit, const float* x, int64_t poses, void* workspace, float* out, void* stream);

/* ---- MyViT lifter, training path (loss.backward() through baselineModel.py:228-362; loop body train_1.py:73-100 with
 * model_lift = MyViT).  No dropout and no BatchNorm in this model: lifter_vit_train_forward is the eval arithmetic that
 * keeps every GEMM input in `saved` (lifter_vit_saved_bytes); lifter_vit_backward turns d loss / d out into the gradient
 * of every trainable parameter (the positional table has requires_grad = False, baselineModel.py:330) and of the input.  vit->fmt is the
 * operand format of forward AND backward (bf16 for the gradient range).  hidden 256 or 512.                       */
typedef struct lifter_vit_grads {            /* fp32, stacked per block exactly like the parameters of lifter_vit_t   */
    float* map;                              /* [1 + in_dim][hidden]: d linear_mapper.bias, then d weight[:, i] per i  */
    float* ln;                               /* [blocks][6][hidden]                                                  */
    float* w_qkv; float* w_out; float* w_mlp1; float* mlp_b1; float* w_mlp2; float* mlp_b2;
    float* w_head1; float* head_b1;
    float* w_head2;                          /* [out_dim][head_dim]                                                  */
    float* head_b2;                          /* [out_dim]                                                            */
} lifter_vit_grads_t;
int64_t lifter_vit_saved_bytes(const lifter_vit_t* vit, int64_t poses);
int64_t lifter_vit_bwd_scratch_bytes(const lifter_vit_t* vit, int64_t poses);
int lifter_vit_train_forward(const lifter_vit_t* vit, const float* x, int64_t poses, void* saved, float* out, void* stream);
/* dx: nullable, [poses][tokens][in_dim] fp32 = d loss / d key points (the phase-5 loop feeds the lifter from a 2D model) */
int lifter_vit_backward(const lifter_vit_t* vit, const float* x, int64_t poses, const void* saved, const float* dout,
                        void* scratch, const lifter_vit_grads_t* grads, float* dx, void* stream);

/* lifter_vit_backward leaves d linear_mapper as [1 + in_dim][hidden] (bias row, then one row per input value); this
 * re-lays it as the module's parameters: d_weight [hidden][in_dim], d_bias [hidden] (the fused trainer's flat gradient). */
int lifter_vit_map_grad(const float* gmap, int in_dim, int hidden, float* d_weight, float* d_bias, void* stream);

/* ---- test hook: one tcgen05 GEMM  D[M][N] = sum_k A(m,k) B(n,k) ------------------------------------ */
typedef struct lifter_gemm_desc {
    const void* A; int64_t lda; const void* B; int64_t ldb;
    int32_t M, N, K, a_mn_major, b_mn_major, fmt;
    int32_t out_mode;            /* 0: 16-bit [M][ldd], 1: fp32 split-K slabs, 2: fp32 [M][n_valid] */
    void* D; int64_t ldd; int32_t n_valid; int32_t splits; int32_t relu;   /* relu: 0 none, 1 ReLU, 2 exact GELU */
    const float* scale; const float* shift; const void* residual; int64_t ldr;
    float* stat_sum; float* stat_sq;
    int32_t cluster;             /* 0 auto, 1 single-CTA tiles (cta_group::1), 2 CTA pairs (cta_group::2, 256-row tiles) */
    void* timeline;              /* debug: uint64 [grid][32] globaltimer stamps of the kernel's phases, or NULL */
    int32_t max_stages;          /* debug: cap on the operand ring depth (0 = all that fit in shared memory) */
} lifter_gemm_desc_t;
int lifter_gemm(const lifter_gemm_desc_t* desc, void* stream);
/* ---- debug hook: per-launch device times.  lifter_trace_begin() makes every later launch of this library record a
 * CUDA event; lifter_trace_end() synchronises and writes one "kernel-name microseconds" line per launch into buf
 * (returns the number of bytes written).  Not for production: the events serialise the launches.             */
int lifter_trace_begin(void* stream);
int lifter_trace_end(char* buf, int len);
int lifter_gemm_slab_rows(int M);
int lifter_gemm_stat_rows(int M);
int lifter_gemm_effective_splits(int K, int splits);

#ifdef __cplusplus
}
#endif
#endif /* LIFTER_B200_H */
