#include "phase5_ops.cuh"

#include "pose_math.cuh"

namespace lifter {

namespace {

// perm[j] = the joint whose value lands on j under flip_pose (utils.py:382-383; an involution)
__constant__ int c_perm[17] = {0, 4, 5, 6, 1, 2, 3, 7, 8, 9, 10, 14, 15, 16, 11, 12, 13};

// One thread per pose, PL_ROWS poses per block; every tensor of the block's rows goes through shared memory with
// coalesced loads / stores (a row is 136 or 204 bytes).  Tiles: the four lifter outputs, pred / gt 2D, pred / gt 3D; a
// gradient is written over the tile of the tensor it belongs to.
__global__ void __launch_bounds__(PL_ROWS) phase5_loss_kernel(Phase5LossArgs a) {
    griddep_launch();
    griddep_wait();
    extern __shared__ float sm[];
    const int J = a.J, O3 = 3 * J, O2 = 2 * J;
    float* s_op = sm;                          // out_pred       -> d_out_pred
    float* s_of = s_op + PL_ROWS * O3;         // out_pred_flip  -> d_out_pred_flip
    float* s_og = s_of + PL_ROWS * O3;         // out_gt         -> d_out_gt
    float* s_ogf = s_og + PL_ROWS * O3;        // out_gt_flip    -> d_out_gt_flip
    float* s_p3 = s_ogf + PL_ROWS * O3;        // pred3d         -> d_pred3d
    float* s_g3 = s_p3 + PL_ROWS * O3;         // gt3d
    float* s_p2 = s_g3 + PL_ROWS * O3;         // pred2d         -> d_pred2d
    float* s_g2 = s_p2 + PL_ROWS * O2;         // gt2d           -> key points in pixel / 1000 (reprojection target)
    float* s_red = s_g2 + PL_ROWS * O2;        // [6][PL_ROWS / 32]
    const long row0 = static_cast<long>(blockIdx.x) * PL_ROWS;
    const int nrows = static_cast<int>(min(static_cast<long>(PL_ROWS), a.rows - row0));
    const bool flip = a.out_pred_flip != nullptr, sep = a.mode == 1, have_gt = a.out_gt != nullptr;
    for (int i = threadIdx.x; i < nrows * O3; i += PL_ROWS) {
        s_op[i] = a.out_pred[row0 * O3 + i];
        if (flip) s_of[i] = a.out_pred_flip[row0 * O3 + i];
        if (have_gt) {
            s_og[i] = a.out_gt[row0 * O3 + i];
            if (flip) s_ogf[i] = a.out_gt_flip[row0 * O3 + i];
        }
        s_p3[i] = a.pred3d[row0 * O3 + i];
        s_g3[i] = a.gt3d[row0 * O3 + i];
    }
    for (int i = threadIdx.x; i < nrows * O2; i += PL_ROWS) {
        s_p2[i] = a.pred2d[row0 * O2 + i];
        s_g2[i] = a.gt2d[row0 * O2 + i];
    }
    __syncthreads();

    const int r = threadIdx.x;
    float l2d = 0.0f, l3d = 0.0f, llift = 0.0f, lgap = 0.0f, l_mp = 0.0f, l_rp = 0.0f, l_bn = 0.0f;
    if (r < nrows) {
        const float inv_rows = 1.0f / static_cast<float>(a.global_rows);
        const float g2 = inv_rows / O2, g3 = inv_rows / O3;        // L1Loss(mean): 1 / element count of the GLOBAL batch
        float* op = s_op + r * O3; float* of = s_of + r * O3; float* og = s_og + r * O3; float* ogf = s_ogf + r * O3;
        float* p3 = s_p3 + r * O3; float* g3p = s_g3 + r * O3; float* p2 = s_p2 + r * O2; float* g2p = s_g2 + r * O2;
        // ---- 2D term, and the reprojection target in pixel / 1000
        for (int i = 0; i < O2; ++i) {
            const float d = p2[i] - g2p[i];
            l2d += fabsf(d);
            p2[i] = pl_sgn(d) * g2;                                               // d_pred2d (own share)
            if (a.mean2d != nullptr) g2p[i] = fmaf(g2p[i], a.std2d[i], a.mean2d[i]);
        }
        // ---- flip averaging: lift = (flip_pose(out_flip) + out) / 2   (backup_june8.py:152-153; flip_pose 3D: x -> -x,
        //      left / right joints swapped).  lp / lg live in registers, joint by joint.
        float LP[PL_MAXJ][3], DLP[PL_MAXJ][3];
#pragma unroll
        for (int j = 0; j < PL_MAXJ; ++j) {
            if (j >= J) break;
            const int pj = (flip && J == 17) ? c_perm[j] : j;
#pragma unroll
            for (int d = 0; d < 3; ++d) {
                const float sgnx = d == 0 ? -1.0f : 1.0f;
                const float lp = flip ? 0.5f * (sgnx * of[3 * pj + d] + op[3 * j + d]) : op[3 * j + d];
                float lg = 0.0f;
                if (have_gt) lg = flip ? 0.5f * (sgnx * ogf[3 * pj + d] + og[3 * j + d]) : og[3 * j + d];
                const float e3 = p3[3 * j + d] - g3p[3 * j + d];
                l3d += fabsf(e3);
                float dp3 = pl_sgn(e3) * g3, dlp, dlg = 0.0f;
                if (!sep) {                                   // TriangleLoss: + L1(lift_pred, pred3d)
                    const float e = lp - p3[3 * j + d];
                    llift += fabsf(e);
                    dlp = pl_sgn(e) * g3;
                    dp3 -= dlp;
                } else {                                      // TriangleLoss_sep: + L1(lift_gt, gt3d) + L1(lift_pred, lift_gt)
                    const float e1 = lg - g3p[3 * j + d], e2 = lp - lg;
                    llift += fabsf(e1);
                    lgap += fabsf(e2);
                    dlp = pl_sgn(e2) * g3;
                    dlg = pl_sgn(e1) * g3 - dlp;
                }
                LP[j][d] = lp; DLP[j][d] = dlp;
                p3[3 * j + d] = dp3;                          // d_pred3d
                if (have_gt) og[3 * j + d] = dlg;             // d_lift_gt, split over the two outputs below
            }
        }
        // (no warp-level sync here: a thread touches only its own row, and rows past the batch skip this branch)
        // ---- unpinned terms on the lifted pose (same per-pose arithmetic as the fused pose-loss kernel)
        if (a.intrinsics != nullptr || a.w_bone != 0.0f) {
            PoseRowArgs ra;
            ra.J = J; ra.mean = a.mean3d; ra.stdv = a.std3d;
            ra.g_mp = 0.0f; ra.g_rp = a.w_reproj * inv_rows / (2 * J); ra.g_bn = a.w_bone * inv_rows / 16.0f;
            ra.reproj = a.intrinsics != nullptr; ra.bone = a.w_bone != 0.0f;
            ra.rx = ra.ry = ra.rz = ra.fx = ra.fy = ra.cx = ra.cy = 0.0f;
            if (a.root != nullptr) { ra.rx = a.root[(row0 + r) * 3]; ra.ry = a.root[(row0 + r) * 3 + 1]; ra.rz = a.root[(row0 + r) * 3 + 2]; }
            if (a.intrinsics != nullptr) {
                const float4 k = *reinterpret_cast<const float4*>(a.intrinsics + (row0 + r) * 4);
                ra.fx = k.x; ra.fy = k.y; ra.cx = k.z; ra.cy = k.w;
            }
            float G[PL_MAXJ][3];
            // lift_pred of this pose as a contiguous row: the output tile of the plain forward is free now
#pragma unroll
            for (int j = 0; j < PL_MAXJ; ++j) {
                if (j >= J) break;
#pragma unroll
                for (int d = 0; d < 3; ++d) op[3 * j + d] = LP[j][d];
            }
            pose_row_terms(ra, op, g3p, g2p, G, l_mp, l_rp, l_bn, nullptr);
#pragma unroll
            for (int j = 0; j < PL_MAXJ; ++j) {
                if (j >= J) break;
#pragma unroll
                for (int d = 0; d < 3; ++d) DLP[j][d] += G[j][d] * (a.std3d != nullptr ? a.std3d[3 * j + d] : 1.0f);
            }
        }
        // ---- back through the flip averaging: d_out = d_lift / 2, d_out_flip[perm(j)] = S d_lift[j] / 2
        if (have_gt && flip) {
#pragma unroll
            for (int j = 0; j < PL_MAXJ; ++j) {
                if (j >= J) break;
                const int pj = J == 17 ? c_perm[j] : j;
#pragma unroll
                for (int d = 0; d < 3; ++d) ogf[3 * pj + d] = (d == 0 ? -0.5f : 0.5f) * og[3 * j + d];
            }
#pragma unroll
            for (int j = 0; j < PL_MAXJ; ++j) {
                if (j >= J) break;
#pragma unroll
                for (int d = 0; d < 3; ++d) og[3 * j + d] *= 0.5f;
            }
        }
#pragma unroll
        for (int j = 0; j < PL_MAXJ; ++j) {
            if (j >= J) break;
            const int pj = (flip && J == 17) ? c_perm[j] : j;
#pragma unroll
            for (int d = 0; d < 3; ++d) {
                op[3 * j + d] = flip ? 0.5f * DLP[j][d] : DLP[j][d];
                if (flip) of[3 * pj + d] = (d == 0 ? -0.5f : 0.5f) * DLP[j][d];
            }
        }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < nrows * O3; i += PL_ROWS) {
        a.d_out_pred[row0 * O3 + i] = s_op[i];
        if (flip) a.d_out_pred_flip[row0 * O3 + i] = s_of[i];
        if (have_gt && a.d_out_gt != nullptr) {
            a.d_out_gt[row0 * O3 + i] = s_og[i];
            if (flip) a.d_out_gt_flip[row0 * O3 + i] = s_ogf[i];
        }
        a.d_pred3d[row0 * O3 + i] = s_p3[i];
    }
    for (int i = threadIdx.x; i < nrows * O2; i += PL_ROWS) a.d_pred2d[row0 * O2 + i] = s_p2[i];
    // block partials: l2d, l3d, lift, gap, reprojection, bone
    float v[6] = {l2d, l3d, llift, lgap, l_rp, l_bn};
#pragma unroll
    for (int k = 0; k < 6; ++k) {
        v[k] = pl_warp_sum(v[k]);
        if ((threadIdx.x & 31) == 0) s_red[k * (PL_ROWS / 32) + (threadIdx.x >> 5)] = v[k];
    }
    __syncthreads();
    if (threadIdx.x < 8) {
        float t = 0.0f;
        if (threadIdx.x < 6)
            for (int w = 0; w < PL_ROWS / 32; ++w) t += s_red[threadIdx.x * (PL_ROWS / 32) + w];
        a.partials[static_cast<long>(blockIdx.x) * 8 + threadIdx.x] = t;
    }
}

// single block: fixed-order fp64 reduction of the block partials, then the means and the total
__global__ void phase5_finalize_kernel(const float* __restrict__ partials, int nblk, Phase5LossArgs a) {
    griddep_launch();
    griddep_wait();
    __shared__ double sh[32][8];
    const int c = threadIdx.x % 8, lane_r = threadIdx.x / 8;             // 8 columns x 32 row lanes
    double s = 0.0;
    for (int b = lane_r; b < nblk; b += 32) s += static_cast<double>(partials[static_cast<long>(b) * 8 + c]);
    sh[lane_r][c] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
        double tot[8];
        for (int k = 0; k < 8; ++k) {
            double t = 0.0;
            for (int r = 0; r < 32; ++r) t += sh[r][k];
            tot[k] = t;
        }
        const double n = static_cast<double>(a.global_rows);
        const double l2d = tot[0] / (n * a.J * 2), l3d = tot[1] / (n * a.J * 3), lift = tot[2] / (n * a.J * 3),
                     gap = tot[3] / (n * a.J * 3), rp = tot[4] / (n * a.J * 2), bn = tot[5] / (n * 16);
        a.out[0] = static_cast<float>(l2d + l3d + lift + gap + a.w_reproj * rp + a.w_bone * bn);
        a.out[1] = static_cast<float>(l2d); a.out[2] = static_cast<float>(l3d); a.out[3] = static_cast<float>(lift);
        a.out[4] = static_cast<float>(gap); a.out[5] = static_cast<float>(rp); a.out[6] = static_cast<float>(bn);
        a.out[7] = 0.0f;
    }
}

__global__ void phase5_input_grad_kernel(const float* __restrict__ d_own, const float* __restrict__ dx,
                                         const float* __restrict__ dx_flip, long n, int J, float* __restrict__ out) {
    griddep_launch();
    griddep_wait();
    const long i = static_cast<long>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float v = d_own[i] + dx[i];
    if (dx_flip != nullptr) {               // x_flip[j] = (1 - x[perm j].x, x[perm j].y): dL/dx[k] += S dx_flip[perm k]
        const int O2 = 2 * J;
        const long row = i / O2;
        const int e = static_cast<int>(i - row * O2), j = e >> 1, d = e & 1;
        const int pj = J == 17 ? c_perm[j] : j;
        v += (d == 0 ? -1.0f : 1.0f) * dx_flip[row * O2 + 2 * pj + d];
    }
    out[i] = v;
}

__global__ void sum_buffers_kernel(float* __restrict__ dst, const float* __restrict__ a, const float* __restrict__ b,
                                   const float* __restrict__ c, const float* __restrict__ d, long n4) {
    griddep_launch();
    griddep_wait();
    const long stride = static_cast<long>(gridDim.x) * blockDim.x;
    for (long i = static_cast<long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n4; i += stride) {
        float4 s = reinterpret_cast<const float4*>(a)[i];
        const float4 t = reinterpret_cast<const float4*>(b)[i];
        s.x += t.x; s.y += t.y; s.z += t.z; s.w += t.w;
        if (c != nullptr) { const float4 u = reinterpret_cast<const float4*>(c)[i]; s.x += u.x; s.y += u.y; s.z += u.z; s.w += u.w; }
        if (d != nullptr) { const float4 u = reinterpret_cast<const float4*>(d)[i]; s.x += u.x; s.y += u.y; s.z += u.z; s.w += u.w; }
        reinterpret_cast<float4*>(dst)[i] = s;
    }
}

}  // namespace

int phase5_partial_rows(int64_t rows) { return static_cast<int>((rows + PL_ROWS - 1) / PL_ROWS); }

int phase5_loss(const Phase5LossArgs& a, cudaStream_t st) {
    LIFTER_REQUIRE(a.J >= 1 && a.J <= PL_MAXJ, "phase5_loss: %d joints unsupported (1..%d)", a.J, PL_MAXJ);
    LIFTER_REQUIRE(a.rows >= 1 && a.global_rows >= a.rows, "phase5_loss: empty batch");
    LIFTER_REQUIRE(a.mode == 0 || a.mode == 1, "phase5_loss: mode %d (0 TriangleLoss, 1 TriangleLoss_sep)", a.mode);
    LIFTER_REQUIRE(a.out_pred && a.pred2d && a.gt2d && a.pred3d && a.gt3d && a.d_out_pred && a.d_pred2d && a.d_pred3d &&
                   a.partials && a.out, "phase5_loss: null pointer");
    LIFTER_REQUIRE(a.out_pred_flip == nullptr || a.d_out_pred_flip != nullptr, "phase5_loss: flipped output without a gradient buffer");
    LIFTER_REQUIRE(a.out_pred_flip == nullptr || a.J == 17, "phase5_loss: flip averaging is defined on the 17-joint H36M skeleton");
    LIFTER_REQUIRE(a.mode == 0 || (a.out_gt != nullptr && a.d_out_gt != nullptr), "phase5_loss: TriangleLoss_sep needs lift(gt 2D) and its gradient buffer");
    LIFTER_REQUIRE(a.out_gt == nullptr || (a.out_gt_flip != nullptr) == (a.out_pred_flip != nullptr), "phase5_loss: flip both lifts or neither");
    LIFTER_REQUIRE(a.out_gt == nullptr || a.out_gt_flip == nullptr || a.d_out_gt == nullptr || a.d_out_gt_flip != nullptr,
                   "phase5_loss: flipped gt output without a gradient buffer");
    LIFTER_REQUIRE((a.mean3d == nullptr) == (a.std3d == nullptr) && (a.mean2d == nullptr) == (a.std2d == nullptr),
                   "phase5_loss: mean and std come together");
    LIFTER_REQUIRE(a.w_bone == 0.0f || a.J == 17, "phase5_loss: the bone-length term is defined on the 17-joint H36M skeleton");
    static PerDeviceOnce once;
    bool& configured = *once.slot();
    const size_t smem = (static_cast<size_t>(PL_ROWS) * (6 * 3 * a.J + 2 * 2 * a.J) + 6 * (PL_ROWS / 32) + 4) * sizeof(float);
    LIFTER_REQUIRE(smem <= 200 * 1024, "phase5_loss: tile does not fit shared memory");
    if (!configured) {
        LIFTER_CHECK_CUDA(cudaFuncSetAttribute(phase5_loss_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        configured = true;
    }
    const int nblk = phase5_partial_rows(a.rows);
    LIFTER_CHECK_CUDA(launch_pdl("phase5_loss_kernel", phase5_loss_kernel, dim3(nblk), dim3(PL_ROWS), smem, st, a));
    LIFTER_CHECK_CUDA(launch_pdl("phase5_finalize_kernel", phase5_finalize_kernel, dim3(1), dim3(256), 0, st, a.partials, nblk, a));
    return LIFTER_OK;
}

int phase5_input_grad(const float* d_own, const float* dx, const float* dx_flip, int64_t rows, int J, float* out,
                      cudaStream_t st) {
    LIFTER_REQUIRE(d_own && dx && out, "phase5_input_grad: null pointer");
    LIFTER_REQUIRE(dx_flip == nullptr || J == 17, "phase5_input_grad: flip_pose is defined on the 17-joint H36M skeleton");
    const long n = rows * 2 * J;
    if (n == 0) return LIFTER_OK;
    LIFTER_CHECK_CUDA(launch_pdl("phase5_input_grad_kernel", phase5_input_grad_kernel, dim3(static_cast<unsigned>((n + 255) / 256)), dim3(256), 0, st,
                                 d_own, dx, dx_flip, n, J, out));
    return LIFTER_OK;
}

int sum_buffers(float* dst, const float* a, const float* b, const float* c, const float* d, int64_t n, cudaStream_t st) {
    LIFTER_REQUIRE(dst && a && b && n > 0 && n % 4 == 0, "sum_buffers: null pointer or length not a multiple of 4");
    long blocks = (n / 4 + 255) / 256;
    const long cap = static_cast<long>(num_sms()) * 8;
    if (blocks > cap) blocks = cap;
    LIFTER_CHECK_CUDA(launch_pdl("sum_buffers_kernel", sum_buffers_kernel, dim3(static_cast<unsigned>(blocks)), dim3(256), 0, st, dst, a, b, c, d,
                                 static_cast<long>(n / 4)));
    return LIFTER_OK;
}

}  // namespace lifter
