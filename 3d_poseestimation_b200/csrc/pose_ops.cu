#include "pose_ops.cuh"

#include "pose_math.cuh"

namespace lifter {

namespace {

// left/right joints of the H36M-17 skeleton (utils.py:382-383): perm[j] = the joint whose value lands on j
__constant__ int c_flip_perm[17] = {0, 4, 5, 6, 1, 2, 3, 7, 8, 9, 10, 14, 15, 16, 11, 12, 13};
// (the 16 unique parent -> child edges of utils.py:37 -- sk_points lists [5,6] twice -- are spelled out in the kernel)

__global__ void flip_pose_kernel(const float* __restrict__ in, float* __restrict__ out, long n, int D, int grad) {
    griddep_launch();
    griddep_wait();
    const long idx = static_cast<long>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (idx >= n) return;
    const int jd = 17 * D;
    const long b = idx / jd;
    const int r = static_cast<int>(idx - b * jd), j = r / D, d = r - j * D;
    float v = in[b * jd + c_flip_perm[j] * D + d];
    if (d == 0) v = (D == 2 && !grad) ? 1.0f - v : -v;
    out[idx] = v;
}

__global__ void affine_cols_kernel(const float* __restrict__ in, float* __restrict__ out, long n, int cols,
                                   const float* __restrict__ a, const float* __restrict__ b, int mode) {
    griddep_launch();
    griddep_wait();
    const long stride = static_cast<long>(gridDim.x) * blockDim.x;
    for (long i = static_cast<long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride) {
        const int c = static_cast<int>(i % cols);
        const float x = in[i];
        float y;
        if (mode == COLS_NORMALISE) y = (x - a[c]) / b[c];
        else if (mode == COLS_DENORMALISE) y = x * b[c] + a[c];
        else if (mode == COLS_MINMAX_NORM) y = 2.0f * x - 1.0f;
        else y = x * (a[c] - b[c]) + b[c];
        out[i] = y;
    }
}

// COCO-17 -> H36M-17 (video2keypoints.py:14-57): each H36M joint is one COCO joint or the midpoint of two; the belly
// (7) is the midpoint of the two midpoints root (0) and neck (8), evaluated in the reference's order of operations.
__constant__ int c_coco_a[17] = {11, 12, 14, 16, 11, 13, 15, -1, 5, 0, 1, 5, 7, 9, 6, 8, 10};
__constant__ int c_coco_b[17] = {12, 12, 14, 16, 11, 13, 15, -1, 6, 0, 2, 5, 7, 9, 6, 8, 10};

__global__ void coco2h36m_kernel(const float* __restrict__ in, int C, float* __restrict__ out, long poses, float scale,
                                 const float* __restrict__ mean, const float* __restrict__ stdv, int mode) {
    griddep_launch();
    griddep_wait();
    const long idx = static_cast<long>(blockIdx.x) * blockDim.x + threadIdx.x;      // one (pose, joint, coord)
    if (idx >= poses * 34) return;
    const long b = idx / 34;
    const int r = static_cast<int>(idx - b * 34), j = r >> 1, d = r & 1;
    const float* x = in + b * 17 * C + d;
    auto mid = [&](int jj) {
        const int a = c_coco_a[jj], bb = c_coco_b[jj];
        return a == bb ? x[a * C] : (x[a * C] + x[bb * C]) * 0.5f;
    };
    float v = j == 7 ? (mid(0) + mid(8)) * 0.5f : mid(j);
    v *= scale;
    if (mode == COLS_NORMALISE) v = (v - mean[r]) / stdv[r];
    else if (mode == COLS_MINMAX_NORM) v = 2.0f * v - 1.0f;
    out[idx] = v;
}

__device__ __forceinline__ float warp_sum(float v) { return pl_warp_sum(v); }

// One thread per pose; the block's rows are staged through shared memory with coalesced loads / stores (rows are
// 204 bytes: a thread-per-row global access pattern would touch 32 sectors per instruction).
__global__ void __launch_bounds__(PL_ROWS) pose_loss_kernel(PoseLossArgs a) {
    griddep_launch();
    griddep_wait();
    extern __shared__ float sm[];
    const int J = a.J, O = 3 * J;
    float* s_pred = sm;                             // [PL_ROWS][O]   (reused for dpred)
    float* s_tgt = s_pred + PL_ROWS * O;            // [PL_ROWS][O]
    float* s_t2d = s_tgt + PL_ROWS * O;             // [PL_ROWS][2J]
    float* s_metric = s_t2d + PL_ROWS * 2 * J;      // [PL_MAXJ]
    float* s_red = s_metric + PL_MAXJ;              // [3][PL_ROWS / 32]
    const long row0 = static_cast<long>(blockIdx.x) * PL_ROWS;
    const int nrows = static_cast<int>(min(static_cast<long>(PL_ROWS), a.rows - row0));
    for (int i = threadIdx.x; i < nrows * O; i += PL_ROWS) {
        s_pred[i] = a.pred[row0 * O + i];
        s_tgt[i] = a.target[row0 * O + i];
    }
    if (a.intrinsics != nullptr)
        for (int i = threadIdx.x; i < nrows * 2 * J; i += PL_ROWS) s_t2d[i] = a.target2d[row0 * 2 * J + i];
    if (threadIdx.x < PL_MAXJ) s_metric[threadIdx.x] = 0.0f;
    __syncthreads();

    const int r = threadIdx.x;
    const bool live = r < nrows;
    float l_mp = 0.0f, l_rp = 0.0f, l_bn = 0.0f;
    float G[PL_MAXJ][3];
    const float inv_rows = 1.0f / static_cast<float>(a.global_rows);
    if (live) {
        PoseRowArgs ra;
        ra.J = J; ra.mean = a.mean; ra.stdv = a.stdv;
        ra.g_mp = a.w_mpjpe * inv_rows / J; ra.g_rp = a.w_reproj * inv_rows / (2 * J); ra.g_bn = a.w_bone * inv_rows / 16.0f;
        ra.reproj = a.intrinsics != nullptr; ra.bone = a.w_bone != 0.0f;
        ra.rx = ra.ry = ra.rz = ra.fx = ra.fy = ra.cx = ra.cy = 0.0f;
        if (a.root != nullptr) { ra.rx = a.root[(row0 + r) * 3]; ra.ry = a.root[(row0 + r) * 3 + 1]; ra.rz = a.root[(row0 + r) * 3 + 2]; }
        if (a.intrinsics != nullptr) {
            const float4 k = *reinterpret_cast<const float4*>(a.intrinsics + (row0 + r) * 4);
            ra.fx = k.x; ra.fy = k.y; ra.cx = k.z; ra.cy = k.w;
        }
        pose_row_terms(ra, s_pred + r * O, s_tgt + r * O, s_t2d + r * 2 * J, G, l_mp, l_rp, l_bn, s_metric);
    }
    __syncthreads();                                // every thread is done reading s_pred
    if (a.dpred != nullptr) {
        if (live) {
#pragma unroll
            for (int j = 0; j < PL_MAXJ; ++j) {
                if (j >= J) break;
#pragma unroll
                for (int d = 0; d < 3; ++d)         // chain rule through joints = v * std + mean
                    s_pred[r * O + 3 * j + d] = G[j][d] * (a.stdv != nullptr ? a.stdv[3 * j + d] : 1.0f);
            }
        }
        __syncthreads();
        for (int i = threadIdx.x; i < nrows * O; i += PL_ROWS) a.dpred[row0 * O + i] = s_pred[i];
    }
    l_mp = warp_sum(l_mp); l_rp = warp_sum(l_rp); l_bn = warp_sum(l_bn);
    if ((threadIdx.x & 31) == 0) {
        s_red[threadIdx.x >> 5] = l_mp;
        s_red[PL_ROWS / 32 + (threadIdx.x >> 5)] = l_rp;
        s_red[2 * (PL_ROWS / 32) + (threadIdx.x >> 5)] = l_bn;
    }
    __syncthreads();
    float* o = a.partials + static_cast<long>(blockIdx.x) * (4 + J);
    if (threadIdx.x < 3) {
        float t = 0.0f;
        for (int w = 0; w < PL_ROWS / 32; ++w) t += s_red[threadIdx.x * (PL_ROWS / 32) + w];
        o[threadIdx.x] = t;
    }
    if (threadIdx.x == 3) o[3] = 0.0f;
    if (threadIdx.x < J) o[4 + threadIdx.x] = s_metric[threadIdx.x];
}

// single block: fixed-order reduction of the per-block partials, then the weighted total
__global__ void pose_loss_finalize_kernel(const float* __restrict__ partials, int nblk, PoseLossArgs a) {
    griddep_launch();
    griddep_wait();
    __shared__ double sh[32][4 + PL_MAXJ];
    const int W = 4 + a.J;
    const int c = threadIdx.x % 32, lane_r = threadIdx.x / 32;
    if (c < W) {
        double s = 0.0;
        for (int b = lane_r; b < nblk; b += 32) s += static_cast<double>(partials[static_cast<long>(b) * W + c]);
        sh[lane_r][c] = s;
    }
    __syncthreads();
    __shared__ double tot[4 + PL_MAXJ];
    if (lane_r == 0 && c < W) {
        double t = 0.0;
        for (int r = 0; r < 32; ++r) t += sh[r][c];
        tot[c] = t;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        const double n = static_cast<double>(a.global_rows);
        const double mp = tot[0] / (n * a.J), rp = tot[1] / (n * a.J * 2), bn = tot[2] / (n * 16);
        a.out[0] = static_cast<float>(a.w_mpjpe * mp + a.w_reproj * rp + a.w_bone * bn);
        a.out[1] = static_cast<float>(mp);
        a.out[2] = static_cast<float>(rp);
        a.out[3] = static_cast<float>(bn);
    }
    if (threadIdx.x < a.J) a.out[4 + threadIdx.x] = static_cast<float>(tot[4 + threadIdx.x]);
}

}  // namespace

int flip_pose(const float* in, float* out, int64_t poses, int J, int D, int grad, cudaStream_t st) {
    LIFTER_REQUIRE(J == 17, "flip_pose: the left/right joint lists are those of the 17-joint H36M skeleton (got %d joints)", J);
    LIFTER_REQUIRE(D == 2 || D == 3, "flip_pose: last dimension must be 2 or 3 (got %d)", D);
    LIFTER_REQUIRE(in != out, "flip_pose: in-place flip is not supported");
    const long n = poses * J * D;
    if (n == 0) return LIFTER_OK;
    LIFTER_CHECK_CUDA(launch_pdl("flip_pose_kernel", flip_pose_kernel, dim3(static_cast<unsigned>((n + 255) / 256)), dim3(256), 0, st, in, out, n, D, grad));
    return LIFTER_OK;
}

int coco2h36m(const float* in, int in_channels, float* out, int64_t poses, float scale, const float* mean,
              const float* stdv, int mode, cudaStream_t st) {
    if (poses <= 0) return LIFTER_OK;
    LIFTER_REQUIRE(in != nullptr && out != nullptr && in != out, "coco2h36m: bad pointers");
    LIFTER_REQUIRE(in_channels >= 2 && in_channels <= 4, "coco2h36m: %d values per keypoint (x, y[, conf])", in_channels);
    LIFTER_REQUIRE(mode == -1 || mode == COLS_MINMAX_NORM || (mode == COLS_NORMALISE && mean && stdv),
                   "coco2h36m: mode %d needs -1 (none), 0 with mean/std, or 2", mode);
    const long n = poses * 34;
    LIFTER_CHECK_CUDA(launch_pdl("coco2h36m_kernel", coco2h36m_kernel, dim3(static_cast<unsigned>((n + 255) / 256)),
                                 dim3(256), 0, st, in, in_channels, out, static_cast<long>(poses), scale, mean, stdv, mode));
    return LIFTER_OK;
}

int affine_cols(const float* in, float* out, int64_t rows, int cols, const float* a, const float* b, int mode,
                cudaStream_t st) {
    LIFTER_REQUIRE(mode >= COLS_NORMALISE && mode <= COLS_MINMAX_DENORM, "affine_cols: unknown mode %d", mode);
    LIFTER_REQUIRE(mode == COLS_MINMAX_NORM || (a != nullptr && b != nullptr), "affine_cols: null statistics");
    const long n = rows * cols;
    if (n == 0) return LIFTER_OK;
    long blocks = (n + 255) / 256;
    const long cap = static_cast<long>(num_sms()) * 16;
    if (blocks > cap) blocks = cap;
    LIFTER_CHECK_CUDA(launch_pdl("affine_cols_kernel", affine_cols_kernel, dim3(static_cast<unsigned>(blocks)), dim3(256), 0, st, in, out, n, cols, a, b, mode));
    return LIFTER_OK;
}

int pose_loss_partial_rows(int64_t rows) { return static_cast<int>((rows + PL_ROWS - 1) / PL_ROWS); }

int pose_loss(const PoseLossArgs& a, cudaStream_t st) {
    LIFTER_REQUIRE(a.J >= 1 && a.J <= PL_MAXJ, "pose_loss: %d joints unsupported (1..%d)", a.J, PL_MAXJ);
    LIFTER_REQUIRE(a.rows >= 1 && a.global_rows >= a.rows, "pose_loss: empty batch");
    LIFTER_REQUIRE(a.pred && a.target && a.partials && a.out, "pose_loss: null pointer");
    LIFTER_REQUIRE((a.mean == nullptr) == (a.stdv == nullptr), "pose_loss: mean and std come together");
    LIFTER_REQUIRE(a.intrinsics == nullptr || a.target2d != nullptr, "pose_loss: reprojection needs the 2D target");
    LIFTER_REQUIRE(a.w_bone == 0.0f || a.J == 17, "pose_loss: the bone-length term is defined on the 17-joint H36M skeleton");
    static PerDeviceOnce once;
    bool& configured = *once.slot();
    const size_t smem = (static_cast<size_t>(PL_ROWS) * (3 * a.J + 3 * a.J + 2 * a.J) + PL_MAXJ + 3 * (PL_ROWS / 32) + 4) * sizeof(float);
    if (!configured) {
        LIFTER_CHECK_CUDA(cudaFuncSetAttribute(pose_loss_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536));
        configured = true;
    }
    const int nblk = pose_loss_partial_rows(a.rows);
    LIFTER_CHECK_CUDA(launch_pdl("pose_loss_kernel", pose_loss_kernel, dim3(nblk), dim3(PL_ROWS), smem, st, a));
    LIFTER_CHECK_CUDA(launch_pdl("pose_loss_finalize_kernel", pose_loss_finalize_kernel, dim3(1), dim3(1024), 0, st, a.partials, nblk, a));
    return LIFTER_OK;
}

}  // namespace lifter
