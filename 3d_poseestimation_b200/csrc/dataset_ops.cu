// Dataset-side keypoint processing: phase1_lifting/main.py:271-355 `process_data` as two
// HBM-bound passes over the [frames][joints][dims] fp32 array instead of Python loops over frames.
//   pass 1  keypoint_stats_kernel    root-centring on the fly, per (joint, coordinate) sum / sum of squares in fp64
//                                    around the first frame's value (shifted one-pass variance), min, max;
//                                    fixed-order block partials -> keypoint_stats_finalize_kernel (no float atomics)
//   pass 2  process_keypoints_kernel root-centring, (x - mean) / std or 2x - 1, root row zeroed or dropped
// Algorithmic traffic: pass 1 reads 4 B per element; pass 2 reads 4 B and writes 4 B per element.
#include "dataset_ops.cuh"

namespace lifter {

constexpr int DS_THREADS = 256;
constexpr int DS_MAXC = 64;          // joints * dims of one frame (17 x 2 = 34, 17 x 3 = 51)

namespace {

__device__ __forceinline__ double centred(const float* __restrict__ row, int c, int d, bool sub_root) {
    const double v = static_cast<double>(row[c]);
    return sub_root ? v - static_cast<double>(row[d]) : v;           // main.py:275-277, exact in fp64
}

// Block = G row groups x cols columns (G = 256 / cols); thread (g, c) walks rows r0 + g, r0 + g + G, ... of column c,
// so the block's loads of one iteration cover G consecutive frames = one contiguous span of the array.
__global__ void __launch_bounds__(DS_THREADS) keypoint_stats_kernel(const float* __restrict__ data, long frames, int cols,
                                                                    int D, int zc, long rows_per_block,
                                                                    double* __restrict__ partials) {
    griddep_launch();
    griddep_wait();
    __shared__ double s_red[4][DS_THREADS];
    const int G = DS_THREADS / cols;
    const int t = threadIdx.x, c = t % cols, g = t / cols, d = c % D;
    const bool sub_root = zc != 0 && c >= D;
    double sum = 0.0, sq = 0.0, mn = CUDART_INF, mx = -CUDART_INF;
    if (g < G) {
        const double shift = centred(data, c, d, sub_root);          // first frame: the same on every block
        const long r0 = static_cast<long>(blockIdx.x) * rows_per_block;
        const long r1 = min(frames, r0 + rows_per_block);
#pragma unroll 8
        for (long r = r0 + g; r < r1; r += G) {
            const double v = centred(data + r * cols, c, d, sub_root);
            const double x = v - shift;
            sum += x;
            sq = fma(x, x, sq);
            mn = fmin(mn, v);
            mx = fmax(mx, v);
        }
    }
    s_red[0][t] = sum; s_red[1][t] = sq; s_red[2][t] = mn; s_red[3][t] = mx;
    __syncthreads();
    if (t < cols) {
        for (int k = 1; k < G; ++k) {                                 // fixed order
            const int o = k * cols + t;
            sum += s_red[0][o]; sq += s_red[1][o];
            mn = fmin(mn, s_red[2][o]); mx = fmax(mx, s_red[3][o]);
        }
        double* p = partials + static_cast<long>(blockIdx.x) * 4 * cols;
        p[t] = sum; p[cols + t] = sq; p[2 * cols + t] = mn; p[3 * cols + t] = mx;
    }
}

// out [4][cols] fp64: mean, population std (main.py:280-289: divide by n_frames), min, max
__global__ void __launch_bounds__(DS_MAXC) keypoint_stats_finalize_kernel(const float* __restrict__ data, long frames,
                                                                          int cols, int D, int zc,
                                                                          const double* __restrict__ partials, int nblk,
                                                                          double* __restrict__ out) {
    griddep_launch();
    griddep_wait();
    const int c = threadIdx.x;
    if (c >= cols) return;
    double sum = 0.0, sq = 0.0, mn = CUDART_INF, mx = -CUDART_INF;
    for (int b = 0; b < nblk; ++b) {
        const double* p = partials + static_cast<long>(b) * 4 * cols;
        sum += p[c]; sq += p[cols + c];
        mn = fmin(mn, p[2 * cols + c]); mx = fmax(mx, p[3 * cols + c]);
    }
    const double shift = centred(data, c, c % D, zc != 0 && c >= D);
    const double n = static_cast<double>(frames);
    const double m = sum / n;
    const double var = fmax(0.0, (sq - sum * m) / n);
    out[c] = shift + m;
    out[cols + c] = sqrt(var);
    out[2 * cols + c] = mn;
    out[3 * cols + c] = mx;
}

// out [frames][Jout][D]; Jout = J - (root_mode == ROOT_DROP)
__global__ void __launch_bounds__(DS_THREADS) process_keypoints_kernel(const float* __restrict__ in, float* __restrict__ out,
                                                                       long frames, int J, int D, int zc,
                                                                       const double* __restrict__ stats, int mode,
                                                                       int root_mode) {
    griddep_launch();
    griddep_wait();
    __shared__ double s_mean[DS_MAXC], s_rstd[DS_MAXC];
    const int cols = J * D;
    if (mode == COLS_NORMALISE) {
        for (int c = threadIdx.x; c < cols; c += blockDim.x) {
            s_mean[c] = stats[c];
            s_rstd[c] = 1.0 / stats[cols + c];
        }
        __syncthreads();
    }
    const int skip = root_mode == ROOT_DROP ? D : 0, ocols = cols - skip;
    const long n = frames * ocols, stride = static_cast<long>(gridDim.x) * blockDim.x;
    // (row, column) of this thread's element advance by a fixed step: no 64-bit division in the loop
    const unsigned first = blockIdx.x * blockDim.x + threadIdx.x;                     // < 2^31 (grid <= 8 * SMs)
    const unsigned step = static_cast<unsigned>(stride);
    long r = first / static_cast<unsigned>(ocols);
    int oc = static_cast<int>(first % static_cast<unsigned>(ocols));
    const int step_r = static_cast<int>(step / static_cast<unsigned>(ocols)), step_c = static_cast<int>(step % static_cast<unsigned>(ocols));
    for (long i = first; i < n; i += stride, r += step_r, oc += step_c) {
        if (oc >= ocols) { oc -= ocols; ++r; }
        const int c = oc + skip, d = c % D;
        double v = centred(in + r * cols, c, d, zc != 0 && c >= D);
        if (mode == COLS_NORMALISE) v = (v - s_mean[c]) * s_rstd[c];               // main.py:334,337
        else if (mode == COLS_MINMAX_NORM) v = 2.0 * v - 1.0;                       // main.py:331
        if (root_mode == ROOT_ZERO && c < D) v = 0.0;                               // main.py:343-344
        out[i] = static_cast<float>(v);
    }
}

}  // namespace

int keypoint_stats_blocks(int64_t frames, int cols) {
    if (frames <= 0 || cols <= 0 || cols > DS_MAXC) return 0;
    const long G = DS_THREADS / cols;
    long blocks = (frames + G * 8 - 1) / (G * 8);                    // at least eight rows per thread
    const long cap = static_cast<long>(num_sms()) * 8;
    if (blocks > cap) blocks = cap;
    return static_cast<int>(blocks < 1 ? 1 : blocks);
}

int keypoint_stats(const float* data, int64_t frames, int J, int D, int zc, double* scratch, double* out,
                   cudaStream_t st) {
    const int cols = J * D;
    LIFTER_REQUIRE(frames >= 1, "keypoint_stats: empty data set");
    LIFTER_REQUIRE(J >= 1 && D >= 1 && cols <= DS_MAXC, "keypoint_stats: %d joints x %d dims unsupported (<= %d values per frame)", J, D, DS_MAXC);
    LIFTER_REQUIRE(data && scratch && out, "keypoint_stats: null pointer");
    const int blocks = keypoint_stats_blocks(frames, cols);
    const long G = DS_THREADS / cols;
    long rpb = (frames + blocks - 1) / blocks;
    rpb = (rpb + G - 1) / G * G;
    LIFTER_CHECK_CUDA(launch_pdl("keypoint_stats_kernel", keypoint_stats_kernel, dim3(blocks), dim3(DS_THREADS), 0, st, data,
                                 static_cast<long>(frames), cols, D, zc, rpb, scratch));
    LIFTER_CHECK_CUDA(launch_pdl("keypoint_stats_finalize_kernel", keypoint_stats_finalize_kernel, dim3(1), dim3(DS_MAXC), 0,
                                 st, data, static_cast<long>(frames), cols, D, zc, static_cast<const double*>(scratch), blocks, out));
    return LIFTER_OK;
}

int process_keypoints(const float* in, float* out, int64_t frames, int J, int D, int zc, const double* stats, int mode,
                      int root_mode, cudaStream_t st) {
    const int cols = J * D;
    LIFTER_REQUIRE(J >= 1 && D >= 1 && cols <= DS_MAXC, "process_keypoints: %d joints x %d dims unsupported", J, D);
    LIFTER_REQUIRE(mode == -1 || mode == COLS_MINMAX_NORM || (mode == COLS_NORMALISE && stats != nullptr),
                   "process_keypoints: mode %d needs -1 (none), 0 with statistics, or 2", mode);
    LIFTER_REQUIRE(root_mode >= ROOT_KEEP && root_mode <= ROOT_DROP, "process_keypoints: root mode %d", root_mode);
    LIFTER_REQUIRE(root_mode != ROOT_DROP || J >= 2, "process_keypoints: nothing left after dropping the root");
    if (frames <= 0) return LIFTER_OK;
    LIFTER_REQUIRE(in && out && in != out, "process_keypoints: bad pointers");
    const long n = frames * (cols - (root_mode == ROOT_DROP ? D : 0));
    long blocks = (n + DS_THREADS * 4 - 1) / (DS_THREADS * 4);
    const long cap = static_cast<long>(num_sms()) * 8;
    if (blocks > cap) blocks = cap;
    LIFTER_CHECK_CUDA(launch_pdl("process_keypoints_kernel", process_keypoints_kernel, dim3(static_cast<unsigned>(blocks)),
                                 dim3(DS_THREADS), 0, st, in, out, static_cast<long>(frames), J, D, zc, stats, mode, root_mode));
    return LIFTER_OK;
}

}  // namespace lifter

// ---- C ABI (include/lifter_b200.h) ---------------------------------------------------------------------------------
#include "../../include/lifter_b200.h"

extern "C" {

int64_t lifter_keypoint_stats_scratch_bytes(int64_t frames, int joints, int dims) {
    if (frames < 1 || joints < 1 || dims < 1 || joints * dims > lifter::DS_MAXC) return LIFTER_ERR_INVALID;
    return static_cast<int64_t>(lifter::keypoint_stats_blocks(frames, joints * dims)) * 4 * joints * dims *
           static_cast<int64_t>(sizeof(double));
}

int lifter_keypoint_stats(const float* data, int64_t frames, int joints, int dims, int zero_centre, void* scratch,
                          double* out, void* stream) {
    return lifter::keypoint_stats(data, frames, joints, dims, zero_centre, static_cast<double*>(scratch), out,
                                  static_cast<cudaStream_t>(stream));
}

int lifter_process_keypoints(const float* in, float* out, int64_t frames, int joints, int dims, int zero_centre,
                             const double* stats, int mode, int root_mode, void* stream) {
    return lifter::process_keypoints(in, out, frames, joints, dims, zero_centre, stats, mode, root_mode,
                                     static_cast<cudaStream_t>(stream));
}

}  // extern "C"
