// Dataset-side keypoint processing: phase1_lifting/main.py:271-355 `process_data` as two
// HBM-bound passes over the [frames][joints][dims] fp32 array instead of Python loops over frames.
//   pass 1  keypoint_stats_kernel    root-centring on the fly, per (joint, coordinate) sum / sum of squares around
//                                    the first frame's value (shifted one-pass variance; fp32 groups of eight folded
//                                    into fp64 accumulators), min, max;
//                                    fixed-order block partials -> keypoint_stats_finalize_kernel (no float atomics)
//   pass 2  process_keypoints_kernel root-centring, (x - mean) / std or 2x - 1, root row zeroed or dropped
// Algorithmic traffic: pass 1 reads 4 B per element; pass 2 reads 4 B and writes 4 B per element.
#include "dataset_ops.cuh"

namespace lifter {

constexpr int DS_THREADS = 256;
constexpr int DS_MAXC = 64;          // joints * dims of one frame (17 x 2 = 34, 17 x 3 = 51)

namespace {

// Thread mapping of both passes: block = G row groups x cols columns (G = 256 / cols); thread (g, c) walks rows
// r0 + g, r0 + g + G, ... of ITS column c, so everything that depends on the column (root coordinate, shift, mean,
// 1 / std) sits in registers, and one iteration of the block covers G consecutive frames = one contiguous span.
// The root coordinate is loaded unconditionally (it is in the same 128-byte lines) and selected, not branched on.

__global__ void __launch_bounds__(DS_THREADS, 2) keypoint_stats_kernel(const float* __restrict__ data, long frames, int cols,
                                                                    int D, int zc, long rows_per_block,
                                                                    double* __restrict__ partials) {
    griddep_launch();
    griddep_wait();
    __shared__ double s_sum[2][DS_THREADS];
    __shared__ float s_mm[2][DS_THREADS];
    const int G = DS_THREADS / cols;
    const int t = threadIdx.x, c = t % cols, g = t / cols, d = c % D;
    const float w = (zc != 0 && c >= D) ? 1.0f : 0.0f;               // main.py:275-277: joints 1.. minus the root
    double sum = 0.0, sq = 0.0;
    float mn = CUDART_INF_F, mx = -CUDART_INF_F;
    if (g < G) {
        // Non-tensor fp64 runs at a few operations per clock per SM on this part (measured: an all-fp64 version of this
        // loop reached 0.76 TB/s), so the stream is fp32: x = (v - root) - shift rounds twice at the magnitude of the
        // CENTRED value (below the input's own quantisation), eight of them are summed in fp32 and each such group
        // is folded into the fp64 accumulators with one add.
        const float shift = data[c] - w * data[d];                    // first frame: any constant near the mean works
        const long r0 = static_cast<long>(blockIdx.x) * rows_per_block;
        const long r1 = min(frames, r0 + rows_per_block);
        const long n_it = r1 > r0 + g ? (r1 - r0 - g + G - 1) / G : 0;
        const float* pv = data + (r0 + g) * cols + c;
        const float* pr = data + (r0 + g) * cols + d;
        const long step = static_cast<long>(G) * cols;
        long it = 0;
        for (; it + 8 <= n_it; it += 8) {
            float v[8], rt[8];
#pragma unroll
            for (int u = 0; u < 8; ++u) { v[u] = pv[(it + u) * step]; rt[u] = pr[(it + u) * step]; }
            float fs = 0.0f, fq = 0.0f;
#pragma unroll
            for (int u = 0; u < 8; ++u) {
                const float vf = v[u] - w * rt[u];
                mn = fminf(mn, vf);
                mx = fmaxf(mx, vf);
                const float x = vf - shift;
                fs += x;
                fq = fmaf(x, x, fq);
            }
            sum += static_cast<double>(fs);
            sq += static_cast<double>(fq);
        }
        float fs = 0.0f, fq = 0.0f;
        for (; it < n_it; ++it) {
            const float vf = pv[it * step] - w * pr[it * step];
            mn = fminf(mn, vf);
            mx = fmaxf(mx, vf);
            const float x = vf - shift;
            fs += x;
            fq = fmaf(x, x, fq);
        }
        sum += static_cast<double>(fs);
        sq += static_cast<double>(fq);
    }
    s_sum[0][t] = sum; s_sum[1][t] = sq; s_mm[0][t] = mn; s_mm[1][t] = mx;
    __syncthreads();
    if (t < cols) {
        for (int k = 1; k < G; ++k) {                                 // fixed order
            const int o = k * cols + t;
            sum += s_sum[0][o]; sq += s_sum[1][o];
            mn = fminf(mn, s_mm[0][o]); mx = fmaxf(mx, s_mm[1][o]);
        }
        double* p = partials + static_cast<long>(blockIdx.x) * 4 * cols;
        p[t] = sum; p[cols + t] = sq; p[2 * cols + t] = static_cast<double>(mn); p[3 * cols + t] = static_cast<double>(mx);
    }
}

// Same pass with the frames staged through shared memory by the bulk-copy engine: one elected thread keeps DS_STAGES
// spans of DS_TR frames (contiguous, 224 * cols bytes each) in flight per block with cp.async.bulk + mbarrier, so the
// bytes in flight no longer depend on registers or occupancy; the 256 threads consume a span from shared memory with the
// same (row group, column) mapping.  Frames past the last full span are read with plain loads by the last block.
constexpr int DS_TR = 56;
constexpr int DS_STAGES = 4;

__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}

__global__ void __launch_bounds__(DS_THREADS, 4) keypoint_stats_tma_kernel(const float* __restrict__ data, long frames, int cols,
                                                                           int D, int zc, long tiles_per_block,
                                                                           double* __restrict__ partials) {
    griddep_launch();
    extern __shared__ __align__(128) unsigned char ds_smem[];
    __shared__ double s_sum[2][DS_THREADS];
    __shared__ float s_mm[2][DS_THREADS];
    __shared__ __align__(8) unsigned long long s_bar[DS_STAGES];
    const int tile_elems = DS_TR * cols;
    const uint32_t tile_bytes = static_cast<uint32_t>(tile_elems) * 4u;
    float* tiles = reinterpret_cast<float*>(ds_smem);
    const int G = DS_THREADS / cols;
    const int t = threadIdx.x, c = t % cols, g = t / cols, d = c % D;
    const float w = (zc != 0 && c >= D) ? 1.0f : 0.0f;
    const long full_tiles = frames / DS_TR;
    const long tile0 = static_cast<long>(blockIdx.x) * tiles_per_block;
    const long tile1 = min(full_tiles, tile0 + tiles_per_block);
    const int n_tiles = tile1 > tile0 ? static_cast<int>(tile1 - tile0) : 0;
    if (t == 0) {
        for (int s = 0; s < DS_STAGES; ++s) mbar_init(smem_u32(&s_bar[s]), 1);
        fence_mbar_init();
    }
    __syncthreads();
    griddep_wait();
    if (t == 0) {
        for (int s = 0; s < DS_STAGES && s < n_tiles; ++s) {
            mbar_arrive_expect_tx(smem_u32(&s_bar[s]), tile_bytes);
            bulk_g2s(smem_u32(tiles + s * tile_elems), data + (tile0 + s) * tile_elems, tile_bytes, smem_u32(&s_bar[s]));
        }
    }
    double sum = 0.0, sq = 0.0;
    float mn = CUDART_INF_F, mx = -CUDART_INF_F;
    const float shift = g < G ? data[c] - w * data[d] : 0.0f;        // first frame: any constant near the mean works
    for (int i = 0; i < n_tiles; ++i) {
        const int slot = i % DS_STAGES;
        mbar_wait(smem_u32(&s_bar[slot]), (i / DS_STAGES) & 1);
        if (g < G) {
            const float* tp = tiles + slot * tile_elems;
            float fs = 0.0f, fq = 0.0f;
#pragma unroll 4
            for (int r = g; r < DS_TR; r += G) {                      // <= 56 terms in fp32, then one fp64 add per span
                const float vf = tp[r * cols + c] - w * tp[r * cols + d];
                mn = fminf(mn, vf);
                mx = fmaxf(mx, vf);
                const float x = vf - shift;
                fs += x;
                fq = fmaf(x, x, fq);
            }
            sum += static_cast<double>(fs);
            sq += static_cast<double>(fq);
        }
        __syncthreads();                                              // every thread is done with this slot
        if (t == 0 && i + DS_STAGES < n_tiles) {
            mbar_arrive_expect_tx(smem_u32(&s_bar[slot]), tile_bytes);
            bulk_g2s(smem_u32(tiles + slot * tile_elems), data + (tile0 + i + DS_STAGES) * tile_elems, tile_bytes, smem_u32(&s_bar[slot]));
        }
    }
    if (blockIdx.x == gridDim.x - 1 && g < G) {                       // frames past the last full span
        float fs = 0.0f, fq = 0.0f;
        for (long r = full_tiles * DS_TR + g; r < frames; r += G) {
            const float vf = data[r * cols + c] - w * data[r * cols + d];
            mn = fminf(mn, vf);
            mx = fmaxf(mx, vf);
            const float x = vf - shift;
            fs += x;
            fq = fmaf(x, x, fq);
        }
        sum += static_cast<double>(fs);
        sq += static_cast<double>(fq);
    }
    s_sum[0][t] = sum; s_sum[1][t] = sq; s_mm[0][t] = mn; s_mm[1][t] = mx;
    __syncthreads();
    if (t < cols) {
        for (int k = 1; k < G; ++k) {                                 // fixed order
            const int o = k * cols + t;
            sum += s_sum[0][o]; sq += s_sum[1][o];
            mn = fminf(mn, s_mm[0][o]); mx = fmaxf(mx, s_mm[1][o]);
        }
        double* p = partials + static_cast<long>(blockIdx.x) * 4 * cols;
        p[t] = sum; p[cols + t] = sq; p[2 * cols + t] = static_cast<double>(mn); p[3 * cols + t] = static_cast<double>(mx);
    }
}

// out [4][cols] fp64: mean, population std (main.py:280-289: divide by n_frames), min, max.  16 groups of 64 lanes
// each fold every 16th partial row, then lane c of group 0 folds the 16 group results: a fixed order again.
constexpr int FIN_GROUPS = 16;
__global__ void __launch_bounds__(FIN_GROUPS * DS_MAXC) keypoint_stats_finalize_kernel(
    const float* __restrict__ data, long frames, int cols, int D, int zc, const double* __restrict__ partials, int nblk,
    double* __restrict__ out) {
    griddep_launch();
    griddep_wait();
    __shared__ double s_part[4][FIN_GROUPS][DS_MAXC];
    const int c = threadIdx.x % DS_MAXC, grp = threadIdx.x / DS_MAXC;
    double sum = 0.0, sq = 0.0, mn = CUDART_INF, mx = -CUDART_INF;
    if (c < cols) {
#pragma unroll 4
        for (int b = grp; b < nblk; b += FIN_GROUPS) {
            const double* p = partials + static_cast<long>(b) * 4 * cols;
            sum += p[c]; sq += p[cols + c];
            mn = fmin(mn, p[2 * cols + c]); mx = fmax(mx, p[3 * cols + c]);
        }
    }
    s_part[0][grp][c] = sum; s_part[1][grp][c] = sq; s_part[2][grp][c] = mn; s_part[3][grp][c] = mx;
    __syncthreads();
    if (grp != 0 || c >= cols) return;
    for (int k = 1; k < FIN_GROUPS; ++k) {
        sum += s_part[0][k][c]; sq += s_part[1][k][c];
        mn = fmin(mn, s_part[2][k][c]); mx = fmax(mx, s_part[3][k][c]);
    }
    const float w = (zc != 0 && c >= D) ? 1.0f : 0.0f;
    const double shift = static_cast<double>(data[c] - w * data[c % D]);          // the kernel's fp32 constant
    const double n = static_cast<double>(frames);
    const double m = sum / n;
    const double var = fmax(0.0, (sq - sum * m) / n);
    out[c] = shift + m;
    out[cols + c] = sqrt(var);
    out[2 * cols + c] = mn;
    out[3 * cols + c] = mx;
}

// out [frames][Jout][D]; Jout = J - (root_mode == ROOT_DROP).  Same thread mapping over the OUTPUT columns.
__global__ void __launch_bounds__(DS_THREADS, 2) process_keypoints_kernel(const float* __restrict__ in, float* __restrict__ out,
                                                                       long frames, int J, int D, int zc,
                                                                       const double* __restrict__ stats, int mode,
                                                                       int root_mode, long rows_per_block) {
    griddep_launch();
    griddep_wait();
    const int cols = J * D, skip = root_mode == ROOT_DROP ? D : 0, ocols = cols - skip;
    const int G = DS_THREADS / ocols;
    const int t = threadIdx.x, oc = t % ocols, g = t / ocols, c = oc + skip, d = c % D;
    if (g >= G) return;
    const float w = (zc != 0 && c >= D) ? 1.0f : 0.0f;
    // out = ((v - root) - a) * b in fp32, a = mean carried as a two-float (hi + lo) so its fp32 rounding does not show
    float a_hi = 0.0f, a_lo = 0.0f, b = 1.0f;
    if (mode == COLS_NORMALISE) {                                                   // main.py:334,337
        const double mean = stats[c];
        a_hi = static_cast<float>(mean);
        a_lo = static_cast<float>(mean - static_cast<double>(a_hi));
        b = static_cast<float>(1.0 / stats[cols + c]);
    } else if (mode == COLS_MINMAX_NORM) { a_hi = 0.5f; b = 2.0f; }                 // main.py:331: 2v - 1
    if (root_mode == ROOT_ZERO && c < D) { a_hi = 0.0f; a_lo = 0.0f; b = 0.0f; }    // main.py:343-344
    const long r0 = static_cast<long>(blockIdx.x) * rows_per_block;
    const long r1 = min(frames, r0 + rows_per_block);
    const long n_it = r1 > r0 + g ? (r1 - r0 - g + G - 1) / G : 0;
    const float* pv = in + (r0 + g) * cols + c;
    const float* pr = in + (r0 + g) * cols + d;
    float* po = out + (r0 + g) * ocols + oc;
    const long istep = static_cast<long>(G) * cols, ostep = static_cast<long>(G) * ocols;
#pragma unroll 8
    for (long it = 0; it < n_it; ++it) {
        const float v = pv[it * istep], rt = w * pr[it * istep];
        po[it * ostep] = (((v - rt) - a_hi) - a_lo) * b;
    }
}

// The apply pass with the same bulk-copy staging of the input frames (outputs go out as plain coalesced stores).
__global__ void __launch_bounds__(DS_THREADS, 4) process_keypoints_tma_kernel(const float* __restrict__ in, float* __restrict__ out,
                                                                              long frames, int J, int D, int zc,
                                                                              const double* __restrict__ stats, int mode,
                                                                              int root_mode, long tiles_per_block) {
    griddep_launch();
    extern __shared__ __align__(128) unsigned char ds_smem[];
    __shared__ __align__(8) unsigned long long s_bar[DS_STAGES];
    const int cols = J * D, skip = root_mode == ROOT_DROP ? D : 0, ocols = cols - skip;
    const int tile_elems = DS_TR * cols;
    const uint32_t tile_bytes = static_cast<uint32_t>(tile_elems) * 4u;
    float* tiles = reinterpret_cast<float*>(ds_smem);
    const int G = DS_THREADS / ocols;
    const int t = threadIdx.x, oc = t % ocols, g = t / ocols, c = oc + skip, d = c % D;
    const long full_tiles = frames / DS_TR;
    const long tile0 = static_cast<long>(blockIdx.x) * tiles_per_block;
    const long tile1 = min(full_tiles, tile0 + tiles_per_block);
    const int n_tiles = tile1 > tile0 ? static_cast<int>(tile1 - tile0) : 0;
    if (t == 0) {
        for (int s = 0; s < DS_STAGES; ++s) mbar_init(smem_u32(&s_bar[s]), 1);
        fence_mbar_init();
    }
    __syncthreads();
    griddep_wait();
    if (t == 0) {
        for (int s = 0; s < DS_STAGES && s < n_tiles; ++s) {
            mbar_arrive_expect_tx(smem_u32(&s_bar[s]), tile_bytes);
            bulk_g2s(smem_u32(tiles + s * tile_elems), in + (tile0 + s) * tile_elems, tile_bytes, smem_u32(&s_bar[s]));
        }
    }
    const float w = (zc != 0 && c >= D) ? 1.0f : 0.0f;
    float a_hi = 0.0f, a_lo = 0.0f, b = 1.0f;                                       // out = ((v - root) - a) * b
    if (mode == COLS_NORMALISE) {                                                   // main.py:334,337
        const double mean = stats[c];
        a_hi = static_cast<float>(mean);
        a_lo = static_cast<float>(mean - static_cast<double>(a_hi));
        b = static_cast<float>(1.0 / stats[cols + c]);
    } else if (mode == COLS_MINMAX_NORM) { a_hi = 0.5f; b = 2.0f; }                 // main.py:331: 2v - 1
    if (root_mode == ROOT_ZERO && c < D) { a_hi = 0.0f; a_lo = 0.0f; b = 0.0f; }    // main.py:343-344
    for (int i = 0; i < n_tiles; ++i) {
        const int slot = i % DS_STAGES;
        mbar_wait(smem_u32(&s_bar[slot]), (i / DS_STAGES) & 1);
        if (g < G) {
            const float* tp = tiles + slot * tile_elems;
            float* po = out + (tile0 + i) * DS_TR * ocols + oc;
#pragma unroll 4
            for (int r = g; r < DS_TR; r += G)
                po[static_cast<long>(r) * ocols] = (((tp[r * cols + c] - w * tp[r * cols + d]) - a_hi) - a_lo) * b;
        }
        __syncthreads();
        if (t == 0 && i + DS_STAGES < n_tiles) {
            mbar_arrive_expect_tx(smem_u32(&s_bar[slot]), tile_bytes);
            bulk_g2s(smem_u32(tiles + slot * tile_elems), in + (tile0 + i + DS_STAGES) * tile_elems, tile_bytes, smem_u32(&s_bar[slot]));
        }
    }
    if (blockIdx.x == gridDim.x - 1 && g < G) {                                     // frames past the last full span
        for (long r = full_tiles * DS_TR + g; r < frames; r += G)
            out[r * ocols + oc] = (((in[r * cols + c] - w * in[r * cols + d]) - a_hi) - a_lo) * b;
    }
}

}  // namespace

int keypoint_stats_blocks(int64_t frames, int cols) {
    if (frames <= 0 || cols <= 0 || cols > DS_MAXC) return 0;
    long blocks = frames / DS_TR;                                    // at least one staged span per block
    const long cap = static_cast<long>(num_sms()) * 4;               // four resident blocks per SM
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
    if (reinterpret_cast<uintptr_t>(data) % 16 == 0) {               // bulk copies need a 16-byte aligned source
        const size_t smem = static_cast<size_t>(DS_STAGES) * DS_TR * cols * sizeof(float);
        static PerDeviceOnce once;
    bool& configured = *once.slot();
        if (!configured) {
            LIFTER_CHECK_CUDA(cudaFuncSetAttribute(keypoint_stats_tma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                                   DS_STAGES * DS_TR * DS_MAXC * static_cast<int>(sizeof(float))));
            configured = true;
        }
        const long full_tiles = frames / DS_TR;
        const long tpb = (full_tiles + blocks - 1) / blocks;
        LIFTER_CHECK_CUDA(launch_pdl("keypoint_stats_tma_kernel", keypoint_stats_tma_kernel, dim3(blocks), dim3(DS_THREADS), smem, st,
                                     data, static_cast<long>(frames), cols, D, zc, tpb, scratch));
    } else {
        const long G = DS_THREADS / cols;
        long rpb = (frames + blocks - 1) / blocks;
        rpb = (rpb + G - 1) / G * G;
        LIFTER_CHECK_CUDA(launch_pdl("keypoint_stats_kernel", keypoint_stats_kernel, dim3(blocks), dim3(DS_THREADS), 0, st, data,
                                     static_cast<long>(frames), cols, D, zc, rpb, scratch));
    }
    LIFTER_CHECK_CUDA(launch_pdl("keypoint_stats_finalize_kernel", keypoint_stats_finalize_kernel, dim3(1), dim3(FIN_GROUPS * DS_MAXC), 0,
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
    const int ocols = cols - (root_mode == ROOT_DROP ? D : 0);
    if (reinterpret_cast<uintptr_t>(in) % 16 == 0) {                 // bulk copies need a 16-byte aligned source
        const size_t smem = static_cast<size_t>(DS_STAGES) * DS_TR * cols * sizeof(float);
        static PerDeviceOnce once;
    bool& configured = *once.slot();
        if (!configured) {
            LIFTER_CHECK_CUDA(cudaFuncSetAttribute(process_keypoints_tma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                                   DS_STAGES * DS_TR * DS_MAXC * static_cast<int>(sizeof(float))));
            configured = true;
        }
        const int blocks = keypoint_stats_blocks(frames, cols);
        const long full_tiles = frames / DS_TR;
        const long tpb = (full_tiles + blocks - 1) / blocks;
        LIFTER_CHECK_CUDA(launch_pdl("process_keypoints_tma_kernel", process_keypoints_tma_kernel, dim3(blocks), dim3(DS_THREADS), smem, st,
                                     in, out, static_cast<long>(frames), J, D, zc, stats, mode, root_mode, tpb));
        return LIFTER_OK;
    }
    const long G = DS_THREADS / ocols;
    long blocks = (frames + G * 8 - 1) / (G * 8);
    const long cap = static_cast<long>(num_sms()) * 8;
    if (blocks > cap) blocks = cap;
    long rpb = (frames + blocks - 1) / blocks;
    rpb = (rpb + G - 1) / G * G;
    LIFTER_CHECK_CUDA(launch_pdl("process_keypoints_kernel", process_keypoints_kernel, dim3(static_cast<unsigned>(blocks)),
                                 dim3(DS_THREADS), 0, st, in, out, static_cast<long>(frames), J, D, zc, stats, mode, root_mode, rpb));
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
