#include "heatmap_ops.cuh"

#include <math.h>

namespace lifter {

namespace {

constexpr int HM_THREADS = 256;
constexpr int HM_VEC_PER_VOLUME = HM_GRID * HM_GRID * HM_GRID / 4;        // float4 stores per key point: 65,536

__device__ __forceinline__ void st_stream_f4(float4* p, const float4& v) {
    asm volatile("st.global.cs.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}

// The volume is written exactly once, 16 bytes per thread and instruction, w fastest (a warp instruction covers 512
// contiguous bytes); the 27 (size^3) non-zero voxels are found by an integer range test per float4, so the kernel is a
// fill with a few integer operations per store.
template <typename T>
__global__ void __launch_bounds__(HM_THREADS) heatmap_kernel(const T* __restrict__ kp, long n, double inv_2s2, int half,
                                                             float* __restrict__ out, int* __restrict__ status) {
    griddep_launch();
    griddep_wait();
    for (long unit = blockIdx.x; unit < n; unit += gridDim.x) {
        const double x = static_cast<double>(kp[unit * 3]), y = static_cast<double>(kp[unit * 3 + 1]),
                     z = static_cast<double>(kp[unit * 3 + 2]);
        if (threadIdx.x == 0 && status != nullptr &&
            (fmin(x, fmin(y, z)) < -1.0 || fmax(x, fmax(y, z)) > 1.0))            // H36_dataset.py:156-157 asserts
            atomicOr(status, 1);
        // xyz_to_uvw (H36_dataset.py:143-144), then the scaling of :163 -- in the key points' own precision, as numpy
        // does it (float32 key points give a float32 k) -- and np.rint (half to even) of :164
        double k0, k1, k2;
        if (sizeof(T) == 4) {
            const float xf = static_cast<float>(kp[unit * 3]), yf = static_cast<float>(kp[unit * 3 + 1]),
                        zf = static_cast<float>(kp[unit * 3 + 2]);
            k0 = static_cast<double>(__fmul_rn(31.5f, __fadd_rn(1.0f, -yf)));
            k1 = static_cast<double>(__fmul_rn(31.5f, __fadd_rn(1.0f, -zf)));
            k2 = static_cast<double>(__fmul_rn(31.5f, __fadd_rn(1.0f, xf)));
        } else {
            k0 = __dmul_rn(31.5, __dadd_rn(1.0, -y)); k1 = __dmul_rn(31.5, __dadd_rn(1.0, -z)); k2 = __dmul_rn(31.5, __dadd_rn(1.0, x));
        }
        const int c0 = static_cast<int>(rint(k0)), c1 = static_cast<int>(rint(k1)), c2 = static_cast<int>(rint(k2));
        float4* vol = reinterpret_cast<float4*>(out) + unit * HM_VEC_PER_VOLUME;
#pragma unroll 4
        for (int i = threadIdx.x; i < HM_VEC_PER_VOLUME; i += HM_THREADS) {
            const int u = i >> 10, v = (i >> 4) & 63, w0 = (i & 15) * 4;
            float4 val = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
            if (abs(u - c0) <= half && abs(v - c1) <= half && w0 + 3 >= c2 - half && w0 <= c2 + half) {
                const double duv = (u - k0) * (u - k0) + (v - k1) * (v - k1);
                float e[4];
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    const int w = w0 + q;
                    const double dw = w - k2;
                    e[q] = abs(w - c2) <= half ? static_cast<float>(exp(-(duv + dw * dw) * inv_2s2)) : 0.0f;
                }
                val = make_float4(e[0], e[1], e[2], e[3]);
            }
            st_stream_f4(vol + i, val);
        }
    }
}

}  // namespace

int keypoint_heatmaps(const void* kp, int kp_is_f64, int64_t n, double sigma, float* out, int* status, cudaStream_t st) {
    if (n <= 0) return LIFTER_OK;
    LIFTER_REQUIRE(kp != nullptr && out != nullptr, "keypoint_heatmaps: null pointer");
    LIFTER_REQUIRE(sigma > 0.0 && sigma <= 5.0, "keypoint_heatmaps: sigma %g outside (0, 5]", sigma);
    LIFTER_REQUIRE((reinterpret_cast<uintptr_t>(out) & 15) == 0, "keypoint_heatmaps: the output must be 16-byte aligned");
    int size = static_cast<int>(ceil(6.0 * sigma));                       // H36_dataset.py:167-170
    if (size % 2 == 0) size += 1;
    const int half = size / 2;
    const double inv_2s2 = 1.0 / (2.0 * (sigma * sigma));
    long blocks = n;
    const long cap = static_cast<long>(num_sms()) * 8;
    if (blocks > cap) blocks = cap;
    if (kp_is_f64)
        LIFTER_CHECK_CUDA(launch_pdl("heatmap_kernel<double>", heatmap_kernel<double>, dim3(static_cast<unsigned>(blocks)), dim3(HM_THREADS), 0, st,
                                     static_cast<const double*>(kp), static_cast<long>(n), inv_2s2, half, out, status));
    else
        LIFTER_CHECK_CUDA(launch_pdl("heatmap_kernel<float>", heatmap_kernel<float>, dim3(static_cast<unsigned>(blocks)), dim3(HM_THREADS), 0, st,
                                     static_cast<const float*>(kp), static_cast<long>(n), inv_2s2, half, out, status));
    return LIFTER_OK;
}

}  // namespace lifter
