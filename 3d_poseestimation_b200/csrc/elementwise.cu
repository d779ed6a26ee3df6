#include "elementwise.cuh"

namespace lifter {

namespace {

__global__ void convert_pad_kernel(const float* __restrict__ src, int rows, int cols, uint16_t* __restrict__ dst,
                                   int rows_p, int cols_p, int fmt) {
    griddep_launch();
    griddep_wait();
    const int groups = cols_p >> 3;
    const long idx = static_cast<long>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (idx >= static_cast<long>(rows_p) * groups) return;
    const int r = static_cast<int>(idx / groups), c0 = static_cast<int>(idx % groups) * 8;
    uint32_t w[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) {
        float a = 0.0f, b = 0.0f;
        if (r < rows) {
            const int c = c0 + 2 * e;
            if (c < cols) a = src[static_cast<long>(r) * cols + c];
            if (c + 1 < cols) b = src[static_cast<long>(r) * cols + c + 1];
        }
        w[e] = pack16(a, b, fmt);
    }
    *reinterpret_cast<uint4*>(dst + static_cast<long>(r) * cols_p + c0) = make_uint4(w[0], w[1], w[2], w[3]);
}

__global__ void fold_bn_kernel(const float* __restrict__ bias, const float* __restrict__ gamma,
                               const float* __restrict__ beta, const float* __restrict__ rm,
                               const float* __restrict__ rv, int H, int use_bn, float eps, float* __restrict__ scale,
                               float* __restrict__ shift) {
    griddep_launch();
    griddep_wait();
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= H) return;
    if (use_bn) {
        const float s = gamma[j] / sqrtf(rv[j] + eps);
        scale[j] = s;
        shift[j] = (bias[j] - rm[j]) * s + beta[j];
    } else {
        scale[j] = 1.0f;
        shift[j] = bias[j];
    }
}

__global__ void prep_input_kernel(const float* __restrict__ x, long rows, int I, int Ip,
                                  const float* __restrict__ mean, const float* __restrict__ stdv,
                                  uint16_t* __restrict__ dst, int fmt, unsigned* __restrict__ zero_words, int n_zero) {
    griddep_launch();
    griddep_wait();
    if (blockIdx.x == 0 && static_cast<int>(threadIdx.x) < n_zero) zero_words[threadIdx.x] = 0u;   // the pass's barrier words
    const int groups = Ip >> 3;
    const long idx = static_cast<long>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (idx >= rows * groups) return;
    const long r = idx / groups;
    const int c0 = static_cast<int>(idx % groups) * 8;
    float v[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) {
        const int c = c0 + e;
        float t = 0.0f;
        if (c < I) {
            t = x[r * I + c];
            if (mean != nullptr) t = (t - mean[c]) / stdv[c];
        }
        v[e] = t;
    }
    *reinterpret_cast<uint4*>(dst + r * Ip + c0) =
        make_uint4(pack16(v[0], v[1], fmt), pack16(v[2], v[3], fmt), pack16(v[4], v[5], fmt), pack16(v[6], v[7], fmt));
}

__global__ void make_out_affine_kernel(const float* __restrict__ bias, const float* __restrict__ mean,
                                       const float* __restrict__ stdv, int O, int Op, float* __restrict__ scale,
                                       float* __restrict__ shift) {
    griddep_launch();
    griddep_wait();
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= Op) return;
    float s = 0.0f, t = 0.0f;
    if (j < O) {
        s = stdv != nullptr ? stdv[j] : 1.0f;
        t = bias[j] * s + (mean != nullptr ? mean[j] : 0.0f);
    }
    scale[j] = s;
    shift[j] = t;
}

}  // namespace

int convert_pad(const float* src, int rows, int cols, void* dst, int rows_p, int cols_p, int fmt, cudaStream_t st) {
    LIFTER_REQUIRE(cols_p % 8 == 0 && rows_p >= rows && cols_p >= cols, "convert_pad: bad padding");
    const long n = static_cast<long>(rows_p) * (cols_p / 8);
    LIFTER_CHECK_CUDA(launch_pdl("convert_pad_kernel", convert_pad_kernel, dim3(static_cast<unsigned>((n + 255) / 256)), dim3(256), 0, st, src, rows, cols,
                                                                             static_cast<uint16_t*>(dst), rows_p, cols_p, fmt));
    return LIFTER_OK;
}

int fold_bn(const float* bias, const float* gamma, const float* beta, const float* rm, const float* rv, int H,
            int use_bn, float eps, float* scale, float* shift, cudaStream_t st) {
    LIFTER_CHECK_CUDA(launch_pdl("fold_bn_kernel", fold_bn_kernel, dim3(ceil_div(H, 256)), dim3(256), 0, st, bias, gamma, beta, rm, rv, H, use_bn, eps, scale, shift));
    return LIFTER_OK;
}

int prep_input(const float* x, int64_t rows, int I, int Ip, const float* mean, const float* stdv, void* dst, int fmt,
               cudaStream_t st, unsigned* zero_words, int n_zero) {
    LIFTER_REQUIRE(Ip % 8 == 0 && Ip >= I, "prep_input: bad padding");
    LIFTER_REQUIRE((mean == nullptr) == (stdv == nullptr), "prep_input: mean and std come together");
    const long n = rows * (Ip / 8);
    LIFTER_CHECK_CUDA(launch_pdl("prep_input_kernel", prep_input_kernel, dim3(static_cast<unsigned>((n + 255) / 256)), dim3(256), 0, st, x, rows, I, Ip, mean, stdv,
                                                                            static_cast<uint16_t*>(dst), fmt, zero_words, zero_words != nullptr ? n_zero : 0));
    return LIFTER_OK;
}

int make_out_affine(const float* bias, const float* mean, const float* stdv, int O, int Op, float* scale, float* shift,
                    cudaStream_t st) {
    LIFTER_CHECK_CUDA(launch_pdl("make_out_affine_kernel", make_out_affine_kernel, dim3(1), dim3(64), 0, st, bias, mean, stdv, O, Op, scale, shift));
    return LIFTER_OK;
}

}  // namespace lifter
