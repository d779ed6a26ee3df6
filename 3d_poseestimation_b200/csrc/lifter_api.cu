// C ABI of the lifter hot path (include/lifter_b200.h): layout queries, weight packing, inference.
// Training entry points live in lifter_train.cu.
#include "../../include/lifter_b200.h"
#include "elementwise.cuh"
#include "gemm_sm100.cuh"
#include "layout.cuh"

using namespace lifter;

namespace lifter {

// Rows per inference chunk.  Measured on B200 (scripts/gemm_sweep.py): a GEMM launch carries ~9 us of fixed cost
// (launch, prologue, first TMA round trip, last epilogue), so short L2-resident chunks (9,472 rows = 21 us per
// layer) lose more to that than they gain from L2 residency; at >= 37,888 rows per launch the H x H GEMM runs at
// >= 1.2 PFLOP/s and its activation traffic (2 x rows x H x 2 B per layer) stays 2.5x under the HBM roof.
// The chunk is therefore as large as a 1 GB workspace allows, in whole waves of 128-row blocks.
int64_t default_eval_chunk(const Layout& lay, int64_t batch) {
    const int n_blocks = ceil_div(lay.H, 256);
    int g = num_sms(), b = n_blocks;
    while (b) { int t = g % b; g = b; b = t; }
    const int unit = num_sms() / g;                                   // m-blocks per whole wave pattern
    const int64_t target_rows = (1ll << 30) / (6ll * lay.H);          // 3 buffers x 2 bytes x H per row
    int64_t m_blocks = (target_rows / 128) / unit * unit;
    if (m_blocks < unit) m_blocks = unit;
    const int64_t chunk = m_blocks * 128;
    return batch < chunk ? batch : chunk;
}

static int64_t eval_ws_bytes(const Layout& lay, int64_t chunk) {
    return static_cast<int64_t>(align_up(static_cast<size_t>(chunk) * lay.Ip * 2, 1024)) +
           3 * static_cast<int64_t>(align_up(static_cast<size_t>(chunk) * lay.H * 2, 1024)) + 1024;
}

}  // namespace lifter

extern "C" {

int lifter_version(void) { return 2; }

// digest of the sources this binary was compiled from (build.py passes it in; the marker lets the loader find it in the
// file without dlopen, so a stale binary is never mapped)
#ifndef LIFTER_SOURCE_DIGEST
#define LIFTER_SOURCE_DIGEST "unknown"
#endif
extern const char lifter_source_digest_marker[] = "LIFTER_SRC_DIGEST=" LIFTER_SOURCE_DIGEST;
const char* lifter_source_digest(void) { return lifter_source_digest_marker + 18; }

long long lifter_launch_count(void) { return launch_count(); }

const char* lifter_last_error(void) { return get_error(); }

int lifter_device_check(void) {
    int dev = 0;
    cudaDeviceProp prop;
    if (cudaGetDevice(&dev) != cudaSuccess || cudaGetDeviceProperties(&prop, dev) != cudaSuccess) {
        set_error("no CUDA device: the lifter kernels are sm_100a only and there is no CPU fallback");
        return LIFTER_ERR_NODEVICE;
    }
    if (prop.major != 10) {
        set_error("device %s is sm_%d%d: the lifter kernels are built for sm_100a only", prop.name, prop.major, prop.minor);
        return LIFTER_ERR_NODEVICE;
    }
    return LIFTER_OK;
}

int lifter_num_param_tensors(const lifter_config_t* cfg) {
    Layout lay;
    if (make_layout(cfg, &lay) != 0) return LIFTER_ERR_INVALID;
    return static_cast<int>(lay.p_off.size());
}
int64_t lifter_param_offset(const lifter_config_t* cfg, int t) {
    Layout lay;
    if (make_layout(cfg, &lay) != 0) return LIFTER_ERR_INVALID;
    if (t < 0 || t >= static_cast<int>(lay.p_off.size())) { set_error("tensor index %d out of range", t); return LIFTER_ERR_INVALID; }
    return lay.p_off[t];
}
int64_t lifter_param_numel(const lifter_config_t* cfg, int t) {
    Layout lay;
    if (make_layout(cfg, &lay) != 0) return LIFTER_ERR_INVALID;
    if (t < 0 || t >= static_cast<int>(lay.p_numel.size())) { set_error("tensor index %d out of range", t); return LIFTER_ERR_INVALID; }
    return lay.p_numel[t];
}
int64_t lifter_param_total(const lifter_config_t* cfg) {
    Layout lay;
    if (make_layout(cfg, &lay) != 0) return LIFTER_ERR_INVALID;
    return lay.p_total;
}
int64_t lifter_operand_bytes(const lifter_config_t* cfg) {
    Layout lay;
    if (make_layout(cfg, &lay) != 0) return LIFTER_ERR_INVALID;
    return lay.w_total * 2;
}
int64_t lifter_eval_workspace_bytes(const lifter_config_t* cfg, int64_t batch) {
    Layout lay;
    if (make_layout(cfg, &lay) != 0) return LIFTER_ERR_INVALID;
    if (batch < 1) { set_error("batch must be >= 1"); return LIFTER_ERR_INVALID; }
    return eval_ws_bytes(lay, default_eval_chunk(lay, batch));
}

int lifter_pack_weights(const lifter_config_t* cfg, const float* params, void* operands, int fmt, void* stream) {
    Layout lay;
    LIFTER_PROPAGATE(make_layout(cfg, &lay));
    LIFTER_REQUIRE(params && operands, "pack_weights: null pointer");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    uint16_t* w = static_cast<uint16_t*>(operands);
    for (int l = 0; l < lay.L; ++l)
        LIFTER_PROPAGATE(convert_pad(params + lay.p_off[lay.t_weight(l)], lay.H, lay.in_dim(l), w + lay.w_off[l], lay.H,
                                     lay.in_pad(l), fmt, st));
    LIFTER_PROPAGATE(convert_pad(params + lay.p_off[lay.t_w2()], lay.O, lay.H, w + lay.w_off[lay.L], lay.Op, lay.H, fmt, st));
    return LIFTER_OK;
}

int lifter_fold_bn(const lifter_config_t* cfg, const float* params, const float* running_mean,
                   const float* running_var, float* scale_shift, void* stream) {
    Layout lay;
    LIFTER_PROPAGATE(make_layout(cfg, &lay));
    LIFTER_REQUIRE(params && scale_shift && running_mean && running_var, "fold_bn: null pointer");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    for (int l = 0; l < lay.L; ++l)
        LIFTER_PROPAGATE(fold_bn(params + lay.p_off[lay.t_bias(l)], params + lay.p_off[lay.t_gamma(l)],
                                 params + lay.p_off[lay.t_beta(l)], running_mean + static_cast<int64_t>(l) * lay.H,
                                 running_var + static_cast<int64_t>(l) * lay.H, lay.H, cfg->use_bn, cfg->bn_eps,
                                 scale_shift + static_cast<int64_t>(2 * l) * lay.H,
                                 scale_shift + static_cast<int64_t>(2 * l + 1) * lay.H, st));
    return LIFTER_OK;
}

int lifter_forward_eval(const lifter_config_t* cfg, const void* operands, const float* scale_shift,
                        const float* w2_bias, const float* x, float* out, int64_t batch, const float* in_mean,
                        const float* in_std, const float* out_mean, const float* out_std, void* workspace,
                        int64_t workspace_bytes, int fmt, int64_t chunk_rows, void* stream) {
    Layout lay;
    LIFTER_PROPAGATE(make_layout(cfg, &lay));
    LIFTER_REQUIRE(operands && scale_shift && w2_bias && x && out && workspace, "forward_eval: null pointer");
    LIFTER_REQUIRE(batch >= 1 && batch < (1ll << 31), "forward_eval: batch %lld out of range", static_cast<long long>(batch));
    LIFTER_REQUIRE(fmt == FMT_F16 || fmt == FMT_BF16, "forward_eval: unknown operand format %d", fmt);
    LIFTER_REQUIRE((out_mean == nullptr) == (out_std == nullptr), "forward_eval: out_mean and out_std come together");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const int64_t chunk = chunk_rows > 0 ? (chunk_rows < batch ? chunk_rows : batch) : default_eval_chunk(lay, batch);
    LIFTER_REQUIRE(workspace_bytes >= eval_ws_bytes(lay, chunk), "forward_eval: workspace too small (%lld < %lld bytes)",
                   static_cast<long long>(workspace_bytes), static_cast<long long>(eval_ws_bytes(lay, chunk)));
    LIFTER_REQUIRE((reinterpret_cast<uintptr_t>(workspace) & 255) == 0, "forward_eval: workspace must be 256-byte aligned");

    uint8_t* ws = static_cast<uint8_t*>(workspace);
    float* out_scale = reinterpret_cast<float*>(ws);
    float* out_shift = out_scale + 64;
    ws += 1024;
    uint16_t* a0 = reinterpret_cast<uint16_t*>(ws);
    ws += align_up(static_cast<size_t>(chunk) * lay.Ip * 2, 1024);
    uint16_t* act[3];
    for (int i = 0; i < 3; ++i) {
        act[i] = reinterpret_cast<uint16_t*>(ws);
        ws += align_up(static_cast<size_t>(chunk) * lay.H * 2, 1024);
    }
    const uint16_t* w = static_cast<const uint16_t*>(operands);
    const int H = lay.H;

    LIFTER_PROPAGATE(make_out_affine(w2_bias, out_mean, out_std, lay.O, lay.Op, out_scale, out_shift, st));

    for (int64_t r0 = 0; r0 < batch; r0 += chunk) {
        const int rows = static_cast<int>(batch - r0 < chunk ? batch - r0 : chunk);
        LIFTER_PROPAGATE(prep_input(x + r0 * lay.I, rows, lay.I, lay.Ip, in_mean, in_std, a0, fmt, st));
        int cur = 0;
        {   // input stage, baselineModel.py:90-94 (dropout is the identity in eval mode)
            GemmArgs g;
            g.A = a0; g.lda = lay.Ip; g.B = w + lay.w_off[0]; g.ldb = lay.Ip;
            g.M = rows; g.N = H; g.K = lay.Ip; g.fmt = fmt; g.out_mode = OUT_16_TMA; g.D = act[cur]; g.ldd = H;
            g.relu = 1; g.scale = scale_shift; g.shift = scale_shift + H;
            LIFTER_PROPAGATE(gemm_launch(g, st));
        }
        for (int s = 0; s < lay.S; ++s) {   // residual block, baselineModel.py:32-47
            const int l1 = 2 * s + 1, l2 = 2 * s + 2;
            const int tmp = 2, nxt = cur ^ 1;
            GemmArgs g;
            g.A = act[cur]; g.lda = H; g.B = w + lay.w_off[l1]; g.ldb = H;
            g.M = rows; g.N = H; g.K = H; g.fmt = fmt; g.out_mode = OUT_16_TMA; g.D = act[tmp]; g.ldd = H;
            g.relu = 1; g.scale = scale_shift + static_cast<int64_t>(2 * l1) * H; g.shift = g.scale + H;
            LIFTER_PROPAGATE(gemm_launch(g, st));
            g.A = act[tmp]; g.B = w + lay.w_off[l2]; g.D = act[nxt];
            g.scale = scale_shift + static_cast<int64_t>(2 * l2) * H; g.shift = g.scale + H;
            g.residual = act[cur]; g.ldr = H;
            LIFTER_PROPAGATE(gemm_launch(g, st));
            cur = nxt;
        }
        {   // output layer, baselineModel.py:100
            GemmArgs g;
            g.A = act[cur]; g.lda = H; g.B = w + lay.w_off[lay.L]; g.ldb = H;
            g.M = rows; g.N = lay.Op; g.K = H; g.fmt = fmt; g.out_mode = OUT_32_DIRECT; g.D = out + r0 * lay.O;
            g.n_valid = lay.O; g.scale = out_scale; g.shift = out_shift;
            LIFTER_PROPAGATE(gemm_launch(g, st));
        }
    }
    return LIFTER_OK;
}

int lifter_gemm(const lifter_gemm_desc_t* d, void* stream) {
    LIFTER_REQUIRE(d != nullptr, "gemm: null descriptor");
    GemmArgs g;
    g.A = d->A; g.lda = d->lda; g.B = d->B; g.ldb = d->ldb;
    g.M = d->M; g.N = d->N; g.K = d->K; g.a_mn_major = d->a_mn_major; g.b_mn_major = d->b_mn_major; g.fmt = d->fmt;
    g.out_mode = d->out_mode; g.D = d->D; g.ldd = d->ldd; g.n_valid = d->n_valid; g.splits = d->splits < 1 ? 1 : d->splits;
    g.relu = d->relu; g.scale = d->scale; g.shift = d->shift; g.residual = d->residual; g.ldr = d->ldr;
    g.stat_sum = d->stat_sum; g.stat_sq = d->stat_sq;
    g.cluster = d->cluster;                                        // test hook: 0 auto, 1 single-CTA, 2 pairs
    g.timeline = static_cast<unsigned long long*>(d->timeline);
    g.max_stages = d->max_stages;
    return gemm_launch(g, static_cast<cudaStream_t>(stream));
}
int lifter_trace_begin(void* stream) { return trace_begin(static_cast<cudaStream_t>(stream)); }
int lifter_trace_end(char* buf, int len) { return trace_end(buf, len); }
int lifter_gemm_slab_rows(int M) { return gemm_slab_rows(M); }
int lifter_gemm_stat_rows(int M) { return gemm_stat_rows(M); }
int lifter_gemm_effective_splits(int K, int splits) { return gemm_effective_splits(K, splits); }

}  // extern "C"
