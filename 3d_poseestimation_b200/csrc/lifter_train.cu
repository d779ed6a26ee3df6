// Training path of the lifter behind the C ABI: forward with saved activations, backward, fused loss, Adam.
//
// Reference being replaced: the autograd graph torch builds for LinearModel.forward in train mode
// (baselineModel.py:87-102 with the blocks of :32-47) and loss.backward() / optimizer.step() at
// train_1.py:94-96.  Per hidden layer l (0 = input stage, 2i+1 / 2i+2 = the two layers of block i):
//
//   forward   z_l = a_{l-1} W_l^T                      tcgen05 GEMM, epilogue stores z (16-bit) + column sums
//             mu, var over the (global) batch           reduce + finalize (SyncBN: all-reduce of the sums between)
//             a_l = drop(relu(gamma (z-mu) rstd + beta)) [+ block input]     bn_apply_fwd, also stores the active bits
//   backward  dy = g * active / (1-p);  dgamma = sum dy zhat;  dbeta = sum dy          bwd_stats (+ reduce)
//             dz = gamma rstd (dy - mean dy - zhat mean(dy zhat))                     bwd_dz
//             dW_l = dz^T a_{l-1}   (split-K, MN-major operands)                      GEMM + wgrad_reduce
//             g_{l-1} = dz W_l [+ gradient of the residual stream]                    GEMM (residual in the epilogue)
#include "../../include/lifter_b200.h"
#include "elementwise.cuh"
#include "gemm_sm100.cuh"
#include "heatmap_ops.cuh"
#include "layout.cuh"
#include "loss_adam.cuh"
#include "phase5_ops.cuh"
#include "pose_ops.cuh"
#include "train_kernels.cuh"

using namespace lifter;

namespace {

struct Carver {
    uint8_t* base; size_t off;
    explicit Carver(void* b) : base(static_cast<uint8_t*>(b)), off(0) {}
    template <typename T> T* take(size_t bytes) {
        T* p = base ? reinterpret_cast<T*>(base + off) : nullptr;
        off += align_up(bytes, 256);
        return p;
    }
};

struct SavedWs {                        // forward workspace kept for backward
    uint16_t* a0;
    std::vector<uint16_t*> z, a;
    std::vector<uint8_t*> bits;
    std::vector<float*> mean_z, rstd, scale, shift;
    std::vector<double*> sums;
    std::vector<size_t> sums_off;
    float *psum, *psq, *out_scale, *out_shift;
    unsigned* bars;                     // grid-barrier words of the forward pass (zeroed by its first kernel)
    size_t bytes;
};

void carve_saved(const Layout& lay, int64_t B, void* base, SavedWs* w) {
    Carver c(base);
    const size_t act = static_cast<size_t>(B) * lay.H * 2;
    w->a0 = c.take<uint16_t>(static_cast<size_t>(B) * lay.Ip * 2);
    w->z.resize(lay.L); w->a.resize(lay.L); w->bits.resize(lay.L);
    w->mean_z.resize(lay.L); w->rstd.resize(lay.L); w->scale.resize(lay.L); w->shift.resize(lay.L);
    w->sums.resize(lay.L); w->sums_off.resize(lay.L);
    for (int l = 0; l < lay.L; ++l) {
        w->z[l] = c.take<uint16_t>(act);
        w->a[l] = c.take<uint16_t>(act);
        w->bits[l] = c.take<uint8_t>(static_cast<size_t>(B) * lay.H / 8);
        w->mean_z[l] = c.take<float>(lay.H * 4);
        w->rstd[l] = c.take<float>(lay.H * 4);
        w->scale[l] = c.take<float>(lay.H * 4);
        w->shift[l] = c.take<float>(lay.H * 4);
        w->sums_off[l] = c.off;
        w->sums[l] = c.take<double>(2 * lay.H * 8);
    }
    const size_t P = gemm_stat_rows(static_cast<int>(B));
    w->psum = c.take<float>(P * lay.H * 4);
    w->psq = c.take<float>(P * lay.H * 4);
    w->out_scale = c.take<float>(64 * 4);
    w->out_shift = c.take<float>(64 * 4);
    w->bars = c.take<unsigned>(GRID_BARRIER_WORDS * 4);
    w->bytes = c.off;
}

// symmetric SyncBN exchange buffer: 2 parities x nsites x world slots x 2 x H values of 16 bytes (LL words)
size_t dp_data_bytes(const Layout& lay, int world) {
    return align_up(static_cast<size_t>(2) * (2 * lay.L) * world * 2 * lay.H * 16, 256);
}
size_t dp_flag_bytes(const Layout&, int) { return 0; }
int fill_exchange(const Layout& lay, const lifter_train_ctx_t* t, int site, SyncExchange* ex) {
    *ex = SyncExchange();
    if (t->dp == nullptr || t->dp->world <= 1) return LIFTER_OK;
    // timing probe only (wrong numerics: local statistics): how much of a data-parallel step is the exchange
    static const bool skip = [] { const char* e = getenv("LIFTER_DP_DEBUG_NO_EXCHANGE"); return e != nullptr && e[0] == '1'; }();
    if (skip) return LIFTER_OK;
    const lifter_dp_t* d = t->dp;
    LIFTER_REQUIRE(d->world <= DP_MAX_WORLD && d->rank >= 0 && d->rank < d->world, "dp: world %d / rank %d out of range",
                   d->world, d->rank);
    LIFTER_REQUIRE(t->step_state != nullptr, "dp: the in-kernel BatchNorm exchange needs the device step state");
    const int calls = t->calls_per_step > 0 ? t->calls_per_step : 1;
    ex->world = d->world; ex->rank = d->rank;
    ex->site = t->call_index * 2 * lay.L + site; ex->nsites = 2 * lay.L * calls;
    for (int p = 0; p < d->world; ++p) {
        LIFTER_REQUIRE(d->sync[p] != nullptr, "dp: null peer buffer %d", p);
        ex->data[p] = d->sync[p];
    }
    ex->epoch = reinterpret_cast<const unsigned*>(t->step_state) + 1;      // lifter_step_state_t.adam_step
    ex->spin_limit = dp_spin_limit_clocks();
    // LIFTER_DP_EXCHANGE_WARP=1: the warp-cooperative polling of bn_device.cuh (measured slower, see there)
    static const bool warp = [] { const char* e = getenv("LIFTER_DP_EXCHANGE_WARP"); return e != nullptr && e[0] == '1'; }();
    ex->serial = warp ? 0 : 1;
    return LIFTER_OK;
}

// LIFTER_FUSE_BN_FWD=1 / LIFTER_FUSE_BN_BWD=1: the BatchNorm passes of the single-call forward / backward run as ONE
// grid-barrier launch per layer instead of separate launches (reduce + finalize, apply; stats, reduce, dz).  Off by
// default: measured on B200 (profiles/README.md, round 2) the barrier kernels are SLOWER inside the step -- a grid that
// must be fully resident cannot start while the side stream's persistent weight-gradient GEMM holds the SMs.
bool fuse_bn_launches(bool forward) {
    static const bool fwd = [] { const char* e = getenv("LIFTER_FUSE_BN_FWD"); return e != nullptr && e[0] == '1'; }();
    static const bool bwd = [] { const char* e = getenv("LIFTER_FUSE_BN_BWD"); return e != nullptr && e[0] == '1'; }();
    return forward ? fwd : bwd;
}

// NVSwitch multicast in the data-parallel optimiser (multimem.ld_reduce / multimem.st) when the caller provides the
// multicast mappings.  Default: for more than two ranks -- with ONE peer it saves no traffic and adds a switch hop
// (measured on 2 B200: 0.563 vs 0.537 ms per step).  LIFTER_DP_MULTICAST=0 / 1 forces it off / on.
bool dp_multicast_enabled(int world) {
    static const int forced = [] { const char* e = getenv("LIFTER_DP_MULTICAST"); return e != nullptr ? (e[0] == '0' ? 0 : 1) : -1; }();
    return forced >= 0 ? forced == 1 : world > 2;
}

int wgrad_splits(int M, int N, int K) {
    const int bn = N <= 64 ? 64 : (N <= 128 ? 128 : 256);
    const int tiles = ceil_div(M, 128) * ceil_div(N, bn);
    int s = num_sms() / tiles;
    if (s < 1) s = 1;
    if (s > 16) s = 16;
    return gemm_effective_splits(K, s);
}

struct ScratchWs {                      // backward workspace
    uint16_t *dO, *G[2], *Gh, *dz[2];   // dz alternates by layer: the weight-gradient GEMM of layer l reads dz[l & 1]
                                        // on the side stream while layer l-1 writes the other one
    float *pA, *pB, *pO, *coef, *slabs;    // pO: output-layer bias partials (read on the side stream)
    std::vector<double*> bsums;
    std::vector<size_t> bsums_off;
    unsigned* bars;                     // grid-barrier words of the backward pass (zeroed by its first kernel)
    size_t bytes;
};

void carve_scratch(const Layout& lay, int64_t B, void* base, ScratchWs* w) {
    Carver c(base);
    const size_t act = static_cast<size_t>(B) * lay.H * 2;
    w->dO = c.take<uint16_t>(static_cast<size_t>(B) * lay.Op * 2);
    w->G[0] = c.take<uint16_t>(act);
    w->G[1] = c.take<uint16_t>(act);
    w->Gh = c.take<uint16_t>(act);
    w->dz[0] = c.take<uint16_t>(act);
    w->dz[1] = c.take<uint16_t>(act);
    size_t R = ew_partial_rows(B, lay.H);
    const size_t Rf = bn_bwd_fused_partial_rows(B, lay.H);
    if (Rf > R) R = Rf;
    const size_t R64 = ew_partial_rows(B, 64);
    size_t pbytes = R * lay.H * 4;
    if (R64 * 64 * 4 > pbytes) pbytes = R64 * 64 * 4;
    w->pA = c.take<float>(pbytes);
    w->pB = c.take<float>(pbytes);
    w->pO = c.take<float>(R64 * 64 * 4);
    w->coef = c.take<float>(3 * lay.H * 4);
    w->bsums.resize(lay.L); w->bsums_off.resize(lay.L);
    for (int l = 0; l < lay.L; ++l) {
        w->bsums_off[l] = c.off;
        w->bsums[l] = c.take<double>(2 * lay.H * 8);
    }
    // split-K slabs: the largest of the wgrad GEMMs
    const int Bk = static_cast<int>(B);
    size_t slab = static_cast<size_t>(wgrad_splits(lay.H, lay.Ip, Bk)) * gemm_slab_rows(lay.H) * lay.Ip * 4;
    const size_t hh = static_cast<size_t>(wgrad_splits(lay.H, lay.H, Bk)) * gemm_slab_rows(lay.H) * lay.H * 4;
    const size_t oh = static_cast<size_t>(wgrad_splits(lay.Op, lay.H, Bk)) * gemm_slab_rows(lay.Op) * lay.H * 4;
    if (lay.S > 0 && hh > slab) slab = hh;
    if (oh > slab) slab = oh;
    w->slabs = c.take<float>(slab);
    w->bars = c.take<unsigned>(GRID_BARRIER_WORDS * 4);
    w->bytes = c.off;
}

int check_ctx(const lifter_config_t* cfg, const lifter_train_ctx_t* t, Layout* lay, bool backward) {
    LIFTER_PROPAGATE(make_layout(cfg, lay));
    LIFTER_REQUIRE(t != nullptr, "train: null context");
    LIFTER_REQUIRE(t->batch >= 1 && t->batch < (1ll << 31), "train: batch %lld out of range", static_cast<long long>(t->batch));
    LIFTER_REQUIRE(t->global_batch >= t->batch, "train: global_batch < batch");
    LIFTER_REQUIRE(t->fmt == FMT_F16 || t->fmt == FMT_BF16, "train: unknown operand format %d", t->fmt);
    LIFTER_REQUIRE(t->calls_per_step >= 0 && t->call_index >= 0 && t->call_index < (t->calls_per_step > 0 ? t->calls_per_step : 1),
                   "train: call_index %d outside calls_per_step %d", t->call_index, t->calls_per_step);
    LIFTER_REQUIRE(t->params && t->operands && t->saved, "train: null pointer");
    LIFTER_REQUIRE((reinterpret_cast<uintptr_t>(t->saved) & 255) == 0, "train: workspace must be 256-byte aligned");
    SavedWs s;
    carve_saved(*lay, t->batch, nullptr, &s);
    LIFTER_REQUIRE(t->saved_bytes >= static_cast<int64_t>(s.bytes), "train: saved workspace too small (%lld < %lld)",
                   static_cast<long long>(t->saved_bytes), static_cast<long long>(s.bytes));
    if (backward) {
        LIFTER_REQUIRE(t->grad_out && t->grad_params && t->scratch, "backward: null pointer");
        LIFTER_REQUIRE((reinterpret_cast<uintptr_t>(t->scratch) & 255) == 0, "backward: scratch must be 256-byte aligned");
        ScratchWs b;
        carve_scratch(*lay, t->batch, nullptr, &b);
        LIFTER_REQUIRE(t->scratch_bytes >= static_cast<int64_t>(b.bytes), "backward: scratch too small (%lld < %lld)",
                       static_cast<long long>(t->scratch_bytes), static_cast<long long>(b.bytes));
    } else {
        LIFTER_REQUIRE(t->x && t->out, "forward: null pointer");
        if (cfg->use_bn && t->bn_mode == 0 && t->global_batch == 1) {
            set_error("Expected more than 1 value per channel when training, got input size torch.Size([1, %d])", lay->H);
            return LIFTER_ERR_BATCH1;
        }
        LIFTER_REQUIRE(!(cfg->use_bn && t->bn_mode == 1) || (t->running_mean && t->running_var),
                       "forward: running statistics required in eval mode");
    }
    return LIFTER_OK;
}

}  // namespace

extern "C" {

int64_t lifter_train_saved_bytes(const lifter_config_t* cfg, int64_t batch) {
    Layout lay;
    if (make_layout(cfg, &lay) != 0) return LIFTER_ERR_INVALID;
    if (batch < 1) { set_error("batch must be >= 1"); return LIFTER_ERR_INVALID; }
    SavedWs s;
    carve_saved(lay, batch, nullptr, &s);
    return static_cast<int64_t>(s.bytes);
}
int64_t lifter_train_scratch_bytes(const lifter_config_t* cfg, int64_t batch) {
    Layout lay;
    if (make_layout(cfg, &lay) != 0) return LIFTER_ERR_INVALID;
    if (batch < 1) { set_error("batch must be >= 1"); return LIFTER_ERR_INVALID; }
    ScratchWs s;
    carve_scratch(lay, batch, nullptr, &s);
    return static_cast<int64_t>(s.bytes);
}
int64_t lifter_dp_sync_bytes(const lifter_config_t* cfg, int world) {
    Layout lay;
    if (make_layout(cfg, &lay) != 0) return LIFTER_ERR_INVALID;
    if (world < 1 || world > DP_MAX_WORLD) { set_error("dp: world must be 1..%d", DP_MAX_WORLD); return LIFTER_ERR_INVALID; }
    return static_cast<int64_t>(dp_data_bytes(lay, world) + dp_flag_bytes(lay, world));
}
int64_t lifter_train_fwd_sums_offset(const lifter_config_t* cfg, int64_t batch, int layer) {
    Layout lay;
    if (make_layout(cfg, &lay) != 0) return LIFTER_ERR_INVALID;
    if (layer < 0 || layer >= lay.L || batch < 1) { set_error("layer/batch out of range"); return LIFTER_ERR_INVALID; }
    SavedWs s;
    carve_saved(lay, batch, nullptr, &s);
    return static_cast<int64_t>(s.sums_off[layer]);
}
int64_t lifter_train_bwd_sums_offset(const lifter_config_t* cfg, int64_t batch, int layer) {
    Layout lay;
    if (make_layout(cfg, &lay) != 0) return LIFTER_ERR_INVALID;
    if (layer < 0 || layer >= lay.L || batch < 1) { set_error("layer/batch out of range"); return LIFTER_ERR_INVALID; }
    ScratchWs s;
    carve_scratch(lay, batch, nullptr, &s);
    return static_cast<int64_t>(s.bsums_off[layer]);
}

// ---------------------------------------------------------------------------------------------- forward
// fused = single-GPU path: the column-sum reduction is merged into the finalize kernel (no all-reduce between)
// LIFTER_FUSE_LAYER=1: the single-call forward runs every hidden layer that fits as ONE launch of the whole-layer kernel
// (GEMM + column statistics + grid barriers + BatchNorm / ReLU / dropout / residual from TMEM; gemm_sm100.cuh).  OFF by
// default: measured on B200 (profiles/r02_layer_timeline_*.txt) the second epilogue pass cannot overlap any MMA and costs
// 5-10 us per layer, so the layer's period (34-40 us) only ties the three launches it replaces (37 us) and the training
// step is 0.516 ms against 0.484 ms.  Read on every call (a getenv), so tests switch it inside one process.
static bool fuse_layer_enabled() {
    const char* e = getenv("LIFTER_FUSE_LAYER");
    return e != nullptr && e[0] == '1';
}
// single-call forward, batch-statistics BatchNorm, shape fits: the whole layer is ONE launch
static bool fwd_whole_layer(const lifter_config_t* cfg, const lifter_train_ctx_t* t, const Layout& lay, int l, bool fused) {
    if (!(fused && cfg->use_bn && t->bn_mode == 0 && fuse_layer_enabled() && 2 * l + 1 < GRID_BARRIER_WORDS)) return false;
    const int drop = t->apply_dropout && cfg->p_dropout > 0.0f;
    return gemm_fused_bn_fits(static_cast<int>(t->batch), lay.H, cfg->p_dropout, drop, t->row_offset);
}

static void fill_finalize(const lifter_config_t* cfg, const lifter_train_ctx_t* t, const Layout& lay, const SavedWs& w, int l,
                          BnFwdFinalize* f) {
    const int H = lay.H;
    f->sums = w.sums[l]; f->count = static_cast<double>(t->global_batch);
    f->bias = t->params + lay.p_off[lay.t_bias(l)];
    f->gamma = t->params + lay.p_off[lay.t_gamma(l)];
    f->beta = t->params + lay.p_off[lay.t_beta(l)];
    f->running_mean = t->running_mean ? t->running_mean + static_cast<int64_t>(l) * H : nullptr;
    f->running_var = t->running_var ? t->running_var + static_cast<int64_t>(l) * H : nullptr;
    f->num_batches_tracked = t->num_batches_tracked ? reinterpret_cast<long long*>(t->num_batches_tracked) + l : nullptr;
    f->momentum = cfg->bn_momentum; f->eps = cfg->bn_eps; f->H = H; f->use_bn = cfg->use_bn; f->bn_mode = t->bn_mode;
    f->mean_z = w.mean_z[l]; f->rstd = w.rstd[l]; f->scale = w.scale[l]; f->shift = w.shift[l];
}

static int fwd_layer_gemm(const lifter_config_t* cfg, const lifter_train_ctx_t* t, int l, void* stream, bool fused) {
    Layout lay;
    LIFTER_PROPAGATE(check_ctx(cfg, t, &lay, false));
    LIFTER_REQUIRE(l >= 0 && l < lay.L, "forward: layer %d out of range", l);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    SavedWs w;
    carve_saved(lay, t->batch, t->saved, &w);
    const int B = static_cast<int>(t->batch), H = lay.H;
    const uint16_t* ops = static_cast<const uint16_t*>(t->operands);
    if (l == 0) LIFTER_PROPAGATE(prep_input(t->x, B, lay.I, lay.Ip, nullptr, nullptr, w.a0, t->fmt, st, w.bars, GRID_BARRIER_WORDS));
    const bool stats = cfg->use_bn && t->bn_mode == 0;
    GemmArgs g;
    g.A = l == 0 ? w.a0 : w.a[l - 1]; g.lda = lay.in_pad(l);
    g.B = ops + lay.w_off[l]; g.ldb = lay.in_pad(l);
    g.M = B; g.N = H; g.K = lay.in_pad(l); g.fmt = t->fmt; g.out_mode = OUT_16_TMA; g.D = w.z[l]; g.ldd = H;
    if (stats) { g.stat_sum = w.psum; g.stat_sq = w.psq; }
    g.tag = "gemm_fwd";
    FusedBnArgs fb;
    if (fwd_whole_layer(cfg, t, lay, l, fused)) {              // GEMM + statistics + BN / ReLU / dropout / residual
        fill_finalize(cfg, t, lay, w, l, &fb.f);
        LIFTER_PROPAGATE(fill_exchange(lay, t, l, &fb.f.ex));
        fb.sums_out = w.sums[l]; fb.a_out = w.a[l]; fb.bits = w.bits[l];
        fb.resid = (l >= 2 && l % 2 == 0) ? w.a[l - 2] : nullptr; fb.ldr = H;
        fb.apply_dropout = t->apply_dropout && cfg->p_dropout > 0.0f; fb.p = cfg->p_dropout; fb.seed = t->seed;
        fb.step = t->step; fb.site = static_cast<unsigned>(l);
        fb.step_mul = t->calls_per_step > 0 ? static_cast<unsigned>(t->calls_per_step) : 1u;
        fb.step_add = static_cast<unsigned>(t->call_index);
        fb.step_dev = t->step_state != nullptr ? reinterpret_cast<const unsigned*>(t->step_state) : nullptr;
        fb.row_offset = t->row_offset;
        fb.barrier = w.bars + 2 * l;
        g.fused_bn = &fb;
        g.tag = "layer_fwd";
    }
    // debug: LIFTER_DEBUG_TIMELINE_PTR=<device address of uint64 [layers][SMs][32]> makes the forward GEMMs stamp their
    // phases with %globaltimer (scripts/layer_timeline.py); never set in production
    if (const char* e = getenv("LIFTER_DEBUG_TIMELINE_PTR"))
        g.timeline = reinterpret_cast<unsigned long long*>(strtoull(e, nullptr, 0)) + static_cast<size_t>(l) * num_sms() * 32;
    LIFTER_PROPAGATE(gemm_launch(g, st));
    if (stats && !fused)
        LIFTER_PROPAGATE(reduce_rows_f64(w.psum, w.psq, gemm_stat_rows(B), H, H, w.sums[l], w.sums[l] + H, st));
    return LIFTER_OK;
}

static int fwd_layer_finish(const lifter_config_t* cfg, const lifter_train_ctx_t* t, int l, void* stream, bool fused) {
    Layout lay;
    LIFTER_PROPAGATE(check_ctx(cfg, t, &lay, false));
    LIFTER_REQUIRE(l >= 0 && l < lay.L, "forward: layer %d out of range", l);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (fwd_whole_layer(cfg, t, lay, l, fused)) return LIFTER_OK;          // done inside the layer kernel
    SavedWs w;
    carve_saved(lay, t->batch, t->saved, &w);
    const int H = lay.H;
    BnFwdFinalize f;
    f.sums = w.sums[l]; f.count = static_cast<double>(t->global_batch);
    f.bias = t->params + lay.p_off[lay.t_bias(l)];
    f.gamma = t->params + lay.p_off[lay.t_gamma(l)];
    f.beta = t->params + lay.p_off[lay.t_beta(l)];
    f.running_mean = t->running_mean ? t->running_mean + static_cast<int64_t>(l) * H : nullptr;
    f.running_var = t->running_var ? t->running_var + static_cast<int64_t>(l) * H : nullptr;
    f.num_batches_tracked = t->num_batches_tracked ? reinterpret_cast<long long*>(t->num_batches_tracked) + l : nullptr;
    f.momentum = cfg->bn_momentum; f.eps = cfg->bn_eps; f.H = H; f.use_bn = cfg->use_bn; f.bn_mode = t->bn_mode;
    f.mean_z = w.mean_z[l]; f.rstd = w.rstd[l]; f.scale = w.scale[l]; f.shift = w.shift[l];
    if (fused) LIFTER_PROPAGATE(fill_exchange(lay, t, l, &f.ex));          // forward site l
    const bool batch_stats = fused && cfg->use_bn && t->bn_mode == 0;
    const bool one_launch = batch_stats && l < GRID_BARRIER_WORDS && fuse_bn_launches(true);
    if (batch_stats && !one_launch)
        LIFTER_PROPAGATE(bn_reduce_finalize_fwd(w.psum, w.psq, gemm_stat_rows(static_cast<int>(t->batch)), f, w.sums[l], st));
    else if (!batch_stats)
        LIFTER_PROPAGATE(bn_finalize_fwd(f, st));
    BnApplyFwd a;
    a.z = w.z[l]; a.scale = w.scale[l]; a.shift = w.shift[l];
    a.resid = (l >= 2 && l % 2 == 0) ? w.a[l - 2] : nullptr;                // out = x + y, baselineModel.py:45
    a.rows = t->batch; a.H = H; a.fmt = t->fmt;
    a.apply_dropout = t->apply_dropout && cfg->p_dropout > 0.0f; a.p = cfg->p_dropout;
    a.seed = t->seed; a.step = t->step; a.site = static_cast<unsigned>(l); a.row_offset = t->row_offset;
    a.step_dev = t->step_state != nullptr ? reinterpret_cast<const unsigned*>(t->step_state) : nullptr;
    a.step_mul = t->calls_per_step > 0 ? static_cast<unsigned>(t->calls_per_step) : 1u;
    a.step_add = static_cast<unsigned>(t->call_index);
    a.a = w.a[l]; a.bits = w.bits[l];
    if (one_launch)     // column reduction + statistics + apply behind one grid barrier: one launch instead of two
        return bn_fwd_fused(w.psum, w.psq, gemm_stat_rows(static_cast<int>(t->batch)), f, w.sums[l], a, w.bars + l, st);
    LIFTER_PROPAGATE(bn_apply_fwd(a, st));
    return LIFTER_OK;
}

int lifter_train_fwd_layer_gemm(const lifter_config_t* cfg, const lifter_train_ctx_t* t, int l, void* stream) {
    return fwd_layer_gemm(cfg, t, l, stream, false);
}
int lifter_train_fwd_layer_finish(const lifter_config_t* cfg, const lifter_train_ctx_t* t, int l, void* stream) {
    return fwd_layer_finish(cfg, t, l, stream, false);
}

int lifter_train_fwd_output(const lifter_config_t* cfg, const lifter_train_ctx_t* t, void* stream) {
    Layout lay;
    LIFTER_PROPAGATE(check_ctx(cfg, t, &lay, false));
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    SavedWs w;
    carve_saved(lay, t->batch, t->saved, &w);
    GemmArgs g;
    g.A = w.a[lay.L - 1]; g.lda = lay.H;
    g.B = static_cast<const uint16_t*>(t->operands) + lay.w_off[lay.L]; g.ldb = lay.H;
    g.M = static_cast<int>(t->batch); g.N = lay.Op; g.K = lay.H; g.fmt = t->fmt; g.out_mode = OUT_32_DIRECT;
    g.D = t->out; g.n_valid = lay.O;
    g.shift = t->params + lay.p_off[lay.t_b2()]; g.shift_valid = lay.O;   // + b2 (columns past o_dim are never stored)
    g.tag = "gemm_fwd_out";
    return gemm_launch(g, st);
}

int lifter_train_forward(const lifter_config_t* cfg, const lifter_train_ctx_t* t, void* stream) {
    Layout lay;
    LIFTER_PROPAGATE(check_ctx(cfg, t, &lay, false));
    for (int l = 0; l < lay.L; ++l) {
        LIFTER_PROPAGATE(fwd_layer_gemm(cfg, t, l, stream, true));
        LIFTER_PROPAGATE(fwd_layer_finish(cfg, t, l, stream, true));
    }
    return lifter_train_fwd_output(cfg, t, stream);
}

// ---------------------------------------------------------------------------------------------- backward
// The weight-gradient GEMMs feed nothing but the optimiser, so the whole-backward entry point runs them on a second
// stream: dW_l overlaps the BatchNorm-backward passes and the data-gradient GEMM of the layers below it.  fork[i] is
// recorded on the caller's stream once dz of layer i exists, done[i] on the side stream once dW_i is reduced; index L
// is the output layer.  Works under stream capture (the side stream joins the caller's before the call returns).
struct SideStream {
    cudaStream_t st = nullptr;
    cudaStream_t opt = nullptr;            // optimiser ranges (lifter_train_backward_step)
    std::vector<cudaEvent_t> fork, done, mdone;   // mdone[i]: the caller's stream is done with layer i's old weights
    cudaEvent_t odone = nullptr;
};
static SideStream* side_stream(int slots) {
    static SideStream per_device[64];
    static const bool off = [] { const char* e = getenv("LIFTER_WGRAD_STREAM"); return e != nullptr && e[0] == '0'; }();
    if (off) return nullptr;
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return nullptr;
    SideStream* s = &per_device[dev];
    if (s->st == nullptr && cudaStreamCreateWithFlags(&s->st, cudaStreamNonBlocking) != cudaSuccess) return nullptr;
    if (s->opt == nullptr && cudaStreamCreateWithFlags(&s->opt, cudaStreamNonBlocking) != cudaSuccess) return nullptr;
    if (s->odone == nullptr && cudaEventCreateWithFlags(&s->odone, cudaEventDisableTiming) != cudaSuccess) return nullptr;
    while (static_cast<int>(s->fork.size()) < slots) {
        cudaEvent_t a, b, c;
        if (cudaEventCreateWithFlags(&a, cudaEventDisableTiming) != cudaSuccess) return nullptr;
        if (cudaEventCreateWithFlags(&b, cudaEventDisableTiming) != cudaSuccess) return nullptr;
        if (cudaEventCreateWithFlags(&c, cudaEventDisableTiming) != cudaSuccess) return nullptr;
        s->fork.push_back(a); s->done.push_back(b); s->mdone.push_back(c);
    }
    return s;
}
static int side_fork(SideStream* s, int slot, cudaStream_t main_st) {
    LIFTER_CHECK_CUDA(cudaEventRecord(s->fork[slot], main_st));
    LIFTER_CHECK_CUDA(cudaStreamWaitEvent(s->st, s->fork[slot], 0));
    return LIFTER_OK;
}

static int wgrad(const Layout& lay, const lifter_train_ctx_t* t, const ScratchWs& b, const uint16_t* dz, int M, long ldz,
                 const uint16_t* a_in, int N, long lda_in, int rows_valid, int cols_valid, float* grad, cudaStream_t st) {
    const int B = static_cast<int>(t->batch);
    GemmArgs g;                                                   // dW[M][N] = dz^T [M][B] * a_in [B][N]
    g.A = dz; g.lda = ldz; g.a_mn_major = 1;
    g.B = a_in; g.ldb = lda_in; g.b_mn_major = 1;
    g.M = M; g.N = N; g.K = B; g.fmt = t->fmt; g.out_mode = OUT_32_TMA; g.D = b.slabs; g.ldd = N;
    g.splits = wgrad_splits(M, N, B);
    g.tag = "gemm_wgrad";
    LIFTER_PROPAGATE(gemm_launch(g, st));
    return wgrad_reduce(b.slabs, gemm_effective_splits(B, g.splits), gemm_slab_rows(M), N, rows_valid, cols_valid, grad, st);
}

static int bwd_begin(const lifter_config_t* cfg, const lifter_train_ctx_t* t, void* stream, SideStream* side) {
    Layout lay;
    LIFTER_PROPAGATE(check_ctx(cfg, t, &lay, true));
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    SavedWs w; ScratchWs b;
    carve_saved(lay, t->batch, t->saved, &w);
    carve_scratch(lay, t->batch, t->scratch, &b);
    const int B = static_cast<int>(t->batch), H = lay.H;
    const uint16_t* ops = static_cast<const uint16_t*>(t->operands);
    // grad_out -> 16-bit padded, bias gradient of the output layer
    LIFTER_PROPAGATE(prep_grad_out(t->grad_out, B, lay.O, lay.Op, t->fmt, b.dO, b.pO, st, b.bars, GRID_BARRIER_WORDS));
    // db2 and dW2 = dOut^T a_{L-1} feed only the optimiser: side stream when there is one
    if (side != nullptr) LIFTER_PROPAGATE(side_fork(side, lay.L, st));
    LIFTER_PROPAGATE(reduce_rows_f32(b.pO, ew_partial_rows(B, lay.Op), lay.O, lay.Op, t->grad_params + lay.p_off[lay.t_b2()],
                                     side != nullptr ? side->st : st));
    LIFTER_PROPAGATE(wgrad(lay, t, b, b.dO, lay.Op, lay.Op, w.a[lay.L - 1], H, H, lay.O, H,
                           t->grad_params + lay.p_off[lay.t_w2()], side != nullptr ? side->st : st));
    if (side != nullptr) LIFTER_CHECK_CUDA(cudaEventRecord(side->done[lay.L], side->st));
    // gradient of the residual stream at the top: g = dOut W2
    GemmArgs g;
    g.A = b.dO; g.lda = lay.Op; g.B = ops + lay.w_off[lay.L]; g.ldb = H; g.b_mn_major = 1;
    g.M = B; g.N = H; g.K = lay.Op; g.fmt = t->fmt; g.out_mode = OUT_16_TMA; g.D = b.G[0]; g.ldd = H;
    g.tag = "gemm_dgrad_top";
    return gemm_launch(g, st);
}
int lifter_train_bwd_begin(const lifter_config_t* cfg, const lifter_train_ctx_t* t, void* stream) {
    return bwd_begin(cfg, t, stream, nullptr);
}

// which buffer holds the incoming gradient g_l of hidden layer l, given that the stream gradient after
// bwd_begin sits in G[0] and flips at every block boundary
static const uint16_t* g_in(const Layout& lay, const ScratchWs& b, int l) {
    if (l % 2 == 1) return b.Gh;                                  // first layer of a block: dh from the second
    const int block_from_top = (lay.L - 1 - l) / 2;               // l even: stream gradient dX
    return b.G[block_from_top & 1];
}

static float inv_keep_of(const lifter_config_t* cfg, const lifter_train_ctx_t* t) {
    const bool drop = t->apply_dropout && cfg->p_dropout > 0.0f;
    return drop ? (cfg->p_dropout < 1.0f ? 1.0f / (1.0f - cfg->p_dropout) : 0.0f) : 1.0f;
}

static BnBwdFinalize bwd_finalize_args(const lifter_config_t* cfg, const lifter_train_ctx_t* t, const Layout& lay,
                                       const SavedWs& w, const ScratchWs& b, int l) {
    BnBwdFinalize f;
    f.sums = b.bsums[l]; f.count = static_cast<double>(t->global_batch);
    f.gamma = t->params + lay.p_off[lay.t_gamma(l)]; f.rstd = w.rstd[l]; f.mean_z = w.mean_z[l];
    f.inv_keep = inv_keep_of(cfg, t);
    f.H = lay.H; f.use_bn = cfg->use_bn; f.bn_mode = t->bn_mode; f.coef = b.coef;
    // batch-statistics BN cancels the Linear bias: its gradient is exactly zero
    f.dbias = (cfg->use_bn && t->bn_mode == 0) ? t->grad_params + lay.p_off[lay.t_bias(l)] : nullptr;
    return f;
}
static int bwd_exchange(const lifter_config_t* cfg, const lifter_train_ctx_t* t, const Layout& lay, int l, bool fused,
                        BnBwdFinalize* f) {
    if (fused && cfg->use_bn && t->bn_mode == 0) return fill_exchange(lay, t, lay.L + l, &f->ex);   // backward site
    return LIFTER_OK;
}

// single-call backward, batch-statistics BatchNorm: stats -> reduce -> dz of a layer run as ONE grid-barrier launch
static bool bwd_one_launch(const lifter_config_t* cfg, const lifter_train_ctx_t* t, int l, bool fused) {
    return fused && cfg->use_bn && t->bn_mode == 0 && 2 * l + 1 < GRID_BARRIER_WORDS && fuse_bn_launches(false);
}

static int bwd_layer_stats(const lifter_config_t* cfg, const lifter_train_ctx_t* t, int l, void* stream, bool fused) {
    Layout lay;
    LIFTER_PROPAGATE(check_ctx(cfg, t, &lay, true));
    LIFTER_REQUIRE(l >= 0 && l < lay.L, "backward: layer %d out of range", l);
    if (!cfg->use_bn) return LIFTER_OK;
    if (bwd_one_launch(cfg, t, l, fused)) return LIFTER_OK;       // done by bwd_layer_finish
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    SavedWs w; ScratchWs b;
    carve_saved(lay, t->batch, t->saved, &w);
    carve_scratch(lay, t->batch, t->scratch, &b);
    const int H = lay.H;
    BwdCommon c;
    c.g = g_in(lay, b, l); c.z = w.z[l]; c.bits = w.bits[l];
    c.rows = t->batch; c.H = H; c.fmt = t->fmt;
    LIFTER_PROPAGATE(bwd_stats(c, b.pA, b.pB, st));
    // local sums (what a data-parallel run all-reduces), dgamma/dbeta from them; fused: also the pass-2 coefficients
    BnBwdFinalize f = bwd_finalize_args(cfg, t, lay, w, b, l);
    LIFTER_PROPAGATE(bwd_exchange(cfg, t, lay, l, fused, &f));
    return bn_bwd_reduce(b.pA, b.pB, ew_partial_rows(t->batch, H), f, b.bsums[l],
                         t->grad_params + lay.p_off[lay.t_gamma(l)], t->grad_params + lay.p_off[lay.t_beta(l)], fused, st);
}

static int bwd_layer_finish(const lifter_config_t* cfg, const lifter_train_ctx_t* t, int l, void* stream, bool fused,
                            SideStream* side = nullptr) {
    Layout lay;
    LIFTER_PROPAGATE(check_ctx(cfg, t, &lay, true));
    LIFTER_REQUIRE(l >= 0 && l < lay.L, "backward: layer %d out of range", l);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    SavedWs w; ScratchWs b;
    carve_saved(lay, t->batch, t->saved, &w);
    carve_scratch(lay, t->batch, t->scratch, &b);
    const int B = static_cast<int>(t->batch), H = lay.H;
    const uint16_t* ops = static_cast<const uint16_t*>(t->operands);

    BnBwdFinalize f = bwd_finalize_args(cfg, t, lay, w, b, l);
    const bool one_launch = bwd_one_launch(cfg, t, l, fused);
    if (!(fused && cfg->use_bn)) LIFTER_PROPAGATE(bn_finalize_bwd(f, st));   // fused: done by bwd_layer_stats
    if (!cfg->use_bn) {                                            // the BN affine exists but is unused: zero gradient
        LIFTER_CHECK_CUDA(cudaMemsetAsync(t->grad_params + lay.p_off[lay.t_gamma(l)], 0, H * 4, st));
        LIFTER_CHECK_CUDA(cudaMemsetAsync(t->grad_params + lay.p_off[lay.t_beta(l)], 0, H * 4, st));
    }

    BwdCommon c;
    c.g = g_in(lay, b, l); c.z = w.z[l]; c.bits = w.bits[l];
    c.rows = t->batch; c.H = H; c.fmt = t->fmt;
    const bool need_bias = f.dbias == nullptr;                     // no BN, or BN on running statistics
    uint16_t* dz = b.dz[l & 1];
    if (side != nullptr && l + 2 < lay.L)                          // dW_{l+2} still reads this buffer on the side stream
        LIFTER_CHECK_CUDA(cudaStreamWaitEvent(st, side->done[l + 2], 0));
    if (one_launch) {
        LIFTER_PROPAGATE(bwd_exchange(cfg, t, lay, l, fused, &f));
        LIFTER_PROPAGATE(bn_bwd_fused(c, b.pA, b.pB, f, b.bsums[l], t->grad_params + lay.p_off[lay.t_gamma(l)],
                                      t->grad_params + lay.p_off[lay.t_beta(l)], dz, b.bars + 2 * l, st));
    } else {
        LIFTER_PROPAGATE(bwd_dz(c, b.coef, dz, need_bias ? b.pA : nullptr, st));
    }
    if (need_bias)
        LIFTER_PROPAGATE(reduce_rows_f32(b.pA, ew_partial_rows(t->batch, H), H, H, t->grad_params + lay.p_off[lay.t_bias(l)], st));

    // dW_l = dz^T a_{l-1}
    const uint16_t* a_in = l == 0 ? w.a0 : w.a[l - 1];
    if (side != nullptr) LIFTER_PROPAGATE(side_fork(side, l, st));
    LIFTER_PROPAGATE(wgrad(lay, t, b, dz, H, H, a_in, lay.in_pad(l), lay.in_pad(l), H, lay.in_dim(l),
                           t->grad_params + lay.p_off[lay.t_weight(l)], side != nullptr ? side->st : st));
    if (side != nullptr) LIFTER_CHECK_CUDA(cudaEventRecord(side->done[l], side->st));

    if (l == 0) {
        if (t->grad_x == nullptr) return LIFTER_OK;
        GemmArgs g;                                                // dx = dz_0 W1   (input gradient, phase-5 call site)
        g.A = dz; g.lda = H; g.B = ops + lay.w_off[0]; g.ldb = lay.Ip; g.b_mn_major = 1;
        g.M = B; g.N = lay.Ip; g.K = H; g.fmt = t->fmt; g.tag = "gemm_dx";
        if (lay.Ip <= 64) {
            g.out_mode = OUT_32_DIRECT; g.D = t->grad_x; g.n_valid = lay.I;
        } else {
            LIFTER_REQUIRE(lay.I % 4 == 0, "input gradient needs i_dim %% 4 == 0 when i_dim > 64 (got %d)", lay.I);
            g.N = lay.I; g.out_mode = OUT_32_TMA; g.D = t->grad_x; g.ldd = lay.I;
        }
        return gemm_launch(g, st);
    }
    GemmArgs g;                                                    // g_{l-1} = dz_l W_l (+ stream gradient)
    g.A = dz; g.lda = H; g.B = ops + lay.w_off[l]; g.ldb = H; g.b_mn_major = 1;
    g.M = B; g.N = H; g.K = H; g.fmt = t->fmt; g.out_mode = OUT_16_TMA; g.ldd = H; g.tag = "gemm_dgrad";
    if (l % 2 == 0) {
        g.D = b.Gh;                                                // second layer of a block -> dh
    } else {
        const int block_from_top = (lay.L - 1 - (l + 1)) / 2;     // the block whose output gradient we consumed
        g.residual = b.G[block_from_top & 1]; g.ldr = H;           // dX_i = dX_{i+1} + dz W
        g.D = b.G[(block_from_top + 1) & 1];
    }
    return gemm_launch(g, st);
}

int lifter_train_bwd_layer_stats(const lifter_config_t* cfg, const lifter_train_ctx_t* t, int l, void* stream) {
    return bwd_layer_stats(cfg, t, l, stream, false);
}
int lifter_train_bwd_layer_finish(const lifter_config_t* cfg, const lifter_train_ctx_t* t, int l, void* stream) {
    return bwd_layer_finish(cfg, t, l, stream, false);
}

int lifter_train_backward(const lifter_config_t* cfg, const lifter_train_ctx_t* t, void* stream) {
    Layout lay;
    LIFTER_PROPAGATE(check_ctx(cfg, t, &lay, true));
    SideStream* side = side_stream(lay.L + 1);
    LIFTER_PROPAGATE(bwd_begin(cfg, t, stream, side));
    for (int l = lay.L - 1; l >= 0; --l) {
        LIFTER_PROPAGATE(bwd_layer_stats(cfg, t, l, stream, true));
        LIFTER_PROPAGATE(bwd_layer_finish(cfg, t, l, stream, true, side));
    }
    if (side != nullptr)                                           // dW_0 is the last thing on the side stream
        LIFTER_CHECK_CUDA(cudaStreamWaitEvent(static_cast<cudaStream_t>(stream), side->done[0], 0));
    return LIFTER_OK;
}

// ---- backward with the optimiser folded in -----------------------------------------------------------------------
static void fill_operand_table(const Layout& lay, AdamArgs* a, void* operands, int fmt) {
    a->operands = operands; a->op_fmt = fmt; a->nw = lay.L + 1;
    for (int l = 0; l <= lay.L; ++l) {
        const int t = l < lay.L ? lay.t_weight(l) : lay.t_w2();
        a->w_p_off[l] = lay.p_off[t]; a->w_numel[l] = lay.p_numel[t]; a->w_off[l] = lay.w_off[l];
        a->w_in[l] = l < lay.L ? lay.in_dim(l) : lay.H;
        a->w_pad[l] = l < lay.L ? lay.in_pad(l) : lay.H;
    }
}

// Adam over the flat-buffer range of layer l (W, b, gamma, beta are consecutive; l == L: the output layer W2, b2)
static int optim_range(const Layout& lay, const lifter_train_ctx_t* t, const lifter_optim_t* o, int l, int phase, bool last,
                       cudaStream_t st) {
    const int64_t lo = lay.p_off[l < lay.L ? lay.t_weight(l) : lay.t_w2()];
    const int64_t hi = l < lay.L ? lay.p_off[l + 1 < lay.L ? lay.t_weight(l + 1) : lay.t_w2()] : lay.p_total;
    const bool dp = t->dp != nullptr && t->dp->world > 1;
    AdamDpArgs d;
    AdamArgs& a = d.a;
    a.p = o->params; a.g = t->grad_params; a.m = o->m; a.v = o->v; a.base = lo; a.n = hi - lo;
    a.lr = 0.0f; a.beta1 = o->beta1; a.beta2 = o->beta2; a.eps = o->eps; a.weight_decay = o->weight_decay;
    a.decoupled = o->decoupled; a.step = 1; a.grad_scale = o->grad_scale;
    a.dyn = reinterpret_cast<const float*>(t->step_state);
    fill_operand_table(lay, &a, o->operands, t->fmt);
    if (!dp) return adam_step(a, st);
    const lifter_dp_t* g = t->dp;
    d.world = g->world; d.rank = g->rank;
    for (int q = 0; q < g->world; ++q) {
        d.g_peer[q] = static_cast<const float*>(g->grad[q]);
        d.p_peer[q] = static_cast<float*>(g->params[q]);
        d.op_peer[q] = g->operands[q];
        d.flag_peer[q] = static_cast<unsigned*>(g->flags[q]);
    }
    d.epoch = reinterpret_cast<const unsigned*>(t->step_state) + 1;
    d.phase = phase; d.last = last ? 1 : 0;
    if (g->mc_grad && g->mc_params && g->mc_operands && dp_multicast_enabled(g->world)) {
        d.mc_g = static_cast<const float*>(g->mc_grad); d.mc_p = static_cast<float*>(g->mc_params); d.mc_op = g->mc_operands;
    }
    return adam_step_dp(d, st);
}

int lifter_train_backward_step(const lifter_config_t* cfg, const lifter_train_ctx_t* t, const lifter_optim_t* o,
                               void* stream) {
    Layout lay;
    LIFTER_PROPAGATE(check_ctx(cfg, t, &lay, true));
    LIFTER_REQUIRE(o != nullptr && o->params && o->operands && o->m && o->v, "backward_step: null optimiser buffer");
    LIFTER_REQUIRE(t->step_state != nullptr, "backward_step: needs the device step state (ctx->step_state)");
    LIFTER_REQUIRE(lay.L + 1 <= ADAM_MAX_WEIGHTS, "backward_step: too many layers for the fused operand refresh");
    LIFTER_REQUIRE(o->params == t->params && o->operands == t->operands,
                   "backward_step: the optimiser must own the buffers the context reads");
    if (t->dp != nullptr && t->dp->world > 1) {
        const lifter_dp_t* g = t->dp;
        LIFTER_REQUIRE(g->world <= ADAM_DP_MAX_WORLD && g->rank >= 0 && g->rank < g->world, "backward_step: bad world/rank");
        for (int q = 0; q < g->world; ++q)
            LIFTER_REQUIRE(g->grad[q] && g->params[q] && g->operands[q] && g->flags[q], "backward_step: null peer buffer %d", q);
        LIFTER_REQUIRE(g->grad[g->rank] == t->grad_params && g->params[g->rank] == o->params &&
                       g->operands[g->rank] == o->operands, "backward_step: this rank's peer entries must be its own buffers");
    }
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    SideStream* side = side_stream(lay.L + 1);
    // range i is safe to update once (a) dW_i is reduced [done[i], side stream] and (b) this stream has consumed the old
    // W_i in the data-gradient GEMM [mdone[i]]
    auto range = [&](int l, int phase, bool last) -> int {
        if (side == nullptr) return optim_range(lay, t, o, l, phase, last, st);
        LIFTER_CHECK_CUDA(cudaEventRecord(side->mdone[l], st));
        LIFTER_CHECK_CUDA(cudaStreamWaitEvent(side->opt, side->mdone[l], 0));
        LIFTER_CHECK_CUDA(cudaStreamWaitEvent(side->opt, side->done[l], 0));
        return optim_range(lay, t, o, l, phase, last, side->opt);
    };
    LIFTER_PROPAGATE(bwd_begin(cfg, t, stream, side));
    LIFTER_PROPAGATE(range(lay.L, 0, false));
    for (int l = lay.L - 1; l >= 0; --l) {
        LIFTER_PROPAGATE(bwd_layer_stats(cfg, t, l, stream, true));
        LIFTER_PROPAGATE(bwd_layer_finish(cfg, t, l, stream, true, side));
        LIFTER_PROPAGATE(range(l, lay.L - l, l == 0));
    }
    if (side != nullptr) {
        LIFTER_CHECK_CUDA(cudaEventRecord(side->odone, side->opt));
        LIFTER_CHECK_CUDA(cudaStreamWaitEvent(st, side->odone, 0));
        LIFTER_CHECK_CUDA(cudaStreamWaitEvent(st, side->done[0], 0));
    }
    return LIFTER_OK;
}

// ---------------------------------------------------------------------------------------------- loss / optimiser
int64_t lifter_loss_scratch_bytes(int64_t rows, int joints) {
    if (rows < 1 || joints < 1) return LIFTER_ERR_INVALID;
    return static_cast<int64_t>(loss_partial_rows(rows)) * (1 + joints) * 4;
}

int lifter_loss(const float* pred, const float* target, int64_t rows, int joints, int kind, int64_t global_rows,
                const float* stdv, float* dpred, float* scratch, float* out, int accumulate_metric, void* stream) {
    LIFTER_REQUIRE(pred && target && scratch && out, "loss: null pointer");
    return fused_loss(pred, target, rows, joints, kind, static_cast<double>(global_rows), stdv, dpred, scratch, out,
                      accumulate_metric, static_cast<cudaStream_t>(stream));
}

int lifter_flip_pose(const float* in, float* out, int64_t poses, int joints, int dims, int grad, void* stream) {
    LIFTER_REQUIRE(in && out, "flip_pose: null pointer");
    return flip_pose(in, out, poses, joints, dims, grad, static_cast<cudaStream_t>(stream));
}

int lifter_coco2h36m(const float* in, int in_channels, float* out, int64_t poses, float scale, const float* mean,
                     const float* stdv, int mode, void* stream) {
    return coco2h36m(in, in_channels, out, poses, scale, mean, stdv, mode, static_cast<cudaStream_t>(stream));
}

int lifter_keypoint_heatmaps(const void* keypoints, int keypoints_are_f64, int64_t n, double sigma, float* out,
                             int* status, void* stream) {
    return keypoint_heatmaps(keypoints, keypoints_are_f64, n, sigma, out, status, static_cast<cudaStream_t>(stream));
}

int lifter_affine_cols(const float* in, float* out, int64_t rows, int cols, const float* a, const float* b, int mode,
                       void* stream) {
    LIFTER_REQUIRE(in && out, "affine_cols: null pointer");
    return affine_cols(in, out, rows, cols, a, b, mode, static_cast<cudaStream_t>(stream));
}

int64_t lifter_pose_loss_scratch_bytes(int64_t rows, int joints) {
    if (rows < 1 || joints < 1) return LIFTER_ERR_INVALID;
    return static_cast<int64_t>(pose_loss_partial_rows(rows)) * (4 + joints) * 4;
}

int lifter_pose_loss(const lifter_pose_loss_t* d, void* stream) {
    LIFTER_REQUIRE(d != nullptr, "pose_loss: null descriptor");
    PoseLossArgs a;
    a.pred = d->pred; a.target = d->target; a.mean = d->mean; a.stdv = d->stdv; a.root = d->root;
    a.intrinsics = d->intrinsics; a.target2d = d->target2d; a.rows = d->rows;
    a.global_rows = d->global_rows > 0 ? d->global_rows : d->rows; a.J = d->joints;
    a.w_mpjpe = d->w_mpjpe; a.w_reproj = d->w_reproj; a.w_bone = d->w_bone;
    a.dpred = d->dpred; a.partials = d->scratch; a.out = d->out;
    return pose_loss(a, static_cast<cudaStream_t>(stream));
}

int64_t lifter_phase5_loss_scratch_bytes(int64_t rows) {
    if (rows < 1) return LIFTER_ERR_INVALID;
    return static_cast<int64_t>(phase5_partial_rows(rows)) * 8 * 4;
}

int lifter_phase5_loss(const lifter_phase5_loss_t* d, void* stream) {
    LIFTER_REQUIRE(d != nullptr, "phase5_loss: null descriptor");
    Phase5LossArgs a;
    a.out_pred = d->out_pred; a.out_pred_flip = d->out_pred_flip; a.out_gt = d->out_gt; a.out_gt_flip = d->out_gt_flip;
    a.pred2d = d->pred2d; a.gt2d = d->gt2d; a.pred3d = d->pred3d; a.gt3d = d->gt3d; a.mode = d->mode;
    a.mean3d = d->mean3d; a.std3d = d->std3d; a.mean2d = d->mean2d; a.std2d = d->std2d; a.root = d->root;
    a.intrinsics = d->intrinsics; a.w_reproj = d->intrinsics != nullptr ? d->w_reproj : 0.0f; a.w_bone = d->w_bone;
    a.rows = d->rows; a.global_rows = d->global_rows > 0 ? d->global_rows : d->rows; a.J = d->joints;
    a.d_out_pred = d->d_out_pred; a.d_out_pred_flip = d->d_out_pred_flip; a.d_out_gt = d->d_out_gt;
    a.d_out_gt_flip = d->d_out_gt_flip; a.d_pred2d = d->d_pred2d; a.d_pred3d = d->d_pred3d;
    a.partials = d->scratch; a.out = d->out;
    return phase5_loss(a, static_cast<cudaStream_t>(stream));
}

int lifter_phase5_input_grad(const float* d_own, const float* dx, const float* dx_flip, int64_t rows, int joints,
                             float* out, void* stream) {
    return phase5_input_grad(d_own, dx, dx_flip, rows, joints, out, static_cast<cudaStream_t>(stream));
}

int lifter_sum_buffers(float* dst, const float* a, const float* b, const float* c, const float* d, int64_t n,
                       void* stream) {
    return sum_buffers(dst, a, b, c, d, n, static_cast<cudaStream_t>(stream));
}

int lifter_l1_sum(const float* a, const float* b, int64_t n, int row_len, int centre, float scale, float* scratch,
                  float* out, void* stream) {
    LIFTER_REQUIRE(a && b && scratch && out, "l1_sum: null pointer");
    return l1_sum(a, b, n, row_len, centre, scale, scratch, out, static_cast<cudaStream_t>(stream));
}

int lifter_l1_grad(const float* a, const float* b, int64_t n, int row_len, int centre, const float* gscale,
                   float scale, float* da, float* db, void* stream) {
    LIFTER_REQUIRE(a && b && da, "l1_grad: null pointer");
    return l1_grad(a, b, n, row_len, centre, gscale, scale, da, db, static_cast<cudaStream_t>(stream));
}

int lifter_step_advance(lifter_step_state_t* state, float beta1, float beta2, float weight_decay, void* stream) {
    return step_advance(state, beta1, beta2, weight_decay, static_cast<cudaStream_t>(stream));
}

int lifter_adam_step_packed(const lifter_config_t* cfg, float* p, const float* g, float* m, float* v, float lr, float beta1,
                            float beta2, float eps, float weight_decay, int decoupled, int step, float grad_scale,
                            const lifter_step_state_t* state, void* operands, int fmt, void* stream) {
    Layout lay;
    LIFTER_PROPAGATE(make_layout(cfg, &lay));
    LIFTER_REQUIRE(p && g && m && v, "adam: null pointer");
    LIFTER_REQUIRE(fmt == FMT_F16 || fmt == FMT_BF16, "adam: unknown operand format %d", fmt);
    AdamArgs a;
    a.p = p; a.g = g; a.m = m; a.v = v; a.n = lay.p_total; a.lr = lr; a.beta1 = beta1; a.beta2 = beta2; a.eps = eps;
    a.weight_decay = weight_decay; a.decoupled = decoupled; a.step = step; a.grad_scale = grad_scale;
    a.dyn = reinterpret_cast<const float*>(state);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (operands != nullptr && lay.L + 1 <= ADAM_MAX_WEIGHTS) {
        a.operands = operands; a.op_fmt = fmt; a.nw = lay.L + 1;
        for (int l = 0; l <= lay.L; ++l) {
            const int t = l < lay.L ? lay.t_weight(l) : lay.t_w2();
            a.w_p_off[l] = lay.p_off[t]; a.w_numel[l] = lay.p_numel[t]; a.w_off[l] = lay.w_off[l];
            a.w_in[l] = l < lay.L ? lay.in_dim(l) : lay.H;
            a.w_pad[l] = l < lay.L ? lay.in_pad(l) : lay.H;
        }
        return adam_step(a, st);
    }
    LIFTER_PROPAGATE(adam_step(a, st));
    if (operands != nullptr) return lifter_pack_weights(cfg, p, operands, fmt, stream);   // very deep models
    return LIFTER_OK;
}

int lifter_adam_step_table(float* p, const float* g, float* m, float* v, int64_t n, float lr, float beta1, float beta2,
                           float eps, float weight_decay, int decoupled, int step, float grad_scale,
                           const lifter_step_state_t* state, void* operands, int fmt, const lifter_adam_table_t* table,
                           void* stream) {
    LIFTER_REQUIRE(p && g && m && v, "adam: null pointer");
    LIFTER_REQUIRE(fmt == FMT_F16 || fmt == FMT_BF16, "adam: unknown operand format %d", fmt);
    AdamArgs a;
    a.p = p; a.g = g; a.m = m; a.v = v; a.n = n; a.lr = lr; a.beta1 = beta1; a.beta2 = beta2; a.eps = eps;
    a.weight_decay = weight_decay; a.decoupled = decoupled; a.step = step; a.grad_scale = grad_scale;
    a.dyn = reinterpret_cast<const float*>(state);
    if (operands != nullptr) {
        LIFTER_REQUIRE(table != nullptr && table->n >= 1 && table->n <= ADAM_MAX_WEIGHTS, "adam: operand table with 1..%d ranges", ADAM_MAX_WEIGHTS);
        a.operands = operands; a.op_fmt = fmt; a.nw = table->n;
        for (int t = 0; t < table->n; ++t) {
            LIFTER_REQUIRE(table->p_off[t] >= 0 && table->numel[t] > 0 && table->p_off[t] + table->numel[t] <= n &&
                           table->p_off[t] % 4 == 0 && table->numel[t] % 4 == 0 && table->op_off[t] % 4 == 0 && table->op_off[t] >= 0,
                           "adam: operand range %d must lie inside the buffer with offsets / length multiples of 4", t);
            a.w_p_off[t] = table->p_off[t]; a.w_numel[t] = table->numel[t]; a.w_off[t] = table->op_off[t];
            a.w_in[t] = 1; a.w_pad[t] = 1;                              // dense: same linear offset in both buffers
        }
    }
    return adam_step(a, static_cast<cudaStream_t>(stream));
}

int lifter_dp_adam_step(const lifter_config_t* cfg, const lifter_dp_t* dp, float* m, float* v, float beta1, float beta2,
                        float eps, float weight_decay, int decoupled, float grad_scale,
                        const lifter_step_state_t* state, int fmt, void* stream) {
    Layout lay;
    LIFTER_PROPAGATE(make_layout(cfg, &lay));
    LIFTER_REQUIRE(dp != nullptr && m && v && state, "dp_adam: null pointer");
    LIFTER_REQUIRE(dp->world >= 2 && dp->world <= ADAM_DP_MAX_WORLD && dp->rank >= 0 && dp->rank < dp->world,
                   "dp_adam: world %d / rank %d out of range", dp->world, dp->rank);
    LIFTER_REQUIRE(fmt == FMT_F16 || fmt == FMT_BF16, "dp_adam: unknown operand format %d", fmt);
    LIFTER_REQUIRE(lay.L + 1 <= ADAM_MAX_WEIGHTS, "dp_adam: too many layers for the fused operand refresh");
    AdamDpArgs d;
    d.world = dp->world; d.rank = dp->rank;
    for (int q = 0; q < dp->world; ++q) {
        LIFTER_REQUIRE(dp->grad[q] && dp->params[q] && dp->operands[q] && dp->flags[q], "dp_adam: null peer buffer %d", q);
        d.g_peer[q] = static_cast<const float*>(dp->grad[q]);
        d.p_peer[q] = static_cast<float*>(dp->params[q]);
        d.op_peer[q] = dp->operands[q];
        d.flag_peer[q] = static_cast<unsigned*>(dp->flags[q]);
    }
    AdamArgs& a = d.a;
    a.p = d.p_peer[d.rank]; a.g = d.g_peer[d.rank]; a.m = m; a.v = v; a.n = lay.p_total;
    a.lr = 0.0f; a.beta1 = beta1; a.beta2 = beta2; a.eps = eps; a.weight_decay = weight_decay; a.decoupled = decoupled;
    a.step = 1; a.grad_scale = grad_scale; a.dyn = reinterpret_cast<const float*>(state);
    a.operands = d.op_peer[d.rank]; a.op_fmt = fmt; a.nw = lay.L + 1;
    for (int l = 0; l <= lay.L; ++l) {
        const int t = l < lay.L ? lay.t_weight(l) : lay.t_w2();
        a.w_p_off[l] = lay.p_off[t]; a.w_numel[l] = lay.p_numel[t]; a.w_off[l] = lay.w_off[l];
        a.w_in[l] = l < lay.L ? lay.in_dim(l) : lay.H;
        a.w_pad[l] = l < lay.L ? lay.in_pad(l) : lay.H;
    }
    d.epoch = reinterpret_cast<const unsigned*>(state) + 1;
    if (dp->mc_grad && dp->mc_params && dp->mc_operands && dp_multicast_enabled(dp->world)) {
        d.mc_g = static_cast<const float*>(dp->mc_grad); d.mc_p = static_cast<float*>(dp->mc_params); d.mc_op = dp->mc_operands;
    }
    return adam_step_dp(d, static_cast<cudaStream_t>(stream));
}

int lifter_dp_adam_step_table(const lifter_dp_t* dp, int64_t n, float* m, float* v, float beta1, float beta2, float eps,
                              float weight_decay, int decoupled, float grad_scale, const lifter_step_state_t* state,
                              int fmt, const lifter_adam_table_t* table, void* stream) {
    LIFTER_REQUIRE(dp != nullptr && m && v && state && table, "dp_adam: null pointer");
    LIFTER_REQUIRE(dp->world >= 2 && dp->world <= ADAM_DP_MAX_WORLD && dp->rank >= 0 && dp->rank < dp->world,
                   "dp_adam: world %d / rank %d out of range", dp->world, dp->rank);
    LIFTER_REQUIRE(fmt == FMT_F16 || fmt == FMT_BF16, "dp_adam: unknown operand format %d", fmt);
    LIFTER_REQUIRE(n > 0 && n % 4 == 0, "dp_adam: the buffer length must be a positive multiple of 4");
    LIFTER_REQUIRE(table->n >= 1 && table->n <= ADAM_MAX_WEIGHTS, "dp_adam: operand table with 1..%d ranges", ADAM_MAX_WEIGHTS);
    AdamDpArgs d;
    d.world = dp->world; d.rank = dp->rank;
    for (int q = 0; q < dp->world; ++q) {
        LIFTER_REQUIRE(dp->grad[q] && dp->params[q] && dp->operands[q] && dp->flags[q], "dp_adam: null peer buffer %d", q);
        d.g_peer[q] = static_cast<const float*>(dp->grad[q]);
        d.p_peer[q] = static_cast<float*>(dp->params[q]);
        d.op_peer[q] = dp->operands[q];
        d.flag_peer[q] = static_cast<unsigned*>(dp->flags[q]);
    }
    AdamArgs& a = d.a;
    a.p = d.p_peer[d.rank]; a.g = d.g_peer[d.rank]; a.m = m; a.v = v; a.n = n;
    a.lr = 0.0f; a.beta1 = beta1; a.beta2 = beta2; a.eps = eps; a.weight_decay = weight_decay; a.decoupled = decoupled;
    a.step = 1; a.grad_scale = grad_scale; a.dyn = reinterpret_cast<const float*>(state);
    a.operands = d.op_peer[d.rank]; a.op_fmt = fmt; a.nw = table->n;
    for (int t = 0; t < table->n; ++t) {
        LIFTER_REQUIRE(table->p_off[t] >= 0 && table->numel[t] > 0 && table->p_off[t] + table->numel[t] <= n &&
                       table->p_off[t] % 4 == 0 && table->numel[t] % 4 == 0 && table->op_off[t] % 4 == 0 && table->op_off[t] >= 0,
                       "dp_adam: operand range %d must lie inside the buffer with offsets / length multiples of 4", t);
        a.w_p_off[t] = table->p_off[t]; a.w_numel[t] = table->numel[t]; a.w_off[t] = table->op_off[t];
        a.w_in[t] = 1; a.w_pad[t] = 1;
    }
    d.epoch = reinterpret_cast<const unsigned*>(state) + 1;
    if (dp->mc_grad && dp->mc_params && dp->mc_operands && dp_multicast_enabled(dp->world)) {
        d.mc_g = static_cast<const float*>(dp->mc_grad); d.mc_p = static_cast<float*>(dp->mc_params); d.mc_op = dp->mc_operands;
    }
    return adam_step_dp(d, static_cast<cudaStream_t>(stream));
}

int lifter_adam_step(float* p, const float* g, float* m, float* v, int64_t n, float lr, float beta1, float beta2,
                     float eps, float weight_decay, int decoupled, int step, float grad_scale, void* stream) {
    LIFTER_REQUIRE(p && g && m && v, "adam: null pointer");
    AdamArgs a;
    a.p = p; a.g = g; a.m = m; a.v = v; a.n = n; a.lr = lr; a.beta1 = beta1; a.beta2 = beta2; a.eps = eps;
    a.weight_decay = weight_decay; a.decoupled = decoupled; a.step = step; a.grad_scale = grad_scale;
    return adam_step(a, static_cast<cudaStream_t>(stream));
}

}  // extern "C"
