// Device-side pieces of the BatchNorm statistics path shared by the elementwise kernels (train_kernels.cu) and the fused
// layer kernel (gemm_sm100.cu): the cross-GPU LL exchange of column sums, the per-column forward finalize (batch
// statistics -> scale / shift, running-statistics update) and the backward coefficients.
#pragma once
#include "train_kernels.cuh"

namespace lifter {

__device__ __forceinline__ float ldg_f32(const float* p) {
    float r;
    asm volatile("ld.global.cg.f32 %0, [%1];" : "=f"(r) : "l"(p));
    return r;
}

// ---------------------------------------------------------------------------------- cross-GPU sum (SyncBN)
// "LL" exchange (the low-latency protocol NCCL uses for small messages, written out here): every double travels as
// one 16-byte store {lo, epoch, hi, epoch} into the writer's slot of every peer's symmetric buffer; each 8-byte half
// validates itself through its epoch word, so no fence and no separate flag round trip are needed -- the receiver
// polls its LOCAL memory until both halves carry this step's epoch.  One NVLink store latency per exchange.
__device__ __forceinline__ void st_ll_f64(uint4* p, double v, unsigned epoch) {
    const unsigned long long b = static_cast<unsigned long long>(__double_as_longlong(v));
    asm volatile("st.volatile.global.v4.u32 [%0], {%1, %2, %3, %4};" ::"l"(p), "r"(static_cast<unsigned>(b)), "r"(epoch),
                 "r"(static_cast<unsigned>(b >> 32)), "r"(epoch)
                 : "memory");
}
__device__ __forceinline__ double ld_ll_f64(const uint4* p, unsigned epoch, long long spin_limit) {
    uint4 w;
    long long t0 = 0;
    unsigned spins = 0;
    while (true) {
        asm volatile("ld.volatile.global.v4.u32 {%0, %1, %2, %3}, [%4];"
                     : "=r"(w.x), "=r"(w.y), "=r"(w.z), "=r"(w.w) : "l"(p) : "memory");
        if (w.y == epoch && w.w == epoch) break;
        if ((++spins & 0x3FF) == 0 && spin_limit > 0) {     // a DEAD peer traps after LIFTER_DP_TIMEOUT_S; a late one is waited for
            const long long now = clock64();
            if (t0 == 0) t0 = now;
            else if (now - t0 > spin_limit) __trap();
        }
    }
    return __longlong_as_double(static_cast<long long>((static_cast<unsigned long long>(w.z) << 32) | w.x));
}

// Called by the first RED_COLS threads of a block (threadIdx.y == 0) holding the local sums (ta, tb) of column
// `col`; returns the sums over all ranks, added in rank order (identical on every GPU).  See SyncExchange.
__device__ __forceinline__ uint4 ld_ll_words(const uint4* p) {
    uint4 w;
    asm volatile("ld.volatile.global.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(w.x), "=r"(w.y), "=r"(w.z), "=r"(w.w) : "l"(p) : "memory");
    return w;
}
__device__ __forceinline__ double ll_value(const uint4& w) {
    return __longlong_as_double(static_cast<long long>((static_cast<unsigned long long>(w.z) << 32) | w.x));
}
// The default form: ONE thread per column stores its pair to every peer and then polls the peers' slots one after another.
__device__ __forceinline__ void exchange_column_sums_serial(const SyncExchange& ex, int H, int col, bool active, double& ta,
                                                     double& tb) {
    if (!active) return;
    const unsigned epoch = *ex.epoch;
    // [parity][site][slot = writer rank][2][H] LL words
    const size_t base = ((static_cast<size_t>(epoch & 1u) * ex.nsites + ex.site) * ex.world) * 2 * H;
    const size_t mine = base + static_cast<size_t>(ex.rank) * 2 * H;
    for (int p = 0; p < ex.world; ++p) {
        if (p == ex.rank) continue;
        uint4* d = reinterpret_cast<uint4*>(ex.data[p]) + mine;
        st_ll_f64(d + col, ta, epoch);
        st_ll_f64(d + H + col, tb, epoch);
    }
    const double la = ta, lb = tb;
    ta = 0.0; tb = 0.0;
    for (int r = 0; r < ex.world; ++r) {
        if (r == ex.rank) { ta += la; tb += lb; continue; }
        const uint4* d = reinterpret_cast<const uint4*>(ex.data[ex.rank]) + base + static_cast<size_t>(r) * 2 * H;
        ta += ld_ll_f64(d + col, epoch, ex.spin_limit);
        tb += ld_ll_f64(d + H + col, epoch, ex.spin_limit);
    }
}

// Warp-cooperative exchange.  Called by ALL 32 lanes of the warp whose lanes are (sub, column) = (lane / 8, lane % 8):
// the sub == 0 lane of a column holds the local sums (ta, tb) and gets the sums over all ranks back, added in rank order
// (identical, bit for bit, on every GPU); `active` = the column exists (the same for the four lanes of a column).
// ex.serial (the default): the sub == 0 lane runs exchange_column_sums_serial.  Otherwise (LIFTER_DP_EXCHANGE_WARP=1) the
// sub == 0 lane stores its pair into its slot on every peer and lane (sub, column) polls the slots of peers `sub` and
// `sub + 4`, both words of a slot in flight together, so that the wait is one L2 round trip for any world size, and the
// sub == 0 lane gathers the values with shuffles.  MEASURED NEGATIVE RESULT (same box, alternating runs, configs[2]): the
// cooperative form is SLOWER, 0.553 vs 0.541 ms per step at 2 ranks and 0.592 vs 0.574 ms at 8
// (profiles/r02_dp_exchange_ab_dp{2,8}.jsonl): the exchange's cost (25 us per step at 2 ranks, 55 us at 8) is not the
// receiver's polling order but the ten cross-rank synchronisation points themselves.
__device__ __forceinline__ void poll_slot(const uint4* pa, const uint4* pb, unsigned epoch, long long spin_limit, double& va,
                                          double& vb) {
    long long t0 = 0;
    unsigned spins = 0;
    while (true) {
        const uint4 wa = ld_ll_words(pa), wb = ld_ll_words(pb);
        if (wa.y == epoch && wa.w == epoch && wb.y == epoch && wb.w == epoch) {
            va = ll_value(wa);
            vb = ll_value(wb);
            return;
        }
        if ((++spins & 0xFF) == 0 && spin_limit > 0) {           // a DEAD peer traps after LIFTER_DP_TIMEOUT_S
            const long long now = clock64();
            if (t0 == 0) t0 = now;
            else if (now - t0 > spin_limit) __trap();
        }
    }
}
__device__ __forceinline__ void exchange_column_sums_warp(const SyncExchange& ex, int H, int col, bool active, int sub,
                                                          double& ta, double& tb) {
    if (ex.serial) {                                                // uniform over the launch
        if (sub == 0) exchange_column_sums_serial(ex, H, col, active, ta, tb);
        return;
    }
    const unsigned epoch = *ex.epoch;
    // [parity][site][slot = writer rank][2][H] LL words
    const size_t base = ((static_cast<size_t>(epoch & 1u) * ex.nsites + ex.site) * ex.world) * 2 * H;
    if (active && sub == 0) {
        const size_t mine = base + static_cast<size_t>(ex.rank) * 2 * H;
        for (int p = 0; p < ex.world; ++p) {
            if (p == ex.rank) continue;
            uint4* d = reinterpret_cast<uint4*>(ex.data[p]) + mine;
            st_ll_f64(d + col, ta, epoch);
            st_ll_f64(d + H + col, tb, epoch);
        }
    }
    const uint4* in = reinterpret_cast<const uint4*>(ex.data[ex.rank]) + base;
    double va[2] = {0.0, 0.0}, vb[2] = {0.0, 0.0};
#pragma unroll
    for (int k = 0; k < 2; ++k) {
        const int r = sub + 4 * k;
        if (active && r < ex.world && r != ex.rank)
            poll_slot(in + static_cast<size_t>(r) * 2 * H + col, in + static_cast<size_t>(r) * 2 * H + H + col, epoch,
                      ex.spin_limit, va[k], vb[k]);
    }
    const double la = ta, lb = tb;
    double sa = 0.0, sb = 0.0;
    const int c = static_cast<int>(lane_id()) & 7;
    for (int r = 0; r < ex.world; ++r) {                            // uniform over the warp: every lane shuffles
        const int src = (r & 3) * 8 + c;
        const double xa = __shfl_sync(0xffffffffu, r < 4 ? va[0] : va[1], src);
        const double xb = __shfl_sync(0xffffffffu, r < 4 ? vb[0] : vb[1], src);
        if (r == ex.rank) { sa += la; sb += lb; } else { sa += xa; sb += xb; }
    }
    if (sub == 0) { ta = sa; tb = sb; }
}

// Per-column inputs of the finalize that do not depend on the reduction: fetched before it so their latency hides.
struct BnColumnInputs { float bias, gamma, beta, rmean, rvar; };
__device__ __forceinline__ BnColumnInputs bn_column_inputs(const BnFwdFinalize& a, int j) {
    BnColumnInputs c = {0.0f, 1.0f, 0.0f, 0.0f, 1.0f};
    c.bias = a.bias[j];
    if (a.use_bn) {
        c.gamma = a.gamma[j]; c.beta = a.beta[j];
        if (a.running_mean != nullptr) { c.rmean = a.running_mean[j]; c.rvar = a.running_var[j]; }
    }
    return c;
}

__device__ __forceinline__ void bn_finalize_fwd_col(const BnFwdFinalize& a, int j, double sum, double sumsq,
                                                    const BnColumnInputs& c) {
    if (!a.use_bn) {
        a.mean_z[j] = 0.0f; a.rstd[j] = 1.0f; a.scale[j] = 1.0f; a.shift[j] = c.bias;
        return;
    }
    float mean, rstd;
    if (a.bn_mode == 0) {
        // the difference sumsq/n - m^2 needs fp64; everything after it is fp32 (no fp64 divide or square root)
        const double inv_n = 1.0 / a.count;                          // uniform: one divide per thread, off the chain
        const double m = sum * inv_n;
        double var = sumsq * inv_n - m * m;                          // biased batch variance (normalisation)
        if (var < 0.0) var = 0.0;
        const float varf = static_cast<float>(var);
        mean = static_cast<float>(m);
        rstd = 1.0f / sqrtf(varf + a.eps);
        if (a.running_mean != nullptr) {                             // torch: momentum 0.1, unbiased running var
            const float unbias = a.count > 1.0 ? static_cast<float>(a.count / (a.count - 1.0)) : 1.0f;
            a.running_mean[j] = (1.0f - a.momentum) * c.rmean + a.momentum * (mean + c.bias);
            a.running_var[j] = (1.0f - a.momentum) * c.rvar + a.momentum * (varf * unbias);
        }
    } else {                                                         // running statistics; z excludes the bias
        mean = c.rmean - c.bias;
        rstd = 1.0f / sqrtf(c.rvar + a.eps);
    }
    const float s = c.gamma * rstd;
    a.mean_z[j] = mean;
    a.rstd[j] = rstd;
    a.scale[j] = s;
    a.shift[j] = c.beta - mean * s;
}

struct BnBwdColumn { float rstd, mean, gamma; };         // fetched ahead of the reduction (latency off the chain)
__device__ __forceinline__ BnBwdColumn bn_bwd_column(const BnBwdFinalize& a, int j) {
    BnBwdColumn c = {1.0f, 0.0f, 1.0f};
    if (a.use_bn) { c.rstd = a.rstd[j]; c.mean = a.mean_z[j]; c.gamma = a.gamma[j]; }
    return c;
}
__device__ __forceinline__ void bn_bwd_coefficients(const BnBwdFinalize& a, int j, double sum_dy, double sum_dyz,
                                                    const BnBwdColumn& cv) {
    float k = 1.0f, c1 = 0.0f, c2 = 0.0f, rstd = 1.0f, mean = 0.0f;
    if (a.use_bn) {
        rstd = cv.rstd;
        mean = cv.mean;
        k = cv.gamma * rstd;
        if (a.bn_mode == 0) {
            const double inv_n = 1.0 / a.count;                 // uniform; the only fp64 divide, off the dependent chain
            c1 = static_cast<float>(sum_dy * inv_n);
            c2 = static_cast<float>(sum_dyz * inv_n);
        }
    }
    const float bz = -k * c2 * rstd;
    a.coef[j] = k * a.inv_keep;
    a.coef[a.H + j] = bz;
    a.coef[2 * a.H + j] = -k * c1 - bz * mean;
    if (a.dbias != nullptr) a.dbias[j] = 0.0f;      // a bias in front of batch-statistics BN has zero gradient
}


}  // namespace lifter
