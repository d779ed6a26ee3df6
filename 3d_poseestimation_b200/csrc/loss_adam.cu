#include "loss_adam.cuh"

#include <math.h>

namespace lifter {

namespace {

constexpr int LOSS_THREADS = 256;
constexpr int LOSS_ROWS_PER_BLOCK = 64;
constexpr int MAX_J = 21;                       // 64 output columns / 3

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int s = 16; s >= 1; s >>= 1) v += __shfl_xor_sync(0xffffffffu, v, s);
    return v;
}

// one thread per (row, joint): coalesced enough for 204-byte rows (consecutive threads = consecutive joints)
__global__ void __launch_bounds__(LOSS_THREADS) fused_loss_kernel(const float* __restrict__ pred,
                                                                  const float* __restrict__ target, long rows, int J,
                                                                  int kind, float inv_count,
                                                                  const float* __restrict__ stdv,
                                                                  float* __restrict__ dpred,
                                                                  float* __restrict__ partials) {
    griddep_launch();
    griddep_wait();
    __shared__ float sh_loss[LOSS_THREADS / 32];
    __shared__ float sh_metric[MAX_J];
    if (threadIdx.x < MAX_J) sh_metric[threadIdx.x] = 0.0f;
    __syncthreads();
    const long row0 = static_cast<long>(blockIdx.x) * LOSS_ROWS_PER_BLOCK;
    const int items = LOSS_ROWS_PER_BLOCK * J;
    float loss = 0.0f;
    const int O = 3 * J;
    const float g_mse = 2.0f * inv_count / O, g_l1 = inv_count / O, g_mp = inv_count / J;
    for (int it = threadIdx.x; it < items; it += LOSS_THREADS) {
        const int r = it / J, j = it - r * J;
        const long row = row0 + r;
        if (row >= rows) break;
        const long o = row * O + 3 * j;
        const float dx = pred[o] - target[o], dy = pred[o + 1] - target[o + 1], dz = pred[o + 2] - target[o + 2];
        float gx, gy, gz;
        if (kind == LOSS_MSE) {
            loss += dx * dx + dy * dy + dz * dz;
            gx = g_mse * dx; gy = g_mse * dy; gz = g_mse * dz;
        } else if (kind == LOSS_L1) {
            loss += fabsf(dx) + fabsf(dy) + fabsf(dz);
            gx = dx > 0 ? g_l1 : (dx < 0 ? -g_l1 : 0.0f);
            gy = dy > 0 ? g_l1 : (dy < 0 ? -g_l1 : 0.0f);
            gz = dz > 0 ? g_l1 : (dz < 0 ? -g_l1 : 0.0f);
        } else {
            const float nrm = sqrtf(dx * dx + dy * dy + dz * dz);
            loss += nrm;
            const float s = nrm > 0.0f ? g_mp / nrm : 0.0f;            // torch.norm backward is 0 at 0
            gx = s * dx; gy = s * dy; gz = s * dz;
        }
        if (dpred != nullptr) { dpred[o] = gx; dpred[o + 1] = gy; dpred[o + 2] = gz; }
        float mx = dx, my = dy, mz = dz;
        if (stdv != nullptr) { mx *= stdv[3 * j]; my *= stdv[3 * j + 1]; mz *= stdv[3 * j + 2]; }
        atomicAdd(&sh_metric[j], sqrtf(mx * mx + my * my + mz * mz));   // shared-memory atomics, <= 64 rows deep
    }
    loss = warp_sum(loss);
    if ((threadIdx.x & 31) == 0) sh_loss[threadIdx.x >> 5] = loss;
    __syncthreads();
    if (threadIdx.x == 0) {
        float t = 0.0f;
        for (int w = 0; w < LOSS_THREADS / 32; ++w) t += sh_loss[w];
        partials[static_cast<long>(blockIdx.x) * (1 + J)] = t;
    }
    if (threadIdx.x < J) partials[static_cast<long>(blockIdx.x) * (1 + J) + 1 + threadIdx.x] = sh_metric[threadIdx.x];
}

// single block: fixed-order reduction of the per-block partials
__global__ void loss_finalize_kernel(const float* __restrict__ partials, int nblk, int J, float loss_scale,
                                     float* __restrict__ out, int accumulate_metric) {
    griddep_launch();
    griddep_wait();
    __shared__ double sh[32][MAX_J + 1];
    const int c = threadIdx.x % 32, lane_r = threadIdx.x / 32;          // 32 columns x 32 row lanes
    double s = 0.0;
    if (c < 1 + J)
        for (int b = lane_r; b < nblk; b += 32) s += static_cast<double>(partials[static_cast<long>(b) * (1 + J) + c]);
    if (c < 1 + J) sh[lane_r][c] = s;
    __syncthreads();
    if (lane_r == 0 && c < 1 + J) {
        double t = 0.0;
        for (int r = 0; r < 32; ++r) t += sh[r][c];
        if (c == 0) out[0] = static_cast<float>(t * loss_scale);
        else if (accumulate_metric) out[c] += static_cast<float>(t);
        else out[c] = static_cast<float>(t);
    }
}

constexpr int L1_THREADS = 256;
constexpr int L1_PER_THREAD = 8;

__global__ void __launch_bounds__(L1_THREADS) l1_sum_kernel(const float* __restrict__ a, const float* __restrict__ b,
                                                            long n, int row_len, int centre,
                                                            float* __restrict__ partials) {
    griddep_launch();
    griddep_wait();
    __shared__ float sh[L1_THREADS / 32];
    const long base = (static_cast<long>(blockIdx.x) * L1_THREADS + threadIdx.x) * L1_PER_THREAD;
    float s = 0.0f;
#pragma unroll
    for (int e = 0; e < L1_PER_THREAD; ++e) {
        const long i = base + e;
        if (i < n) {
            float d = a[i] - b[i];
            if (centre && i >= row_len) { const long c = i % row_len; d -= a[c] - b[c]; }
            s += fabsf(d);
        }
    }
    s = warp_sum(s);
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
        float t = 0.0f;
        for (int w = 0; w < L1_THREADS / 32; ++w) t += sh[w];
        partials[blockIdx.x] = t;
    }
}

__global__ void l1_finalize_kernel(const float* __restrict__ partials, int nblk, float scale, float* __restrict__ out) {
    griddep_launch();
    griddep_wait();
    __shared__ double sh[256];
    double s = 0.0;
    for (int b = threadIdx.x; b < nblk; b += 256) s += static_cast<double>(partials[b]);
    sh[threadIdx.x] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
        double t = 0.0;
        for (int i = 0; i < 256; ++i) t += sh[i];
        out[0] = static_cast<float>(t * scale);
    }
}

__global__ void l1_grad_kernel(const float* __restrict__ a, const float* __restrict__ b, long n, int row_len,
                               int centre, const float* __restrict__ gscale, float scale, float* __restrict__ da,
                               float* __restrict__ db) {
    griddep_launch();
    griddep_wait();
    const long i = static_cast<long>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float gs = (gscale != nullptr ? gscale[0] : 1.0f) * scale;
    float g;
    if (!centre || i >= row_len) {
        float d = a[i] - b[i];
        if (centre) { const long c = i % row_len; d -= a[c] - b[c]; }
        g = d > 0 ? gs : (d < 0 ? -gs : 0.0f);
    } else {                                   // element of row 0: own term minus every other row's gradient
        const float d0 = a[i] - b[i];
        g = d0 > 0 ? gs : (d0 < 0 ? -gs : 0.0f);
        const long rows = n / row_len;
        for (long r = 1; r < rows; ++r) {
            const float d = (a[r * row_len + i] - b[r * row_len + i]) - d0;
            g -= d > 0 ? gs : (d < 0 ? -gs : 0.0f);
        }
    }
    da[i] = g;
    if (db != nullptr) db[i] = -g;
}

__global__ void __launch_bounds__(256) adam_kernel(AdamArgs a, float step_size, float inv_sqrt_bc2, float decay) {
    griddep_launch();
    griddep_wait();
    if (a.dyn != nullptr) { step_size = a.dyn[3]; inv_sqrt_bc2 = a.dyn[4]; decay = a.dyn[5]; }
    const long n4 = a.n >> 2, b4 = a.base >> 2;
    const long stride = static_cast<long>(gridDim.x) * blockDim.x;
    uint16_t* ops = static_cast<uint16_t*>(a.operands);
    for (long i = b4 + static_cast<long>(blockIdx.x) * blockDim.x + threadIdx.x; i < b4 + n4; i += stride) {
        float4 p = reinterpret_cast<float4*>(a.p)[i];
        const float4 g4 = reinterpret_cast<const float4*>(a.g)[i];
        float4 m = reinterpret_cast<float4*>(a.m)[i];
        float4 v = reinterpret_cast<float4*>(a.v)[i];
        float* pp = &p.x; const float* gp = &g4.x; float* mp = &m.x; float* vp = &v.x;
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            float g = gp[e] * a.grad_scale;
            if (a.decoupled) pp[e] *= decay;                            // AdamW: p *= 1 - lr * wd
            else if (a.weight_decay != 0.0f) g = fmaf(a.weight_decay, pp[e], g);
            mp[e] = a.beta1 * mp[e] + (1.0f - a.beta1) * g;
            vp[e] = a.beta2 * vp[e] + (1.0f - a.beta2) * g * g;
            const float denom = sqrtf(vp[e]) * inv_sqrt_bc2 + a.eps;
            pp[e] -= step_size * (mp[e] / denom);
        }
        reinterpret_cast<float4*>(a.p)[i] = p;
        reinterpret_cast<float4*>(a.m)[i] = m;
        reinterpret_cast<float4*>(a.v)[i] = v;
        if (ops != nullptr) {                                           // 16-bit operand copy of a Linear weight
            const long idx = i * 4;
            for (int t = 0; t < a.nw; ++t) {
                const long e0 = idx - a.w_p_off[t];
                if (e0 < 0 || e0 >= a.w_numel[t]) continue;
                if (a.w_in[t] == a.w_pad[t]) {                          // dense rows: same linear offset
                    *reinterpret_cast<uint2*>(ops + a.w_off[t] + e0) =
                        make_uint2(pack16(p.x, p.y, a.op_fmt), pack16(p.z, p.w, a.op_fmt));
                } else {                                                // padded rows (the i_dim-wide first layer)
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        const long el = e0 + e;
                        const long r = el / a.w_in[t];
                        const int c = static_cast<int>(el - r * a.w_in[t]);
                        ops[a.w_off[t] + r * a.w_pad[t] + c] = cvt16(pp[e], a.op_fmt);
                    }
                }
                break;
            }
        }
    }
}

__device__ __forceinline__ void st_release_sys_u32(unsigned* p, unsigned v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned ld_acquire_sys_u32(const unsigned* p) {
    unsigned v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void wait_epoch(const unsigned* f, unsigned epoch, long long spin_limit) {
    long long t0 = 0;
    unsigned spins = 0;
    while (static_cast<int>(ld_acquire_sys_u32(f) - epoch) < 0) {         // a DEAD peer traps after LIFTER_DP_TIMEOUT_S
        if ((++spins & 0xFF) == 0 && spin_limit > 0) {
            const long long now = clock64();
            if (t0 == 0) t0 = now;
            else if (now - t0 > spin_limit) __trap();
        }
    }
}
__device__ __forceinline__ float4 ld_peer_f4(const float* p) {
    float4 r;
    asm volatile("ld.relaxed.sys.global.v4.f32 {%0, %1, %2, %3}, [%4];"
                 : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p) : "memory");
    return r;
}

// NVSwitch multicast (multimem.*, sm_90+): one load that returns the SUM of the addressed float4 over every rank's copy
// (reduced inside the switch), one store that lands in every rank's copy.  Weak operations: the flag handshakes around
// the kernel order them against the peers' kernels.
__device__ __forceinline__ float4 multimem_ld_reduce_f4(const float* mc) {
    float4 r;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
                 : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(mc) : "memory");
    return r;
}
__device__ __forceinline__ void multimem_st_f4(float* mc, const float4& v) {
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(mc), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w)
                 : "memory");
}
__device__ __forceinline__ void multimem_st_b64(void* mc, uint32_t a, uint32_t b) {     // two 32-bit words of packed 16-bit data
    asm volatile("multimem.st.relaxed.sys.global.v2.f32 [%0], {%1, %2};" ::"l"(mc), "f"(__uint_as_float(a)), "f"(__uint_as_float(b))
                 : "memory");
}

constexpr int DP_UNROLL = 2;
__global__ void __launch_bounds__(256) adam_dp_kernel(AdamDpArgs d, float step_size, float inv_sqrt_bc2, float decay) {
    griddep_launch();
    griddep_wait();
    const AdamArgs& a = d.a;
    if (a.dyn != nullptr) { step_size = a.dyn[3]; inv_sqrt_bc2 = a.dyn[4]; decay = a.dyn[5]; }
    const unsigned epoch = *d.epoch * static_cast<unsigned>(ADAM_DP_PHASES) + static_cast<unsigned>(d.phase);
    unsigned* my_flags = d.flag_peer[d.rank];
    // ---- barrier 1: every rank's gradient is complete (my predecessor kernels finished: this kernel runs)
    // (d.external_flags: a one-block dp_flag_kernel in front of this launch did the handshake -- no block of this grid spins)
    if (!d.external_flags && threadIdx.x < d.world) {
        if (blockIdx.x == 0) st_release_sys_u32(d.flag_peer[threadIdx.x] + d.rank, epoch);
        wait_epoch(my_flags + threadIdx.x, epoch, d.spin_limit);
    }
    __syncthreads();
    // ---- my slice: sum of the gradients in rank order, Adam, parameters to every rank
    const long n4 = a.n >> 2, b4 = a.base >> 2;
    const long slice = (n4 + d.world - 1) / d.world;
    const long lo = b4 + slice * d.rank, hi = min(lo + slice, b4 + n4);
    const long stride = static_cast<long>(gridDim.x) * blockDim.x;
    for (long i0 = lo + static_cast<long>(blockIdx.x) * blockDim.x + threadIdx.x; i0 < hi; i0 += stride * DP_UNROLL) {
        float4 gs[DP_UNROLL][ADAM_DP_MAX_WORLD];
#pragma unroll
        for (int u = 0; u < DP_UNROLL; ++u) {
            const long i = i0 + u * stride;
            if (d.mc_g != nullptr) {                                          // reduced in the switch: one load
                if (i < hi) gs[u][0] = multimem_ld_reduce_f4(d.mc_g + i * 4);
            } else {
#pragma unroll
                for (int q = 0; q < ADAM_DP_MAX_WORLD; ++q)
                    if (q < d.world && i < hi) gs[u][q] = ld_peer_f4(d.g_peer[q] + i * 4);
            }
        }
#pragma unroll
        for (int u = 0; u < DP_UNROLL; ++u) {
            const long i = i0 + u * stride;
            if (i >= hi) break;
            float4 g4 = gs[u][0];
            if (d.mc_g == nullptr) {
#pragma unroll
                for (int q = 1; q < ADAM_DP_MAX_WORLD; ++q)
                    if (q < d.world) { g4.x += gs[u][q].x; g4.y += gs[u][q].y; g4.z += gs[u][q].z; g4.w += gs[u][q].w; }
            }
            float4 p = reinterpret_cast<float4*>(a.p)[i];
            float4 m = reinterpret_cast<float4*>(a.m)[i];
            float4 v = reinterpret_cast<float4*>(a.v)[i];
            float* pp = &p.x; const float* gp = &g4.x; float* mp = &m.x; float* vp = &v.x;
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                float g = gp[e] * a.grad_scale;
                if (a.decoupled) pp[e] *= decay;
                else if (a.weight_decay != 0.0f) g = fmaf(a.weight_decay, pp[e], g);
                mp[e] = a.beta1 * mp[e] + (1.0f - a.beta1) * g;
                vp[e] = a.beta2 * vp[e] + (1.0f - a.beta2) * g * g;
                const float denom = sqrtf(vp[e]) * inv_sqrt_bc2 + a.eps;
                pp[e] -= step_size * (mp[e] / denom);
            }
            reinterpret_cast<float4*>(a.m)[i] = m;
            reinterpret_cast<float4*>(a.v)[i] = v;
            if (d.mc_p != nullptr) multimem_st_f4(d.mc_p + i * 4, p);
            else for (int q = 0; q < d.world; ++q) reinterpret_cast<float4*>(d.p_peer[q])[i] = p;
            if (a.operands != nullptr) {
                const long idx = i * 4;
                for (int t = 0; t < a.nw; ++t) {
                    const long e0 = idx - a.w_p_off[t];
                    if (e0 < 0 || e0 >= a.w_numel[t]) continue;
                    if (a.w_in[t] == a.w_pad[t]) {
                        const uint2 w2 = make_uint2(pack16(p.x, p.y, a.op_fmt), pack16(p.z, p.w, a.op_fmt));
                        if (d.mc_op != nullptr)
                            multimem_st_b64(static_cast<uint16_t*>(d.mc_op) + a.w_off[t] + e0, w2.x, w2.y);
                        else
                            for (int q = 0; q < d.world; ++q)
                                *reinterpret_cast<uint2*>(static_cast<uint16_t*>(d.op_peer[q]) + a.w_off[t] + e0) = w2;
                    } else {
#pragma unroll
                        for (int e = 0; e < 4; ++e) {
                            const long el = e0 + e;
                            const long r = el / a.w_in[t];
                            const int c = static_cast<int>(el - r * a.w_in[t]);
                            const uint16_t h = cvt16(pp[e], a.op_fmt);
                            for (int q = 0; q < d.world; ++q)
                                static_cast<uint16_t*>(d.op_peer[q])[a.w_off[t] + r * a.w_pad[t] + c] = h;
                        }
                    }
                    break;
                }
            }
        }
    }
    // ---- barrier 2: the last block of this rank tells every peer "my stores are done" and waits for theirs
    if (!d.last || d.external_flags) return;   // an earlier range of the step / the closing handshake is its own launch
    __threadfence_system();
    __syncthreads();
    __shared__ unsigned last;
    unsigned* counter = my_flags + 2 * d.world;
    if (threadIdx.x == 0) last = (atomicAdd(counter, 1u) == gridDim.x - 1) ? 1u : 0u;
    __syncthreads();
    if (last) {
        if (threadIdx.x == 0) *counter = 0;                                // ready for the next step
        if (threadIdx.x < d.world) {
            st_release_sys_u32(d.flag_peer[threadIdx.x] + d.world + d.rank, epoch);
            wait_epoch(my_flags + d.world + threadIdx.x, epoch, d.spin_limit);
        }
    }
}

// One-block handshake of the data-parallel optimiser: rank r raises flag[slot + r] on every peer and waits until every
// peer has raised its flag here.  Runs as its own launch in front of (slot 0: "every rank's gradient of this range is
// complete") and behind (slot world: "every rank's parameter stores of this step have landed") the range kernels, so
// that the stream order carries the dependency and no compute block ever spins on a system-scope flag beside the
// backward GEMMs (with the handshake inside the range kernel, each of its ~590 blocks polled until the peer arrived).
__global__ void dp_flag_kernel(AdamDpArgs d, int slot) {
    griddep_launch();
    griddep_wait();
    const unsigned epoch = *d.epoch * static_cast<unsigned>(ADAM_DP_PHASES) + static_cast<unsigned>(d.phase);
    if (threadIdx.x < d.world) {
        __threadfence_system();
        st_release_sys_u32(d.flag_peer[threadIdx.x] + slot + d.rank, epoch);
        wait_epoch(d.flag_peer[d.rank] + slot + threadIdx.x, epoch, d.spin_limit);
    }
}

// state: [0] u32 dropout_step, [1] i32 adam_step, [2] f32 lr, [3] step_size, [4] inv_sqrt_bc2, [5] decay
__global__ void step_advance_kernel(unsigned int* state, float beta1, float beta2, float weight_decay) {
    griddep_launch();
    griddep_wait();
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    const int t = static_cast<int>(state[1]) + 1;
    state[0] = static_cast<unsigned int>(t - 1);                       // dropout stream of this step
    state[1] = static_cast<unsigned int>(t);
    const float lr = __uint_as_float(state[2]);
    const double bc1 = 1.0 - pow(static_cast<double>(beta1), static_cast<double>(t));
    const double bc2 = 1.0 - pow(static_cast<double>(beta2), static_cast<double>(t));
    state[3] = __float_as_uint(static_cast<float>(static_cast<double>(lr) / bc1));
    state[4] = __float_as_uint(static_cast<float>(1.0 / sqrt(bc2)));
    state[5] = __float_as_uint(1.0f - lr * weight_decay);
}

}  // namespace

int loss_partial_rows(int64_t rows) { return static_cast<int>((rows + LOSS_ROWS_PER_BLOCK - 1) / LOSS_ROWS_PER_BLOCK); }

int fused_loss(const float* pred, const float* target, int64_t rows, int J, int kind, double count, const float* stdv,
               float* dpred, float* partials, float* out, int accumulate_metric, cudaStream_t st) {
    LIFTER_REQUIRE(J >= 1 && J <= MAX_J, "fused_loss: %d joints unsupported (1..%d)", J, MAX_J);
    LIFTER_REQUIRE(kind >= LOSS_MSE && kind <= LOSS_MPJPE, "fused_loss: unknown loss kind %d", kind);
    LIFTER_REQUIRE(rows >= 1 && count >= 1.0, "fused_loss: empty batch");
    const int nblk = loss_partial_rows(rows);
    const float inv_count = static_cast<float>(1.0 / count);
    LIFTER_CHECK_CUDA(launch_pdl("fused_loss_kernel", fused_loss_kernel, dim3(nblk), dim3(LOSS_THREADS), 0, st, pred, target, rows, J, kind, inv_count, stdv, dpred, partials));
    const double denom = kind == LOSS_MPJPE ? count * J : count * 3.0 * J;
    LIFTER_CHECK_CUDA(launch_pdl("loss_finalize_kernel", loss_finalize_kernel, dim3(1), dim3(1024), 0, st, partials, nblk, J, static_cast<float>(1.0 / denom), out, accumulate_metric));
    return LIFTER_OK;
}

int l1_sum(const float* a, const float* b, int64_t n, int row_len, int centre, float scale, float* partials, float* out,
           cudaStream_t st) {
    LIFTER_REQUIRE(n >= 1 && row_len >= 1 && n % row_len == 0, "l1_sum: bad shape");
    const int nblk = static_cast<int>((n + L1_THREADS * L1_PER_THREAD - 1) / (L1_THREADS * L1_PER_THREAD));
    LIFTER_CHECK_CUDA(launch_pdl("l1_sum_kernel", l1_sum_kernel, dim3(nblk), dim3(L1_THREADS), 0, st, a, b, n, row_len, centre, partials));
    LIFTER_CHECK_CUDA(launch_pdl("l1_finalize_kernel", l1_finalize_kernel, dim3(1), dim3(256), 0, st, partials, nblk, scale, out));
    return LIFTER_OK;
}

int l1_grad(const float* a, const float* b, int64_t n, int row_len, int centre, const float* gscale, float scale,
            float* da, float* db, cudaStream_t st) {
    LIFTER_REQUIRE(n >= 1 && row_len >= 1 && n % row_len == 0, "l1_grad: bad shape");
    LIFTER_CHECK_CUDA(launch_pdl("l1_grad_kernel", l1_grad_kernel, dim3(static_cast<unsigned>((n + 255) / 256)), dim3(256), 0, st, a, b, n, row_len, centre, gscale, scale, da, db));
    return LIFTER_OK;
}

int adam_step(const AdamArgs& a, cudaStream_t st) {
    LIFTER_REQUIRE(a.n % 4 == 0 && a.base % 4 == 0 && a.n > 0, "adam: range start / length must be multiples of 4");
    LIFTER_REQUIRE(a.step >= 1 || a.dyn != nullptr, "adam: step is 1-based");
    LIFTER_REQUIRE(a.operands == nullptr || (a.nw >= 1 && a.nw <= ADAM_MAX_WEIGHTS), "adam: bad operand table");
    const int t = a.step >= 1 ? a.step : 1;
    const double bc1 = 1.0 - pow(static_cast<double>(a.beta1), t);
    const double bc2 = 1.0 - pow(static_cast<double>(a.beta2), t);
    const float step_size = static_cast<float>(a.lr / bc1);
    const float inv_sqrt_bc2 = static_cast<float>(1.0 / sqrt(bc2));
    const float decay = 1.0f - a.lr * a.weight_decay;
    const long n4 = a.n / 4;
    long blocks = (n4 + 255) / 256;
    const long cap = static_cast<long>(num_sms()) * 8;             // grid = multiple of the SM count, grid-stride
    if (blocks > cap) blocks = cap;
    LIFTER_CHECK_CUDA(launch_pdl("adam_kernel", adam_kernel, dim3(static_cast<unsigned>(blocks)), dim3(256), 0, st, a, step_size, inv_sqrt_bc2, decay));
    return LIFTER_OK;
}

int adam_step_dp(const AdamDpArgs& d_in, cudaStream_t st) {
    AdamDpArgs d = d_in;
    d.spin_limit = dp_spin_limit_clocks();
    const AdamArgs& a = d.a;
    LIFTER_REQUIRE(a.n % 4 == 0 && a.base % 4 == 0 && a.n > 0, "adam: range start / length must be multiples of 4");
    LIFTER_REQUIRE(d.phase >= 0 && d.phase < ADAM_DP_PHASES, "adam_dp: phase %d out of range", d.phase);
    LIFTER_REQUIRE(d.world >= 2 && d.world <= ADAM_DP_MAX_WORLD && d.rank >= 0 && d.rank < d.world, "adam_dp: bad world/rank");
    LIFTER_REQUIRE(a.dyn != nullptr && d.epoch != nullptr, "adam_dp: needs the device step state");
    const long slice4 = (a.n / 4 + d.world - 1) / d.world;
    long blocks = (slice4 + 256 * DP_UNROLL - 1) / (256 * DP_UNROLL);
    // LIFTER_DP_INKERNEL_FLAGS=1: the round-1 form (handshakes inside the range kernel, all blocks co-resident)
    static const bool inkernel = [] { const char* e = getenv("LIFTER_DP_INKERNEL_FLAGS"); return e != nullptr && e[0] == '1'; }();
    d.external_flags = inkernel ? 0 : 1;
    const long cap = static_cast<long>(num_sms()) * (inkernel ? 4 : 2);
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    if (!inkernel) LIFTER_CHECK_CUDA(launch_pdl("dp_flag_kernel", dp_flag_kernel, dim3(1), dim3(32), 0, st, d, 0));
    LIFTER_CHECK_CUDA(launch_pdl("adam_dp_kernel", adam_dp_kernel, dim3(static_cast<unsigned>(blocks)), dim3(256), 0, st, d, 0.0f, 1.0f, 1.0f));
    if (!inkernel && d.last) LIFTER_CHECK_CUDA(launch_pdl("dp_flag_kernel", dp_flag_kernel, dim3(1), dim3(32), 0, st, d, d.world));
    return LIFTER_OK;
}

int step_advance(void* state, float beta1, float beta2, float weight_decay, cudaStream_t st) {
    LIFTER_REQUIRE(state != nullptr, "step_advance: null state");
    LIFTER_CHECK_CUDA(launch_pdl("step_advance_kernel", step_advance_kernel, dim3(1), dim3(32), 0, st,
                                 static_cast<unsigned int*>(state), beta1, beta2, weight_decay));
    return LIFTER_OK;
}

// ---------------------------------------------------------------------------------- no-allocation variant of the loss
// lifter_loss_static (include/lifter_b200.h): the fused loss with its per-block partial sums in a file-scope device array
// instead of caller-provided scratch -- for callers that cannot allocate (CUDA-graph capture, serving loops, harnesses
// that forbid cudaMalloc).  Every GPU that loads the library holds the array (88 KB) until the process exits.
constexpr long LOSS_STATIC_MAX_ROWS = LIFTER_LOSS_STATIC_MAX_ROWS;                       // 65,536 rows per call
constexpr long LOSS_STATIC_FLOATS = (LOSS_STATIC_MAX_ROWS / LOSS_ROWS_PER_BLOCK) * (1 + MAX_J);   // 1,024 blocks x 22
static __device__ __align__(256) float g_loss_static_scratch[LOSS_STATIC_FLOATS];

// In host code `g_loss_static_scratch` is a host-side stand-in, not a device address: the kernels get the address that
// cudaGetSymbolAddress reports for the CURRENT device, looked up once per device and cached.
static int loss_static_scratch(float** out) {
    static float* per_device[64] = {};
    int dev = 0;
    LIFTER_CHECK_CUDA(cudaGetDevice(&dev));
    LIFTER_REQUIRE(dev >= 0 && dev < 64, "loss_static: device index %d out of range", dev);
    if (per_device[dev] == nullptr) {
        void* ptr = nullptr;
        LIFTER_CHECK_CUDA(cudaGetSymbolAddress(&ptr, g_loss_static_scratch));
        LIFTER_REQUIRE(ptr != nullptr, "loss_static: cudaGetSymbolAddress returned a null address");
        per_device[dev] = static_cast<float*>(ptr);
    }
    *out = per_device[dev];
    return LIFTER_OK;
}

}  // namespace lifter

extern "C" int lifter_loss_static(const float* pred, const float* target, int64_t rows, int joints, int kind,
                                  int64_t global_rows, const float* stdv, float* dpred, float* out,
                                  int accumulate_metric, void* stream) {
    using namespace lifter;
    LIFTER_REQUIRE(pred && target && out, "loss_static: null pointer");
    LIFTER_REQUIRE(rows >= 1 && rows <= LOSS_STATIC_MAX_ROWS,
                   "loss_static: %lld rows exceed the static scratch (at most %ld rows per call; use lifter_loss with "
                   "caller-provided scratch for larger batches)", static_cast<long long>(rows), LOSS_STATIC_MAX_ROWS);
    LIFTER_REQUIRE(joints >= 1 && joints <= MAX_J, "loss_static: %d joints unsupported (1..%d)", joints, MAX_J);
    float* scratch = nullptr;
    LIFTER_PROPAGATE(loss_static_scratch(&scratch));
    // nothing in the scratch has to start at zero: block b of the loss kernel overwrites all 1 + joints floats of its
    // partial row and the finalize kernel reads only the rows this call wrote, so the last call's contents are inert
    return fused_loss(pred, target, rows, joints, kind, static_cast<double>(global_rows), stdv, dpred, scratch, out,
                      accumulate_metric, static_cast<cudaStream_t>(stream));
}
