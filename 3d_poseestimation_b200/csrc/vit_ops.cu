// MyViT lifter, eval-mode forward (phase1_lifting/baselineModel.py:228-254, 286-362).
// Tokens = joints: a pose is 17 tokens of width `hidden` (256).  All Linear layers over the flattened [poses * 17][.]
// token matrix run on the tcgen05 GEMM of gemm_sm100.cu (bias, ReLU / GELU and the residual add in its epilogue); the
// kernels here are what sits between the GEMMs:
//   vit_embed_kernel      Linear(2 -> hidden) + positional embedding -> residual stream (16-bit)            [HBM]
//   vit_layernorm_kernel  one or two LayerNorms in a row (block.norm1 then mhsa.norm; block.norm2)          [HBM]
//   vit_attention_kernel  softmax(q k^T / sqrt(d_head)) v for the 17 x 17 token pairs of one (pose, head)   [FP32 pipe]
#include "vit_ops.cuh"

#include "elementwise.cuh"
#include "gemm_sm100.cuh"
#include "vit_common.cuh"

namespace lifter {

namespace {

// baselineModel.py:349-354: tokens = linear_mapper(x) + pos_embed -> h (16-bit residual stream), and in the same pass the
// first block's two LayerNorms (norm1, mhsa.norm; :303, :244) taken from the fp32 values -> u (g1 == nullptr: h only).
// x [T][in_d] fp32, w [Hd][in_d], pos [tokens][Hd].  IND = in_d when the lane's weight slice is loaded as float4s
// (in_d == 2, the reference's input), 0 = any in_d <= 4 with scalar loads.
template <int NV, int IND>
__global__ void __launch_bounds__(256) vit_embed_kernel(const float* __restrict__ x, long T, int tokens, int in_d,
                                                        const float* __restrict__ w, const float* __restrict__ bias,
                                                        const float* __restrict__ pos, void* __restrict__ h,
                                                        const float* __restrict__ g1, const float* __restrict__ b1,
                                                        const float* __restrict__ g2, const float* __restrict__ b2, float eps,
                                                        void* __restrict__ u, int fmt, float* __restrict__ h32) {
    griddep_launch();
    griddep_wait();
    const int lane = threadIdx.x & 31;
    const long t0 = (static_cast<long>(blockIdx.x) * (blockDim.x >> 5) + (threadIdx.x >> 5)) * VIT_RPW;
    if (t0 >= T) return;
    RowParams<NV> ln1, ln2;
    if (g1 != nullptr) { load_row_params<NV>(ln1, g1, b1, lane); load_row_params<NV>(ln2, g2, b2, lane); }
    float bs[NV][8], w0[NV][8], w1[NV][8];                            // IND == 2: bias and the two weight columns
#pragma unroll
    for (int k = 0; k < NV; ++k) {
        const int c0 = k * 256 + lane * 8;
        const float4 q0 = *reinterpret_cast<const float4*>(bias + c0), q1 = *reinterpret_cast<const float4*>(bias + c0 + 4);
        bs[k][0] = q0.x; bs[k][1] = q0.y; bs[k][2] = q0.z; bs[k][3] = q0.w; bs[k][4] = q1.x; bs[k][5] = q1.y; bs[k][6] = q1.z; bs[k][7] = q1.w;
        if constexpr (IND == 2) {
            const float4* wv = reinterpret_cast<const float4*>(w + c0 * 2);
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                const float4 ww = wv[e];                              // columns c0 + 2e, c0 + 2e + 1
                w0[k][2 * e] = ww.x; w1[k][2 * e] = ww.y; w0[k][2 * e + 1] = ww.z; w1[k][2 * e + 1] = ww.w;
            }
        }
    }
    const long t1 = min(T, t0 + VIT_RPW);
    int n = static_cast<int>(t0 % tokens);
    for (long t = t0; t < t1; ++t) {
        float v[NV][8];
#pragma unroll
        for (int k = 0; k < NV; ++k) {
            const int c0 = k * 256 + lane * 8;
            const float4 p0 = *reinterpret_cast<const float4*>(pos + n * (NV * 256) + c0);
            const float4 p1 = *reinterpret_cast<const float4*>(pos + n * (NV * 256) + c0 + 4);
            const float pp[8] = {p0.x, p0.y, p0.z, p0.w, p1.x, p1.y, p1.z, p1.w};
            if constexpr (IND == 2) {
                const float2 xi = *reinterpret_cast<const float2*>(x + t * 2);
#pragma unroll
                for (int e = 0; e < 8; ++e) v[k][e] = fmaf(xi.y, w1[k][e], fmaf(xi.x, w0[k][e], bs[k][e] + pp[e]));
            } else {
#pragma unroll
                for (int e = 0; e < 8; ++e) {
                    float acc = bs[k][e] + pp[e];
                    for (int i = 0; i < in_d; ++i) acc = fmaf(x[t * in_d + i], w[(c0 + e) * in_d + i], acc);
                    v[k][e] = acc;
                }
            }
        }
        if (h32 != nullptr) store_row32<NV>(h32, t, lane, v);      // fp32 residual stream (eval path)
        else store_row16<NV>(h, t, lane, v, fmt);
        if (g1 != nullptr) {
            warp_layernorm<NV>(v, ln1, eps);
            warp_layernorm<NV>(v, ln2, eps);
            store_row16<NV>(u, t, lane, v, fmt);
        }
        if (++n == tokens) n = 0;
    }
}

// out = LN2(LN1(in)) (g2 == nullptr: one LayerNorm).  in / out 16-bit [T][Hd]; statistics and affine in fp32, no
// 16-bit rounding between the two norms.
constexpr int LN_DEPTH = 4;                          // input rows a warp keeps in flight (cp.async ring, private to the warp)
template <int NV, bool IN32 = false>
__global__ void __launch_bounds__(256) vit_layernorm_kernel(const void* __restrict__ in, long T, long rows_per_warp,
                                                            const float* __restrict__ g1, const float* __restrict__ b1,
                                                            const float* __restrict__ g2, const float* __restrict__ b2, float eps,
                                                            void* __restrict__ out, int fmt) {
    griddep_launch();
    griddep_wait();
    extern __shared__ __align__(16) unsigned char ln_smem[];
    // IN32: the input rows are the fp32 residual stream (eval path): two 16-byte vectors per lane and 256 columns
    constexpr int VPR = IN32 ? 2 : 1;                                 // uint4 vectors per lane per 256 columns
    constexpr int ROWB = NV * 512 * VPR;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const long t0 = (static_cast<long>(blockIdx.x) * 8 + wid) * rows_per_warp;
    if (t0 >= T) return;
    const long t1 = min(T, t0 + rows_per_warp);
    unsigned char* ring = ln_smem + wid * (LN_DEPTH * ROWB) + lane * (16 * VPR);
    const uint32_t ring_addr = smem_u32(ring);
    auto issue = [&](long row, int slot) {            // lane l copies, and later reads, the same bytes: no barrier needed
        if (row < t1) {
#pragma unroll
            for (int k = 0; k < NV; ++k)
#pragma unroll
                for (int q = 0; q < VPR; ++q)
                    cp_async16(ring_addr + slot * ROWB + k * 512 * VPR + q * 16,
                               static_cast<const char*>(in) + row * static_cast<long>(ROWB) + k * 512 * VPR + lane * (16 * VPR) + q * 16);
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    };
#pragma unroll
    for (int k = 0; k < LN_DEPTH; ++k) issue(t0 + k, k);
    RowParams<NV> ln1, ln2;
    load_row_params<NV>(ln1, g1, b1, lane);
    if (g2 != nullptr) load_row_params<NV>(ln2, g2, b2, lane);
    int slot = 0;
    for (long t = t0; t < t1; ++t) {
        asm volatile("cp.async.wait_group %0;" :: "n"(LN_DEPTH - 1) : "memory");
        float v[NV][8];
#pragma unroll
        for (int k = 0; k < NV; ++k) {
            const unsigned char* src = ring + slot * ROWB + k * 512 * VPR;
            if constexpr (IN32) {
                const uint4 a = *reinterpret_cast<const uint4*>(src), b = *reinterpret_cast<const uint4*>(src + 16);
                v[k][0] = __uint_as_float(a.x); v[k][1] = __uint_as_float(a.y); v[k][2] = __uint_as_float(a.z); v[k][3] = __uint_as_float(a.w);
                v[k][4] = __uint_as_float(b.x); v[k][5] = __uint_as_float(b.y); v[k][6] = __uint_as_float(b.z); v[k][7] = __uint_as_float(b.w);
            } else {
                const uint4 c = *reinterpret_cast<const uint4*>(src);
                const uint32_t wd[4] = {c.x, c.y, c.z, c.w};
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    const float2 f = unpack16(wd[e], fmt);
                    v[k][2 * e] = f.x;
                    v[k][2 * e + 1] = f.y;
                }
            }
        }
        issue(t + LN_DEPTH, slot);
        slot = slot + 1 == LN_DEPTH ? 0 : slot + 1;
        warp_layernorm<NV>(v, ln1, eps);
        if (g2 != nullptr) warp_layernorm<NV>(v, ln2, eps);
        store_row16<NV>(out, t, lane, v, fmt);
    }
    asm volatile("cp.async.wait_group 0;" ::: "memory");
}

// baselineModel.py:243-253.  qkv 16-bit [T][3 * Hd] = [q | k | v], each (head, d) major; o 16-bit [T][Hd].
// One warp per (pose, head), on the warp-level tensor-core MMA (mma.sync m16n8k16, fp32 accumulate; a 17 x 17 x 64
// problem is far below a tcgen05 tile): the head's q, k, v slices are staged in shared memory with coalesced 16-byte loads
// (rows past the token count zeroed), S = q k^T by ldmatrix + MMA per 16-row tile, softmax in fp32 on the accumulator
// fragments (columns past the token count masked), P re-used in registers as the A operand of P v, the output rows
// normalised and staged through the (now dead) q rows for coalesced 16-byte stores.
constexpr int ATT_WARPS = 4;

__device__ __forceinline__ void ldsm_x4(uint32_t (&r)[4], uint32_t addr) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0, %1, %2, %3}, [%4];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}
__device__ __forceinline__ void ldsm_x4_trans(uint32_t (&r)[4], uint32_t addr) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0, %1, %2, %3}, [%4];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}
template <int FMT>
__device__ __forceinline__ void mma16816(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    if constexpr (FMT == FMT_BF16)
        asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, {%0, %1, %2, %3};"
                     : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3]) : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
    else
        asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, {%0, %1, %2, %3};"
                     : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3]) : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

template <int DH, int FMT>
__global__ void __launch_bounds__(ATT_WARPS * 32) vit_attention_kernel(const void* __restrict__ qkv, long pose_heads,
                                                                       int tokens, int heads, float scale_log2e,
                                                                       void* __restrict__ o) {
    constexpr int fmt = FMT;
    griddep_launch();
    griddep_wait();
    constexpr int LDS = DH + 8;                        // halfs per staged row: 16-byte aligned, conflict-free for ldmatrix
    constexpr int VPR = DH / 8;                        // 16-byte vectors per row slice
    extern __shared__ __align__(16) unsigned char att_smem[];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const long ph = static_cast<long>(blockIdx.x) * ATT_WARPS + wid;
    if (ph >= pose_heads) return;
    const long pose = ph / heads;
    const int head = static_cast<int>(ph - pose * heads);
    const int Hd = heads * DH;
    const int R = tokens + 1;                          // staged rows per slice: the tokens and ONE zero row that every
    const int MT = (tokens + 15) / 16, NT = (tokens + 7) / 8;   // MMA-tile row past the token count aliases (less shared
                                                       // memory per warp = more resident warps for a latency-bound kernel)
    unsigned short* sq = reinterpret_cast<unsigned short*>(att_smem) + static_cast<size_t>(wid) * (3 * R * LDS);
    const unsigned short* src = static_cast<const unsigned short*>(qkv) + pose * tokens * (3L * Hd) + head * DH;
    constexpr int RPP = 32 / VPR;                      // staged rows per warp-wide pass of 16-byte vectors
    const int vec = lane % VPR, rsub = lane / VPR;
#pragma unroll
    for (int slice = 0; slice < 3; ++slice)
        for (int n = rsub; n < R; n += RPP) {
            unsigned short* dst = sq + (slice * R + n) * LDS + vec * 8;
            if (n < tokens) cp_async16(smem_u32(dst), src + n * (3L * Hd) + slice * Hd + vec * 8);
            else *reinterpret_cast<uint4*>(dst) = make_uint4(0u, 0u, 0u, 0u);
        }
    cp_async_wait_all();
    __syncwarp();
    const uint32_t q_addr = smem_u32(sq), k_addr = q_addr + R * LDS * 2, v_addr = k_addr + R * LDS * 2;
    const int g = lane >> 2, t4 = lane & 3;
    for (int mt = 0; mt < MT; ++mt) {
        // ---- S = q k^T for query rows mt * 16 .. + 15 ----------------------------------------------------------
        float sc[4][4];
#pragma unroll
        for (int nt = 0; nt < 4; ++nt)
#pragma unroll
            for (int e = 0; e < 4; ++e) sc[nt][e] = 0.0f;
#pragma unroll
        for (int ks = 0; ks < DH / 16; ++ks) {
            uint32_t a[4];
            ldsm_x4(a, q_addr + (min(mt * 16 + (lane & 7) + ((lane >> 3) & 1) * 8, tokens) * LDS + ks * 16 + (lane >> 4) * 8) * 2);
#pragma unroll
            for (int np = 0; np < 2; ++np) {           // key tiles 2 np, 2 np + 1 (keys np * 16 .. + 15)
                if (np * 2 < NT) {
                    uint32_t bk[4];                    // [0], [1]: tile 2 np (k lo, k hi); [2], [3]: tile 2 np + 1
                    ldsm_x4(bk, k_addr + (min(np * 16 + (lane & 7) + (lane >> 4) * 8, tokens) * LDS + ks * 16 + ((lane >> 3) & 1) * 8) * 2);
                    mma16816<FMT>(sc[np * 2], a, bk[0], bk[1]);
                    if (np * 2 + 1 < NT) mma16816<FMT>(sc[np * 2 + 1], a, bk[2], bk[3]);
                }
            }
        }
        // ---- softmax over the keys, rows g and g + 8 of the tile ----------------------------------------------------
        float mx[2] = {-CUDART_INF_F, -CUDART_INF_F};
#pragma unroll
        for (int nt = 0; nt < 4; ++nt)
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                const int col = nt * 8 + t4 * 2 + (e & 1);
                sc[nt][e] = col < tokens ? sc[nt][e] * scale_log2e : -CUDART_INF_F;
                mx[e >> 1] = fmaxf(mx[e >> 1], sc[nt][e]);
            }
        float den[2] = {0.0f, 0.0f};
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            mx[h] = fmaxf(mx[h], __shfl_xor_sync(0xffffffffu, mx[h], 1));
            mx[h] = fmaxf(mx[h], __shfl_xor_sync(0xffffffffu, mx[h], 2));
        }
#pragma unroll
        for (int nt = 0; nt < 4; ++nt)
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                sc[nt][e] = exp2f(sc[nt][e] - mx[e >> 1]);          // masked columns: exp2(-inf) = 0
                den[e >> 1] += sc[nt][e];
            }
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            den[h] += __shfl_xor_sync(0xffffffffu, den[h], 1);
            den[h] += __shfl_xor_sync(0xffffffffu, den[h], 2);
            den[h] = 1.0f / den[h];
        }
        // ---- O = P v ------------------------------------------------------------------------------------------------
        float oc[DH / 8][4];
#pragma unroll
        for (int dn = 0; dn < DH / 8; ++dn)
#pragma unroll
            for (int e = 0; e < 4; ++e) oc[dn][e] = 0.0f;
#pragma unroll
        for (int kt = 0; kt < 2; ++kt) {               // keys kt * 16 .. + 15 = score tiles 2 kt, 2 kt + 1
            if (kt * 2 < NT) {
                uint32_t pa[4];
                pa[0] = pack16(sc[kt * 2][0], sc[kt * 2][1], fmt);
                pa[1] = pack16(sc[kt * 2][2], sc[kt * 2][3], fmt);
                pa[2] = pack16(sc[kt * 2 + 1][0], sc[kt * 2 + 1][1], fmt);
                pa[3] = pack16(sc[kt * 2 + 1][2], sc[kt * 2 + 1][3], fmt);
#pragma unroll
                for (int dp = 0; dp < DH / 16; ++dp) {                 // value column tiles 2 dp, 2 dp + 1
                    uint32_t bv[4];
                    ldsm_x4_trans(bv, v_addr + (min(kt * 16 + (lane & 7) + ((lane >> 3) & 1) * 8, tokens) * LDS + dp * 16 + (lane >> 4) * 8) * 2);
                    mma16816<FMT>(oc[dp * 2], pa, bv[0], bv[1]);
                    mma16816<FMT>(oc[dp * 2 + 1], pa, bv[2], bv[3]);
                }
            }
        }
        // ---- normalise, stage through the q rows of this tile (no longer read) -------------------------------------------
        __syncwarp();
#pragma unroll
        for (int dn = 0; dn < DH / 8; ++dn) {
            if (mt * 16 + g < tokens)
                *reinterpret_cast<uint32_t*>(sq + (mt * 16 + g) * LDS + dn * 8 + t4 * 2) = pack16(oc[dn][0] * den[0], oc[dn][1] * den[0], fmt);
            if (mt * 16 + g + 8 < tokens)
                *reinterpret_cast<uint32_t*>(sq + (mt * 16 + g + 8) * LDS + dn * 8 + t4 * 2) = pack16(oc[dn][2] * den[1], oc[dn][3] * den[1], fmt);
        }
    }
    __syncwarp();
    unsigned short* dst = static_cast<unsigned short*>(o) + pose * tokens * static_cast<long>(Hd) + head * DH;
    for (int n = rsub; n < tokens; n += RPP)
        *reinterpret_cast<uint4*>(dst + n * static_cast<long>(Hd) + vec * 8) = *reinterpret_cast<const uint4*>(sq + n * LDS + vec * 8);
}

size_t att_smem_bytes(int tokens, int dh) {
    return static_cast<size_t>(ATT_WARPS) * 3 * (tokens + 1) * (dh + 8) * 2;
}

}  // namespace

int vit_validate(const VitDesc& d) {
    LIFTER_REQUIRE(d.tokens >= 1 && d.tokens <= VIT_MAX_TOKENS, "vit: %d tokens per pose unsupported (1..%d)", d.tokens, VIT_MAX_TOKENS);
    LIFTER_REQUIRE(d.in_dim >= 1 && d.in_dim <= 4, "vit: %d input values per token unsupported (1..4)", d.in_dim);
    LIFTER_REQUIRE(d.hidden >= 256 && d.hidden <= 1024 && d.hidden % 256 == 0, "vit: hidden width %d unsupported (256, 512, 768, 1024)", d.hidden);
    LIFTER_REQUIRE(d.heads >= 1 && d.hidden % d.heads == 0 && (d.hidden / d.heads == 32 || d.hidden / d.heads == 64),
                   "vit: %d heads over width %d unsupported (head width must be 32 or 64)", d.heads, d.hidden);
    LIFTER_REQUIRE(d.blocks >= 0 && d.blocks <= 64, "vit: %d blocks", d.blocks);
    LIFTER_REQUIRE(d.mlp_dim >= 192 && d.mlp_dim % 64 == 0, "vit: MLP width %d must be a multiple of 64, at least 192", d.mlp_dim);
    LIFTER_REQUIRE(d.head_dim >= 64 && d.head_dim % 64 == 0, "vit: head MLP width %d must be a multiple of 64", d.head_dim);
    LIFTER_REQUIRE(d.out_dim >= 1 && d.out_dim <= 64, "vit: %d outputs per token unsupported (1..64)", d.out_dim);
    LIFTER_REQUIRE(d.fmt == FMT_F16 || d.fmt == FMT_BF16, "vit: unknown operand format %d", d.fmt);
    return LIFTER_OK;
}

// ---- launch wrappers shared with the training path (vit_train.cu) ---------------------------------------------------------
int vit_embed(const VitDesc& d, const float* x, long T, void* h, const float* ln0, void* u, cudaStream_t st) {
    const int H = d.hidden, nv = H / 256;
    const unsigned row_blocks = static_cast<unsigned>((T + 8 * VIT_RPW - 1) / (8 * VIT_RPW));   // 8 warps x VIT_RPW token rows per block
    const float* l1 = ln0 ? ln0 + H : nullptr;
    const float* l2 = ln0 ? ln0 + 2 * H : nullptr;
    const float* l3 = ln0 ? ln0 + 3 * H : nullptr;
    if (d.in_dim == 2) {
        VIT_DISPATCH_NV(nv, LIFTER_CHECK_CUDA(launch_pdl("vit_embed_kernel", vit_embed_kernel<NV, 2>, dim3(row_blocks), dim3(256), 0, st, x, T,
                                                         d.tokens, d.in_dim, d.map_w, d.map_b, d.pos, h, ln0, l1, l2, l3, d.ln_eps, u, d.fmt, static_cast<float*>(nullptr))));
    } else {
        VIT_DISPATCH_NV(nv, LIFTER_CHECK_CUDA(launch_pdl("vit_embed_kernel", vit_embed_kernel<NV, 0>, dim3(row_blocks), dim3(256), 0, st, x, T,
                                                         d.tokens, d.in_dim, d.map_w, d.map_b, d.pos, h, ln0, l1, l2, l3, d.ln_eps, u, d.fmt, static_cast<float*>(nullptr))));
    }
    return LIFTER_OK;
}

template <int NV, bool IN32>
static int layernorm_launch(const void* in, long T, const float* g1, const float* b1, const float* g2, const float* b2, float eps, void* out,
                            int fmt, cudaStream_t st) {
    constexpr int SMEM = 8 * LN_DEPTH * NV * 512 * (IN32 ? 2 : 1);
    static PerDeviceOnce once;
    bool* configured = once.slot();
    if (!*configured) {
        LIFTER_CHECK_CUDA(cudaFuncSetAttribute(vit_layernorm_kernel<NV, IN32>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM));
        *configured = true;
    }
    // about four blocks per SM, at least VIT_RPW rows per warp (the ring's fill is paid once per warp)
    long rows_per_warp = (T + 4L * num_sms() * 8 - 1) / (4L * num_sms() * 8);
    if (rows_per_warp < VIT_RPW) rows_per_warp = VIT_RPW;
    const unsigned blocks = static_cast<unsigned>((T + rows_per_warp * 8 - 1) / (rows_per_warp * 8));
    LIFTER_CHECK_CUDA(launch_pdl(IN32 ? "vit_layernorm_kernel<fp32 in>" : "vit_layernorm_kernel", vit_layernorm_kernel<NV, IN32>, dim3(blocks), dim3(256),
                                 static_cast<size_t>(SMEM), st, in, T, rows_per_warp, g1, b1, g2, b2, eps, out, fmt));
    return LIFTER_OK;
}

int vit_layernorm(const void* in, long T, const float* g1, const float* b1, const float* g2, const float* b2, float eps, void* out,
                  int H, int fmt, cudaStream_t st) {
    VIT_DISPATCH_NV(H / 256, return (layernorm_launch<NV, false>(in, T, g1, b1, g2, b2, eps, out, fmt, st)));
    return LIFTER_OK;
}

// LayerNorm(s) of the fp32 residual stream -> 16-bit GEMM operand (eval path)
static int vit_layernorm32(const float* in, long T, const float* g1, const float* b1, const float* g2, const float* b2, float eps,
                           void* out, int H, int fmt, cudaStream_t st) {
    VIT_DISPATCH_NV(H / 256, return (layernorm_launch<NV, true>(static_cast<const void*>(in), T, g1, b1, g2, b2, eps, out, fmt, st)));
    return LIFTER_OK;
}

template <int DH, int FMT>
static int attention_launch(const void* qkv, long pose_heads, int tokens, int heads, float scale_log2e, void* o, size_t smem, cudaStream_t st) {
    static PerDeviceOnce once;
    bool* configured = once.slot();
    if (!*configured) {
        LIFTER_CHECK_CUDA(cudaFuncSetAttribute(vit_attention_kernel<DH, FMT>, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024));
        *configured = true;
    }
    const dim3 grid(static_cast<unsigned>((pose_heads + ATT_WARPS - 1) / ATT_WARPS));
    LIFTER_CHECK_CUDA(launch_pdl("vit_attention_kernel", vit_attention_kernel<DH, FMT>, grid, dim3(ATT_WARPS * 32), smem, st, qkv, pose_heads, tokens,
                                 heads, scale_log2e, o));
    return LIFTER_OK;
}

int vit_attention(const void* qkv, long pose_heads, int tokens, int heads, int dh, void* o, int fmt, cudaStream_t st) {
    const size_t smem = att_smem_bytes(tokens, dh);
    LIFTER_REQUIRE(smem <= 64 * 1024, "vit_attention: %d tokens of width %d need %zu bytes of shared memory per block", tokens, dh, smem);
    const float scale_log2e = 1.4426950408889634f / sqrtf(static_cast<float>(dh));
    if (dh == 64) return fmt == FMT_BF16 ? attention_launch<64, FMT_BF16>(qkv, pose_heads, tokens, heads, scale_log2e, o, smem, st)
                                         : attention_launch<64, FMT_F16>(qkv, pose_heads, tokens, heads, scale_log2e, o, smem, st);
    return fmt == FMT_BF16 ? attention_launch<32, FMT_BF16>(qkv, pose_heads, tokens, heads, scale_log2e, o, smem, st)
                           : attention_launch<32, FMT_F16>(qkv, pose_heads, tokens, heads, scale_log2e, o, smem, st);
}

// Workspace per token, in 16-bit units: the residual stream h0, h1 in FP32 (2 x 2 hidden), then 16-bit u, o (hidden
// each), qkv (3 hidden), m (mlp_dim), r (head_dim).  The residual stream is fp32 in the eval path: of the fifteen 16-bit
// rounding points of the round-1 forward, the two per block on `h` were the largest error term (0.094 of 0.143 mm rms
// after de-normalisation, scripts/vit_rounding_ablation.py) and put ~0.5 % of the coordinates outside north_star's
// 0.5 mm + 1e-2 band; GEMM operands (u, o, m, the head input) and weights stay 16-bit.
static long vit_ws_halfs_per_token(const VitDesc& d) { return 9L * d.hidden + d.mlp_dim + d.head_dim; }

int64_t vit_workspace_bytes(const VitDesc& d, int64_t poses) {
    if (vit_validate(d) != LIFTER_OK || poses < 0) return LIFTER_ERR_INVALID;
    return poses * d.tokens * vit_ws_halfs_per_token(d) * 2 + 256;
}

int vit_forward(const VitDesc& d, const float* x, int64_t poses, void* workspace, float* out, cudaStream_t st) {
    LIFTER_PROPAGATE(vit_validate(d));
    if (poses <= 0) return LIFTER_OK;
    LIFTER_REQUIRE(x && workspace && out, "vit_forward: null pointer");
    LIFTER_REQUIRE(d.map_w && d.map_b && d.pos && d.head_b1 && d.head_b2 && d.w_head1 && d.w_head2, "vit_forward: null parameter");
    LIFTER_REQUIRE(d.blocks == 0 || (d.ln && d.mlp_b1 && d.mlp_b2 && d.w_qkv && d.w_out && d.w_mlp1 && d.w_mlp2), "vit_forward: null block parameter");
    LIFTER_REQUIRE(reinterpret_cast<uintptr_t>(workspace) % 16 == 0, "vit_forward: workspace must be 16-byte aligned");
    const long T = poses * d.tokens;
    LIFTER_REQUIRE(T <= 0x7fffffffL, "vit_forward: %ld tokens in one call; process the batch in chunks", T);
    const int H = d.hidden, nv = H / 256, dh = H / d.heads, M = static_cast<int>(T);
    float* h0 = static_cast<float*>(workspace);                      // fp32 residual stream, two buffers
    float* h1 = h0 + T * H;
    unsigned short* u = reinterpret_cast<unsigned short*>(h1 + T * H);
    unsigned short* o = u + T * H;
    unsigned short* qkv = o + T * H;
    unsigned short* m = qkv + T * 3L * H;
    unsigned short* r = m + T * static_cast<long>(d.mlp_dim);
    const unsigned row_blocks = static_cast<unsigned>((T + 8 * VIT_RPW - 1) / (8 * VIT_RPW));   // 8 warps x VIT_RPW token rows per block

    {   // embedding -> h0 (fp32) (+ the first block's LayerNorm pair -> u)
        const float* ln0 = d.blocks > 0 ? d.ln : nullptr;
        const float* l1 = ln0 ? ln0 + H : nullptr;
        const float* l2 = ln0 ? ln0 + 2 * H : nullptr;
        const float* l3 = ln0 ? ln0 + 3 * H : nullptr;
        if (d.in_dim == 2) {
            VIT_DISPATCH_NV(nv, LIFTER_CHECK_CUDA(launch_pdl("vit_embed_kernel", vit_embed_kernel<NV, 2>, dim3(row_blocks), dim3(256), 0, st, x, T,
                                                             d.tokens, d.in_dim, d.map_w, d.map_b, d.pos, static_cast<void*>(nullptr), ln0, l1, l2, l3,
                                                             d.ln_eps, static_cast<void*>(u), d.fmt, h0)));
        } else {
            VIT_DISPATCH_NV(nv, LIFTER_CHECK_CUDA(launch_pdl("vit_embed_kernel", vit_embed_kernel<NV, 0>, dim3(row_blocks), dim3(256), 0, st, x, T,
                                                             d.tokens, d.in_dim, d.map_w, d.map_b, d.pos, static_cast<void*>(nullptr), ln0, l1, l2, l3,
                                                             d.ln_eps, static_cast<void*>(u), d.fmt, h0)));
        }
    }
    const long pose_heads = poses * d.heads;
    const size_t esz = 2;
    for (int b = 0; b < d.blocks; ++b) {
        const float* ln = d.ln + static_cast<long>(b) * 6 * H;
        const char* wqkv = static_cast<const char*>(d.w_qkv) + static_cast<long>(b) * 3 * H * H * esz;
        const char* wout = static_cast<const char*>(d.w_out) + static_cast<long>(b) * H * H * esz;
        const char* w1 = static_cast<const char*>(d.w_mlp1) + static_cast<long>(b) * d.mlp_dim * H * esz;
        const char* w2 = static_cast<const char*>(d.w_mlp2) + static_cast<long>(b) * H * d.mlp_dim * esz;
        // u = mhsa.norm(norm1(h0))   (block 0: done by the embedding kernel)      baselineModel.py:303, 244
        if (b > 0) LIFTER_PROPAGATE(vit_layernorm32(h0, T, ln, ln + H, ln + 2 * H, ln + 3 * H, d.ln_eps, u, H, d.fmt, st));
        GemmArgs g;                                                             // qkv = u Wqkv^T (no bias)   :246
        g.A = u; g.lda = H; g.B = wqkv; g.ldb = H; g.M = M; g.N = 3 * H; g.K = H; g.fmt = d.fmt;
        g.out_mode = OUT_16_TMA; g.D = qkv; g.ldd = 3 * H; g.tag = "vit_qkv";
        LIFTER_PROPAGATE(gemm_launch(g, st));
        LIFTER_PROPAGATE(vit_attention(qkv, pose_heads, d.tokens, d.heads, dh, o, d.fmt, st));
        GemmArgs go;                                                            // h1 = h0 + o Wout^T  (fp32 in, fp32 out)   :254, 303
        go.A = o; go.lda = H; go.B = wout; go.ldb = H; go.M = M; go.N = H; go.K = H; go.fmt = d.fmt;
        go.out_mode = OUT_32_TMA; go.D = h1; go.ldd = H; go.residual32 = h0; go.ldr32 = H; go.tag = "vit_attn_out";
        LIFTER_PROPAGATE(gemm_launch(go, st));
        LIFTER_PROPAGATE(vit_layernorm32(h1, T, ln + 4 * H, ln + 5 * H, nullptr, nullptr, d.ln_eps, u, H, d.fmt, st));
        GemmArgs g1;                                                            // m = GELU(u W1^T + b1)     :296-298
        g1.A = u; g1.lda = H; g1.B = w1; g1.ldb = H; g1.M = M; g1.N = d.mlp_dim; g1.K = H; g1.fmt = d.fmt;
        g1.out_mode = OUT_16_TMA; g1.D = m; g1.ldd = d.mlp_dim; g1.shift = d.mlp_b1 + static_cast<long>(b) * d.mlp_dim; g1.relu = 2;
        g1.tag = "vit_mlp1";
        LIFTER_PROPAGATE(gemm_launch(g1, st));
        GemmArgs g2;                                                            // h0 = h1 + m W2^T + b2  (fp32)     :299, 304
        g2.A = m; g2.lda = d.mlp_dim; g2.B = w2; g2.ldb = d.mlp_dim; g2.M = M; g2.N = H; g2.K = d.mlp_dim; g2.fmt = d.fmt;
        g2.out_mode = OUT_32_TMA; g2.D = h0; g2.ldd = H; g2.shift = d.mlp_b2 + static_cast<long>(b) * H; g2.residual32 = h1; g2.ldr32 = H;
        g2.tag = "vit_mlp2";
        if (b == d.blocks - 1) {       // only the head reads the final residual stream, as a 16-bit operand: store it as one (u)
            g2.out_mode = OUT_16_TMA; g2.D = u;
        }
        LIFTER_PROPAGATE(gemm_launch(g2, st));
    }
    // no blocks: the head's first Linear takes the embedding (fp32) as a 16-bit operand through one conversion pass
    if (d.blocks == 0) LIFTER_PROPAGATE(prep_input(h0, T, H, H, nullptr, nullptr, u, d.fmt, st));
    GemmArgs gh;                                                                // r = ReLU(h0 Wh1^T + b)    :336-338
    gh.A = u; gh.lda = H; gh.B = d.w_head1; gh.ldb = H; gh.M = M; gh.N = d.head_dim; gh.K = H; gh.fmt = d.fmt;
    gh.out_mode = OUT_16_TMA; gh.D = r; gh.ldd = d.head_dim; gh.shift = d.head_b1; gh.relu = 1; gh.tag = "vit_head1";
    LIFTER_PROPAGATE(gemm_launch(gh, st));
    GemmArgs gf;                                                                // out = r Wh2^T + b (fp32)  :339
    gf.A = r; gf.lda = d.head_dim; gf.B = d.w_head2; gf.ldb = d.head_dim; gf.M = M; gf.N = 64; gf.K = d.head_dim; gf.fmt = d.fmt;
    gf.out_mode = OUT_32_DIRECT; gf.D = out; gf.n_valid = d.out_dim; gf.shift = d.head_b2; gf.tag = "vit_head2";
    LIFTER_PROPAGATE(gemm_launch(gf, st));
    return LIFTER_OK;
}

}  // namespace lifter

// ---- C ABI (include/lifter_b200.h) ---------------------------------------------------------------------------------
#include "../../include/lifter_b200.h"

lifter::VitDesc lifter_vit_to_desc(const lifter_vit_t* v) {
    lifter::VitDesc d;
    d.tokens = v->tokens; d.in_dim = v->in_dim; d.hidden = v->hidden; d.heads = v->heads; d.blocks = v->blocks;
    d.mlp_dim = v->mlp_dim; d.head_dim = v->head_dim; d.out_dim = v->out_dim; d.fmt = v->fmt; d.ln_eps = v->ln_eps;
    d.map_w = v->map_w; d.map_b = v->map_b; d.pos = v->pos; d.ln = v->ln; d.mlp_b1 = v->mlp_b1; d.mlp_b2 = v->mlp_b2;
    d.head_b1 = v->head_b1; d.head_b2 = v->head_b2; d.w_qkv = v->w_qkv; d.w_out = v->w_out; d.w_mlp1 = v->w_mlp1;
    d.w_mlp2 = v->w_mlp2; d.w_head1 = v->w_head1; d.w_head2 = v->w_head2;
    return d;
}

extern "C" {

int64_t lifter_vit_workspace_bytes(const lifter_vit_t* vit, int64_t poses) {
    if (vit == nullptr) return LIFTER_ERR_INVALID;
    return lifter::vit_workspace_bytes(lifter_vit_to_desc(vit), poses);
}

int lifter_vit_forward(const lifter_vit_t* vit, const float* x, int64_t poses, void* workspace, float* out, void* stream) {
    LIFTER_REQUIRE(vit != nullptr, "vit_forward: null descriptor");
    return lifter::vit_forward(lifter_vit_to_desc(vit), x, poses, workspace, out, static_cast<cudaStream_t>(stream));
}

}  // extern "C"
