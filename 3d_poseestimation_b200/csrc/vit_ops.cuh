// MyViT lifter (phase1_lifting/baselineModel.py:307-362), eval-mode forward over the tcgen05 GEMM.
#pragma once
#include <math_constants.h>

#include "common.cuh"

namespace lifter {

struct VitDesc {
    int tokens = 17, in_dim = 2, hidden = 256, heads = 4, blocks = 2, mlp_dim = 1024, head_dim = 128, out_dim = 3;
    int fmt = FMT_F16;
    float ln_eps = 1e-5f;
    // fp32 parameters
    const float* map_w = nullptr; const float* map_b = nullptr; const float* pos = nullptr;   // [hidden][in_dim], [hidden], [tokens][hidden]
    const float* ln = nullptr;          // [blocks][6][hidden]: norm1 w, b; mhsa.norm w, b; norm2 w, b
    const float* mlp_b1 = nullptr;      // [blocks][mlp_dim]
    const float* mlp_b2 = nullptr;      // [blocks][hidden]
    const float* head_b1 = nullptr;     // [head_dim]
    const float* head_b2 = nullptr;     // [64], zero padded past out_dim
    // 16-bit GEMM operands, [out][in] row-major as torch stores Linear.weight
    const void* w_qkv = nullptr;        // [blocks][3 hidden][hidden]
    const void* w_out = nullptr;        // [blocks][hidden][hidden]
    const void* w_mlp1 = nullptr;       // [blocks][mlp_dim][hidden]
    const void* w_mlp2 = nullptr;       // [blocks][hidden][mlp_dim]
    const void* w_head1 = nullptr;      // [head_dim][hidden]
    const void* w_head2 = nullptr;      // [64][head_dim], rows past out_dim zero
};

int vit_validate(const VitDesc& d);
int64_t vit_workspace_bytes(const VitDesc& d, int64_t poses);
// x [poses][tokens][in_dim] fp32 -> out [poses][tokens][out_dim] fp32.  Activations live in `workspace`.
int vit_forward(const VitDesc& d, const float* x, int64_t poses, void* workspace, float* out, cudaStream_t st);

// launch wrappers (vit_ops.cu) re-used by the training path
int vit_embed(const VitDesc& d, const float* x, long T, void* h, const float* ln0, void* u, cudaStream_t st);
int vit_layernorm(const void* in, long T, const float* g1, const float* b1, const float* g2, const float* b2, float eps, void* out,
                  int H, int fmt, cudaStream_t st);
int vit_attention(const void* qkv, long pose_heads, int tokens, int heads, int dh, void* o, int fmt, cudaStream_t st);

// ---- training path (vit_train.cu) ----
struct VitGrads {                        // fp32 gradients, stacked per block like the parameters of VitDesc
    float* map = nullptr;                // [1 + in_dim][hidden]: d bias, then d weight column by input
    float* ln = nullptr;                 // [blocks][6][hidden]
    float* w_qkv = nullptr; float* w_out = nullptr; float* w_mlp1 = nullptr; float* mlp_b1 = nullptr;
    float* w_mlp2 = nullptr; float* mlp_b2 = nullptr;
    float* w_head1 = nullptr; float* head_b1 = nullptr;
    float* w_head2 = nullptr;            // [out_dim][head_dim]
    float* head_b2 = nullptr;            // [out_dim]
};
int64_t vit_saved_bytes(const VitDesc& d, int64_t poses);
int64_t vit_bwd_scratch_bytes(const VitDesc& d, int64_t poses);
// forward that keeps every GEMM input in `saved`; out [poses][tokens][out_dim] fp32
int vit_train_forward(const VitDesc& d, const float* x, int64_t poses, void* saved, float* out, cudaStream_t st);
// dout [poses][tokens][out_dim] fp32 -> gradients of every parameter, and (dx non-null) of the key points [poses][tokens][in_dim]
int vit_backward(const VitDesc& d, const float* x, int64_t poses, const void* saved, const float* dout, void* scratch,
                 const VitGrads& G, float* dx, cudaStream_t st);

}  // namespace lifter
