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

}  // namespace lifter
