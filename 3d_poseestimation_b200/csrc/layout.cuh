// Host-side memory layout of the lifter (flat parameter buffer, 16-bit operand copies, workspaces).
#pragma once
#include <vector>

#include "../../include/lifter_b200.h"
#include "common.cuh"

namespace lifter {

struct Layout {
    int I, O, H, S, L;       // i_dim, o_dim, linear_size, num_stage, hidden layers = 2 S + 1
    int Ip, Op;              // GEMM-padded input / output widths (multiples of 64)
    std::vector<int64_t> p_off, p_numel;     // flat fp32 parameters, reference parameters() order
    int64_t p_total;
    std::vector<int64_t> w_off;              // operand buffer: element offset of layer l's weight; w_off[L] = w2
    int64_t w_total;                         // 16-bit elements

    int in_dim(int l) const { return l == 0 ? I : H; }
    int in_pad(int l) const { return l == 0 ? Ip : H; }
    // tensor indices in the flat buffer
    int t_weight(int l) const { return 4 * l; }
    int t_bias(int l) const { return 4 * l + 1; }
    int t_gamma(int l) const { return 4 * l + 2; }
    int t_beta(int l) const { return 4 * l + 3; }
    int t_w2() const { return 4 * L; }
    int t_b2() const { return 4 * L + 1; }
};

// validates the constructor arguments; returns LIFTER_OK and fills `lay`
int make_layout(const lifter_config_t* cfg, Layout* lay);

}  // namespace lifter
