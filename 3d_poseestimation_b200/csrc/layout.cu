#include "layout.cuh"

namespace lifter {

int make_layout(const lifter_config_t* cfg, Layout* lay) {
    LIFTER_REQUIRE(cfg != nullptr, "null config");
    const int H = cfg->linear_size;
    LIFTER_REQUIRE(H >= 64 && H <= 2048 && (H & (H - 1)) == 0,
                   "linear_size=%d unsupported: the sm_100a kernels need a power of two in [64, 2048]", H);
    LIFTER_REQUIRE(cfg->i_dim >= 1 && cfg->i_dim <= 4096, "i_dim=%d out of range [1, 4096]", cfg->i_dim);
    LIFTER_REQUIRE(cfg->o_dim >= 1 && cfg->o_dim <= 64, "o_dim=%d unsupported: at most 64 output features", cfg->o_dim);
    LIFTER_REQUIRE(cfg->num_stage >= 0 && cfg->num_stage <= 64, "num_stage=%d out of range [0, 64]", cfg->num_stage);
    LIFTER_REQUIRE(cfg->p_dropout >= 0.0f && cfg->p_dropout <= 1.0f,
                   "dropout probability has to be between 0 and 1, but got %g", cfg->p_dropout);
    lay->I = cfg->i_dim; lay->O = cfg->o_dim; lay->H = H; lay->S = cfg->num_stage; lay->L = 2 * cfg->num_stage + 1;
    lay->Ip = ceil_div(lay->I, 64) * 64;
    lay->Op = 64;
    lay->p_off.clear(); lay->p_numel.clear(); lay->w_off.clear();
    int64_t off = 0, woff = 0;
    auto push = [&](int64_t n) {
        lay->p_off.push_back(off);
        lay->p_numel.push_back(n);
        off += static_cast<int64_t>(align_up(static_cast<size_t>(n), 32));
    };
    for (int l = 0; l < lay->L; ++l) {
        push(static_cast<int64_t>(H) * lay->in_dim(l));
        push(H); push(H); push(H);
        lay->w_off.push_back(woff);
        woff += static_cast<int64_t>(H) * lay->in_pad(l);
    }
    push(static_cast<int64_t>(lay->O) * H);
    push(lay->O);
    lay->w_off.push_back(woff);
    woff += static_cast<int64_t>(lay->Op) * H;
    lay->p_total = off;
    lay->w_total = woff;
    return LIFTER_OK;
}

}  // namespace lifter
