// Philox4x32-10 counter RNG and the dropout-mask stream (bit-exact twin: oracle/philox.py).
//
// The reference draws dropout masks from torch's generator (baselineModel.py:21-22,81).  Here one Philox call
// yields eight 16-bit lanes = the keep decisions of eight consecutive columns:
//     counter = (row, col / 8, site, step)   key = (seed_lo, seed_hi)
//     keep(e) = u16[e] >= round(p * 65536)   u16[e] = (r[e / 2] >> (16 * (e % 2))) & 0xFFFF
// `row` is the GLOBAL row index, so the shards of a data-parallel step reproduce the single-process masks.
#pragma once
#include <stdint.h>

namespace lifter {

struct Philox4 { uint32_t x, y, z, w; };

__device__ __forceinline__ Philox4 philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0,
                                                 uint32_t k1) {
    constexpr uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        const uint32_t hi0 = __umulhi(M0, c0), lo0 = M0 * c0;
        const uint32_t hi1 = __umulhi(M1, c2), lo1 = M1 * c2;
        const uint32_t n0 = hi1 ^ c1 ^ k0, n2 = hi0 ^ c3 ^ k1;
        c0 = n0; c1 = lo1; c2 = n2; c3 = lo0;
        k0 += W0; k1 += W1;
    }
    return {c0, c1, c2, c3};
}

// bit e of the result = keep decision of column 8 * group + e
__device__ __forceinline__ uint32_t dropout_keep_bits8(uint32_t row, uint32_t group, uint32_t site, uint32_t step,
                                                       uint64_t seed, uint32_t threshold) {
    const Philox4 r = philox4x32_10(row, group, site, step, static_cast<uint32_t>(seed),
                                    static_cast<uint32_t>(seed >> 32));
    const uint32_t w[4] = {r.x, r.y, r.z, r.w};
    uint32_t bits = 0;
#pragma unroll
    for (int e = 0; e < 8; ++e) {
        const uint32_t u16 = (w[e >> 1] >> (16 * (e & 1))) & 0xFFFFu;
        bits |= (u16 >= threshold ? 1u : 0u) << e;
    }
    return bits;
}

static inline uint32_t dropout_threshold(float p) {
    double t = static_cast<double>(p) * 65536.0;
    long r = static_cast<long>(t + 0.5);          // round half up (p >= 0)
    // python's round() is banker's rounding; p * 65536 is an exact half only for p = k/131072 -- avoided by callers
    if (r < 0) r = 0;
    if (r > 65536) r = 65536;
    return static_cast<uint32_t>(r);
}

}  // namespace lifter
