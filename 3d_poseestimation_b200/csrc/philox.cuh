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

// The ten round keys depend on the seed only, so the host expands them once per launch and they reach the kernel as
// constant-bank operands (no per-round key arithmetic on the device).
struct PhiloxKeys { uint32_t k0[10], k1[10]; };
static inline PhiloxKeys philox_round_keys(uint64_t seed) {
    PhiloxKeys k;
    uint32_t a = static_cast<uint32_t>(seed), b = static_cast<uint32_t>(seed >> 32);
    for (int r = 0; r < 10; ++r) { k.k0[r] = a; k.k1[r] = b; a += 0x9E3779B9u; b += 0xBB67AE85u; }
    return k;
}

// One round = two 32x32->64 multiplies (IMAD.WIDE) and two three-input XORs.
__device__ __forceinline__ Philox4 philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3,
                                                 const PhiloxKeys& k) {
    constexpr uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u;
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        const uint64_t p0 = static_cast<uint64_t>(M0) * c0;
        const uint64_t p1 = static_cast<uint64_t>(M1) * c2;
        const uint32_t n0 = static_cast<uint32_t>(p1 >> 32) ^ c1 ^ k.k0[r];
        const uint32_t n2 = static_cast<uint32_t>(p0 >> 32) ^ c3 ^ k.k1[r];
        c1 = static_cast<uint32_t>(p1); c3 = static_cast<uint32_t>(p0);
        c0 = n0; c2 = n2;
    }
    return {c0, c1, c2, c3};
}

// keep decision of column 8 * group + e from the four words of one call: lane e is the low (e even) or high (e odd)
// half of word e / 2, compared in the top 16 bits so that neither half needs a mask:
//     u16 >= t   <=>   (u16 << 16) >= (t << 16)   (t <= 65535; t = 65536 means p = 1, where the caller's 1/(1-p) = 0
//     scaling already forces every output to zero)
__device__ __forceinline__ bool dropout_keep_lane(const uint32_t (&w)[4], int e, uint32_t threshold_hi) {
    const uint32_t v = (e & 1) ? w[e >> 1] : (w[e >> 1] << 16);
    return v >= threshold_hi;
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
