// Philox4x32-10 counter RNG and the dropout-mask stream (bit-exact twin: oracle/philox.py).
//
// The reference draws dropout masks from torch's generator (baselineModel.py:21-22,81).  Here one Philox call yields
// 128 random bits, cut into lanes of k bits; a lane is one keep decision.  k is the smallest of {1, 8, 16} that
// represents p exactly (p = 0.5, the reference's value: k = 1, so one call decides 128 elements; p a multiple of 1/256:
// k = 8; otherwise k = 16 and p is rounded to 1/65536).  A call covers R = 16 / k consecutive ROWS of 8 consecutive
// columns (R = 16, 2, 1):
//     counter = (row / R, col / 8, site, step)        key = (seed_lo, seed_hi)
//     lane    = (row % R) * 8 + col % 8               value = bits [lane k, lane k + k) of r0 | r1 << 32 | ...
//     keep    = value >= round(p * 2^k)
// `row` is the GLOBAL row index, so the shards of a data-parallel step reproduce the single-process masks.
// (Round 1 used sixteen-bit lanes for every p -- 8 decisions per call -- and the apply kernel was instruction-bound
// on Philox; for k = 16 the stream is unchanged.)
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

// How the stream is cut for a given p (host).
struct DropoutLanes {
    int bits;                 // k: 1, 8 or 16
    int rows_shift;           // log2(R): 4, 1, 0 -- rows covered by one call
    uint32_t threshold;       // round(p * 2^k): keep <=> lane value >= threshold   (threshold == 2^k: nothing kept)
};
static inline DropoutLanes dropout_lanes(float p) {
    DropoutLanes d;
    const double pd = static_cast<double>(p);
    const double p256 = pd * 256.0;
    if (p == 0.5f) { d.bits = 1; d.rows_shift = 4; }
    else if (p256 == static_cast<double>(static_cast<long>(p256))) { d.bits = 8; d.rows_shift = 1; }
    else { d.bits = 16; d.rows_shift = 0; }
    long r = static_cast<long>(pd * static_cast<double>(1l << d.bits) + 0.5);      // round half up (p >= 0)
    if (r < 0) r = 0;
    if (r > (1l << d.bits)) r = 1l << d.bits;
    d.threshold = static_cast<uint32_t>(r);
    return d;
}

// 8-bit keep mask (bit e = column 8 * group + e is kept) of global row `row` from the four words of the call that covers
// it.  `bits` is uniform over the launch, so the branches do not diverge.
__device__ __forceinline__ uint32_t dropout_keep8(const Philox4& r, uint32_t row, int bits, uint32_t threshold) {
    const uint32_t w[4] = {r.x, r.y, r.z, r.w};
    if (bits == 1) {                                          // 16 rows per call: byte (row % 16) of the 128 bits
        const uint32_t byte = (w[(row >> 2) & 3] >> ((row & 3) * 8)) & 0xFFu;
        return threshold == 1 ? byte : (threshold == 0 ? 0xFFu : 0u);
    }
    uint32_t m = 0;
    if (bits == 8) {                                          // 2 rows per call: bytes (row % 2) * 8 + e
        const uint32_t lo = w[(row & 1) * 2], hi = w[(row & 1) * 2 + 1];
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            m |= (((lo >> (8 * e)) & 0xFFu) >= threshold) ? (1u << e) : 0u;
            m |= (((hi >> (8 * e)) & 0xFFu) >= threshold) ? (1u << (4 + e)) : 0u;
        }
        return m;
    }
#pragma unroll
    for (int e = 0; e < 8; ++e)                               // one row per call: sixteen-bit lane e
        m |= (((w[e >> 1] >> (16 * (e & 1))) & 0xFFFFu) >= threshold) ? (1u << e) : 0u;
    return m;
}

}  // namespace lifter
