// Per-pose arithmetic shared by the fused pose-loss kernel (pose_ops.cu) and the phase-5 loss kernel (phase5_ops.cu):
// de-standardisation, MPJPE, pinhole reprojection and bone-length terms of ONE pose and their gradient.
#pragma once
#include "common.cuh"

namespace lifter {

constexpr int PL_ROWS = 64;             // rows (= threads) per block
constexpr int PL_MAXJ = 21;

__device__ __forceinline__ float pl_warp_sum(float v) {
#pragma unroll
    for (int s = 16; s >= 1; s >>= 1) v += __shfl_xor_sync(0xffffffffu, v, s);
    return v;
}
__device__ __forceinline__ float pl_sgn(float d) { return d > 0.0f ? 1.0f : (d < 0.0f ? -1.0f : 0.0f); }

struct PoseRowArgs {
    int J;
    const float* mean; const float* stdv;          // [J*3] nullable: joints = v * std + mean
    float g_mp, g_rp, g_bn;                        // gradient scale of each term (weight / its element count)
    bool reproj, bone;
    float rx, ry, rz, fx, fy, cx, cy;              // root offset and intrinsics of this pose
};

// pred / tgt: this pose's J*3 standardised values, t2d: its J*2 target key points (pixel / 1000; read when reproj).
// G receives d(total) / d(joints) BEFORE the chain rule through the de-standardisation; s_metric (nullable, shared
// memory): per-joint sums of the MPJPE distances, added with shared-memory atomics.
__device__ __forceinline__ void pose_row_terms(const PoseRowArgs& a, const float* pred, const float* tgt, const float* t2d,
                                               float (&G)[PL_MAXJ][3], float& l_mp, float& l_rp, float& l_bn,
                                               float* s_metric) {
    const int J = a.J;
    float P[PL_MAXJ][3], T[PL_MAXJ][3];
#pragma unroll
    for (int j = 0; j < PL_MAXJ; ++j) {
        if (j >= J) break;
#pragma unroll
        for (int d = 0; d < 3; ++d) {
            const float sd = a.stdv != nullptr ? a.stdv[3 * j + d] : 1.0f, mu = a.mean != nullptr ? a.mean[3 * j + d] : 0.0f;
            P[j][d] = fmaf(pred[3 * j + d], sd, mu);
            T[j][d] = fmaf(tgt[3 * j + d], sd, mu);
        }
        // MPJPE (losses.py:3-7 per joint; two2three.py:389-392 mean): torch.norm's gradient is 0 at 0
        const float dx = P[j][0] - T[j][0], dy = P[j][1] - T[j][1], dz = P[j][2] - T[j][2];
        const float e = sqrtf(dx * dx + dy * dy + dz * dz);
        l_mp += e;
        if (s_metric != nullptr) atomicAdd(&s_metric[j], e);
        const float s = e > 0.0f ? a.g_mp / e : 0.0f;
        G[j][0] = s * dx; G[j][1] = s * dy; G[j][2] = s * dz;
        if (a.reproj) {          // pinhole reprojection, pixel / 1000 (PARITY UNPINNED)
            const float X = P[j][0] + a.rx, Y = P[j][1] + a.ry, Z = P[j][2] + a.rz;
            const float iz = 1.0f / Z;
            const float u = (a.fx * X * iz + a.cx) * 1e-3f, v = (a.fy * Y * iz + a.cy) * 1e-3f;
            const float du = u - t2d[2 * j], dv = v - t2d[2 * j + 1];
            l_rp += fabsf(du) + fabsf(dv);
            const float su = pl_sgn(du) * a.g_rp * 1e-3f * a.fx * iz, sv = pl_sgn(dv) * a.g_rp * 1e-3f * a.fy * iz;
            G[j][0] += su;
            G[j][1] += sv;
            G[j][2] -= (su * X + sv * Y) * iz;
        }
    }
    if (a.bone && J == 17) {          // bone lengths over the 16 skeleton edges (PARITY UNPINNED)
        // (compile-time joint indices keep P / T / G in registers)
#define LIFTER_BONE(p, c)                                                                                     \
    {                                                                                                         \
        const float bx = P[c][0] - P[p][0], by = P[c][1] - P[p][1], bz = P[c][2] - P[p][2];                   \
        const float tx = T[c][0] - T[p][0], ty = T[c][1] - T[p][1], tz = T[c][2] - T[p][2];                   \
        const float lp = sqrtf(bx * bx + by * by + bz * bz), lt = sqrtf(tx * tx + ty * ty + tz * tz);         \
        l_bn += fabsf(lp - lt);                                                                               \
        const float s = lp > 0.0f ? pl_sgn(lp - lt) * a.g_bn / lp : 0.0f;                                     \
        G[c][0] += s * bx; G[c][1] += s * by; G[c][2] += s * bz;                                              \
        G[p][0] -= s * bx; G[p][1] -= s * by; G[p][2] -= s * bz;                                              \
    }
        LIFTER_BONE(0, 1) LIFTER_BONE(1, 2) LIFTER_BONE(2, 3) LIFTER_BONE(0, 4) LIFTER_BONE(4, 5) LIFTER_BONE(5, 6)
        LIFTER_BONE(0, 7) LIFTER_BONE(7, 8) LIFTER_BONE(8, 9) LIFTER_BONE(9, 10) LIFTER_BONE(8, 11)
        LIFTER_BONE(11, 12) LIFTER_BONE(12, 13) LIFTER_BONE(8, 14) LIFTER_BONE(14, 15) LIFTER_BONE(15, 16)
#undef LIFTER_BONE
    }
}

}  // namespace lifter
