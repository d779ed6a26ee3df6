// Volumetric Gaussian heat-map targets of the dataset (phase3_direct/my_HybrIK/H36_dataset.py:143-202).
#pragma once
#include "common.cuh"

namespace lifter {

constexpr int HM_GRID = 64;                         // H36_dataset.py:160: np.zeros((64, 64, 64))

// kp [n][3] (x, y, z) in [-1, 1], fp32 or fp64 -> out [n][64][64][64] fp32.  Per key point (H36_dataset.py:143-194):
// (u, v, w) = (-y, -z, x); k = 31.5 (1 + uvw); c = rint(k); a cubic window of `size` = ceil(6 sigma) made odd voxels
// around c holds exp(-|voxel - k|^2 / (2 sigma^2)), clipped to the volume; everything else is zero.  The arithmetic is
// fp64 like the reference's numpy, rounded to fp32 on the store.  status (nullable device int): bit 0 is set when a
// coordinate lies outside [-1, 1] (the reference asserts).  One streaming pass over the output: 1 MB per key point.
int keypoint_heatmaps(const void* kp, int kp_is_f64, int64_t n, double sigma, float* out, int* status, cudaStream_t st);

}  // namespace lifter
