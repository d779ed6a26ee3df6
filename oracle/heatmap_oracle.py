"""CPU restatement of the dataset's volumetric heat-map targets.

TEST INFRASTRUCTURE (see ``oracle/__init__.py``).  Restates, vectorised over key points,
``phase3_direct/my_HybrIK/H36_dataset.py``:
* ``xyz_to_uvw`` (:143-144): ``(u, v, w) = (-y, -z, x)``
* ``_keypoint_to_heatmap_3D`` (:148-194): ``k = 31.5 (1 + uvw)``; ``c = rint(k)``; a window of
  ``size = ceil(6 sigma)`` (made odd) voxels per axis around ``c`` holds
  ``exp(-((uu - k0)^2 + (vv - k1)^2 + (ww - k2)^2) / (2 sigma^2))`` (float64, stored as float32),
  clipped to the 64^3 volume; zero elsewhere.  The reference asserts every coordinate in [-1, 1].
* ``keypoints_to_heatmap_3D`` (:196-202): one volume per joint, stacked.
Pinned by ``tests/golden/heatmaps.npz``, which ``oracle/make_golden.py heatmaps`` produced by exec of the
reference's own source lines.
"""
import math

import numpy as np

GRID = 64


def window(sigma):
    size = int(math.ceil(6 * sigma))
    if not size % 2:
        size += 1
    return size


def keypoints_to_heatmap_3d(keypoints, sigma=0.5):
    """keypoints (..., 3) in [-1, 1] -> float32 (..., 64, 64, 64)."""
    kp = np.asarray(keypoints)
    assert kp.shape[-1] == 3
    flat = kp.reshape(-1, 3)                    # dtype kept: numpy scales float32 key points in float32 (k below),
    assert flat.size == 0 or (flat.min() >= -1 and flat.max() <= 1)      # the Gaussian itself is float64 either way
    out = np.zeros((flat.shape[0], GRID, GRID, GRID), dtype=np.float32)
    half = window(sigma) // 2
    for n in range(flat.shape[0]):
        uvw = np.array([-flat[n, 1], -flat[n, 2], flat[n, 0]])
        k = 31.5 * (1 + uvw)                    # float32 in, float32 out (numpy's weak scalar promotion)
        c = np.rint(k).astype(int)
        axes = [np.arange(max(0, c[a] - half), min(GRID, c[a] + half + 1)) for a in range(3)]
        if any(len(a) == 0 for a in axes):
            continue
        uu, vv, ww = np.meshgrid(*axes, indexing="ij", sparse=True)
        z = np.exp(-((uu - k[0]) ** 2 + (vv - k[1]) ** 2 + (ww - k[2]) ** 2) / (2 * (sigma ** 2)))
        out[n, axes[0][0]:axes[0][-1] + 1, axes[1][0]:axes[1][-1] + 1, axes[2][0]:axes[2][-1] + 1] = z
    return out.reshape(kp.shape[:-1] + (GRID, GRID, GRID))


def to_sparse(vol):
    """(n, 64, 64, 64) -> (key point index, flat voxel index, value) of the non-zero voxels (golden storage)."""
    n = vol.shape[0]
    flat = vol.reshape(n, -1)
    kp_i, vox_i = np.nonzero(flat)
    return kp_i.astype(np.int32), vox_i.astype(np.int32), flat[kp_i, vox_i]
