"""CPU restatement of the dataset-side keypoint processing, numpy float64.

TEST INFRASTRUCTURE (see ``oracle/__init__.py``).

Pinned against the reference's own ``process_data`` (``phase1_lifting/main.py:271-355``, run by
``oracle/make_golden.py process_data`` through an exec of the source slice; vectors in
``tests/golden/process_data.npz``):

* ``root_centre``      <- ``main.py:275-277``  every joint but the root minus the frame's root joint
* ``keypoint_stats``   <- ``main.py:280-289``  per (joint, coordinate) mean and POPULATION standard deviation
  (divide by n_frames) over the frames, taken AFTER root-centring and BEFORE the root is zeroed; the min / max
  rows are the statistics of ``phase5_loop/train_5.py:35-40`` (min/max de-normalisation)
* ``process_data``     <- the whole method: centring, statistics, ``(x - mean) / std`` (``main.py:334,337``) or
  ``2x - 1`` (``main.py:331``), then the root row zeroed (``main.py:343-344``) or dropped for the 16-joint
  variant (``main.py:340-342``)
"""
import numpy as np


def root_centre(dataset):
    out = np.array(dataset, dtype=np.float64, copy=True)
    out[:, 1:] -= out[:, :1]
    return out


def keypoint_stats(dataset, z_c=True):
    """Returns mean, std, min, max, each (J, D) float64."""
    d = root_centre(dataset) if z_c else np.asarray(dataset, dtype=np.float64)
    mean = d.sum(axis=0) / d.shape[0]
    std = np.sqrt(((d - mean) ** 2).sum(axis=0) / d.shape[0])
    return mean, std, d.min(axis=0), d.max(axis=0)


def process_data(dataset, mean=None, std=None, *, standardize=False, z_c=True, minmax=False, num_of_joints=17):
    """``mean``/``std`` None: statistics of this data set (``is_train=True``).  Returns (out, mean, std)."""
    d = root_centre(dataset) if z_c else np.array(dataset, dtype=np.float64, copy=True)
    if mean is None:
        mean, std, _, _ = keypoint_stats(dataset, z_c)
    if standardize:
        if minmax and d.shape[2] == 2:
            d = 2 * d - 1
        else:
            d = (d - mean) / std
    if num_of_joints == 16:
        d = d[:, 1:, :].copy()
    elif z_c:
        d[:, :1, :] *= 0
    return d, mean, std
