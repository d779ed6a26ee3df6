"""CPU oracle for the phase1_lifting ``LinearModel`` hot path.

TEST INFRASTRUCTURE ONLY.  Nothing in ``3d_poseestimation_b200/`` may import this
package: it is the checker used by ``tests/``, ``__graft_entry__.smoke()`` and the
``cpu_baseline`` / ``--impl reference`` legs of ``bench.py``.  It restates, in plain
PyTorch fp32 on the CPU (the reference's own arithmetic library), the algorithm of

* ``phase1_lifting/baselineModel.py:10-102``  (weight_init, Linear block, LinearModel)
* ``phase5_loop/losses.py:3-119``             (loss_MPJPE, TriangleLoss, TriangleLoss_sep)
* ``phase3_direct/my_HybrIK/utils.py:372-396`` (flip_pose)
* ``phase1_lifting/main.py:334-337,473-479``  (normalise / denormalise)
* ``torch.optim.Adam / AdamW`` as called at ``phase1_lifting/train_1.py:39``

Parity pin: every function here is checked against outputs of the *imported reference
itself* (``oracle/make_golden.py`` ran ``/root/reference`` in the build container and
committed the vectors under ``tests/golden/``); see ``tests/test_oracle_golden.py``.
Two terms have no reference implementation and are marked PARITY UNPINNED in
``oracle/losses_oracle.py``: the pinhole reprojection loss and the bone-length loss.
"""
