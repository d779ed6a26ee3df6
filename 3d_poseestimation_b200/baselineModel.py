"""Drop-in for the reference's ``phase1_lifting/baselineModel.py`` import name.

The reference's scripts do ``sys.path.append("../phase1_lifting/")`` and
``from baselineModel import LinearModel`` (train_1.py:15-17, train_5.py:22-24, backup_june8.py:19).
Point that path at this directory instead and the same line imports the sm_100a implementation;
inside the package it is ``3d_poseestimation_b200.baselineModel``.
"""
try:
    from .lifter import Linear, LinearModel, weight_init
    from .vit import Attention, MyViT, MyViTBlock, get_positional_embeddings
except ImportError:                      # imported as top-level ``baselineModel`` through sys.path
    import importlib as _il
    import os as _os
    import sys as _sys
    _here = _os.path.dirname(_os.path.abspath(__file__))
    if _os.path.dirname(_here) not in _sys.path:
        _sys.path.insert(0, _os.path.dirname(_here))
    _m = _il.import_module(_os.path.basename(_here) + ".lifter")
    Linear, LinearModel, weight_init = _m.Linear, _m.LinearModel, _m.weight_init
    _v = _il.import_module(_os.path.basename(_here) + ".vit")
    Attention, MyViT, MyViTBlock, get_positional_embeddings = _v.Attention, _v.MyViT, _v.MyViTBlock, _v.get_positional_embeddings

__all__ = ["LinearModel", "Linear", "weight_init", "MyViT", "MyViTBlock", "Attention", "get_positional_embeddings"]
