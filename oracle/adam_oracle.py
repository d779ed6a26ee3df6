"""CPU restatement of ``torch.optim.Adam`` / ``torch.optim.AdamW`` single-tensor update.

TEST INFRASTRUCTURE (see ``oracle/__init__.py``).  The reference creates the optimiser at
``phase1_lifting/train_1.py:39`` (AdamW, torch default weight_decay 0.01), ``main.py:407`` (Adam) and
``two2three.py:419`` (Adam, weight_decay 1e-8).  Defaults: betas (0.9, 0.999), eps 1e-8, bias correction,
no amsgrad.  Pinned against torch.optim itself in ``tests/golden/adam.npz``.
"""
from __future__ import annotations

import math

import torch


def adam_step(p, g, m, v, step, lr, beta1=0.9, beta2=0.999, eps=1e-8, weight_decay=0.0, decoupled=False):
    """One update; ``step`` is 1-based.  Returns (p, m, v) new tensors (fp32)."""
    if decoupled:                       # AdamW: p *= 1 - lr*wd
        p = p * (1.0 - lr * weight_decay)
    elif weight_decay != 0.0:           # Adam: L2 folded into the gradient
        g = g + weight_decay * p
    m = beta1 * m + (1.0 - beta1) * g
    v = beta2 * v + (1.0 - beta2) * g * g
    bc1 = 1.0 - beta1 ** step
    bc2 = 1.0 - beta2 ** step
    denom = v.sqrt() / math.sqrt(bc2) + eps
    p = p - (lr / bc1) * (m / denom)
    return p, m, v


class ReduceLROnPlateauOracle:
    """``ReduceLROnPlateau(factor=0.7, patience=3, cooldown=2, min_lr=5e-6)`` host logic, train_1.py:41,106
    (mode 'min', threshold 1e-4 relative -- torch defaults)."""

    def __init__(self, lr, factor=0.7, patience=3, cooldown=2, min_lr=5e-6, threshold=1e-4):
        self.lr, self.factor, self.patience, self.cooldown, self.min_lr = lr, factor, patience, cooldown, min_lr
        self.threshold = threshold
        self.best = math.inf
        self.num_bad = 0
        self.cooldown_counter = 0

    def step(self, metric):
        if metric < self.best * (1.0 - self.threshold):
            self.best = metric
            self.num_bad = 0
        else:
            self.num_bad += 1
        if self.cooldown_counter > 0:
            self.cooldown_counter -= 1
            self.num_bad = 0
        if self.num_bad > self.patience:
            new_lr = max(self.lr * self.factor, self.min_lr)
            if self.lr - new_lr > 1e-8:
                self.lr = new_lr
            self.cooldown_counter = self.cooldown
            self.num_bad = 0
        return self.lr
