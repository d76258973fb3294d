"""Seeding / determinism switch of the reference (reference resnet/main.py:16-21 ``set_random_seeds``)."""
import random

import numpy as np
import torch


def set_random_seeds(seed=0):
    """Seed torch / numpy / random and ask for deterministic algorithms where libraries are still in play.

    The reference sets the cuDNN flags here; our hot path has no cuDNN, and its kernels are deterministic by
    construction (two-pass reductions, fixed-rank-order all-reduce, deterministic split-K), so the flags only
    matter for the stock-op oracle path. They are kept so that path behaves like the reference.
    """
    torch.manual_seed(seed)
    torch.backends.cudnn.deterministic = True
    torch.backends.cudnn.benchmark = False
    np.random.seed(seed)
    random.seed(seed)
