"""Helpers shared by the test modules (imported as a top-level module; conftest puts tests/ on sys.path)."""
import numpy as np
import torch


def t32(a):
    return torch.from_numpy(np.asarray(a)).float()
