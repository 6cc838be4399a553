"""Stub of pytorch_lightning for importing the reference in oracle/make_golden.py."""
import torch.nn as _nn


class LightningModule(_nn.Module):
    pass


class LightningDataModule:
    pass


class Trainer:
    def __init__(self, *a, **k):
        pass


class Callback:
    pass


def seed_everything(seed, *a, **k):
    import torch, random
    import numpy as np
    random.seed(seed); np.random.seed(seed); torch.manual_seed(seed)
