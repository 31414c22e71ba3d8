"""Test-only stand-in for `pytorch_lightning` (absent from this image)."""
import torch


class LightningModule(torch.nn.Module):
    pass


class LightningDataModule:
    pass


class Trainer:
    pass
