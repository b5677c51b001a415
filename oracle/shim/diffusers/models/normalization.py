"""`diffusers.models.normalization.RMSNorm` is imported by reference src/models/dit_3b/normalization.py:16
but never instantiated for the configs in scope (norm='fusedrms' -> CustomRMSNorm).  Placeholder only."""
from torch import nn


class RMSNorm(nn.Module):
    def __init__(self, *a, **k):
        super().__init__()
        raise NotImplementedError("diffusers RMSNorm is not on the SeedVR2 hot path")
