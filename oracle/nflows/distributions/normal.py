"""nflows.distributions.normal.StandardNormal (restated).  Imported by
src/ahsd/models/flows.py:22 but unused on the LeanNPE path (PSDScaledNormal is the base)."""
import math

import torch

from .base import Distribution


class StandardNormal(Distribution):
    def __init__(self, shape):
        super().__init__()
        self._shape = torch.Size(shape)
        n = 1
        for s in shape:
            n *= int(s)
        self.register_buffer("_log_z", torch.tensor(0.5 * n * math.log(2 * math.pi), dtype=torch.float64),
                             persistent=False)

    def _log_prob(self, inputs, context):
        neg_energy = -0.5 * inputs.reshape(inputs.shape[0], -1).pow(2).sum(dim=1)
        return neg_energy - self._log_z.to(inputs.dtype)

    def _sample(self, num_samples, context):
        if context is None:
            return torch.randn(num_samples, *self._shape, device=self._log_z.device)
        b = context.shape[0]
        return torch.randn(b, num_samples, *self._shape, device=context.device)
