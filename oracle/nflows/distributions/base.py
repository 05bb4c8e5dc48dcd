"""nflows.distributions.base subset (restated).  Test infrastructure only."""
import torch.nn as nn


class Distribution(nn.Module):
    def forward(self, *args):
        raise RuntimeError("Forward method cannot be called for a Distribution object.")

    def log_prob(self, inputs, context=None):
        return self._log_prob(inputs, context)

    def sample(self, num_samples, context=None):
        return self._sample(num_samples, context)
