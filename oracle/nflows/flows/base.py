"""nflows.flows.base.Flow (restated).  The reference builds Flow(transform, base_dist) at
src/ahsd/models/flows.py:532, which is why the state dict holds every flow tensor twice
(`transform._transforms.*` and `flow._transform._transforms.*`)."""
import torch.nn as nn

from ..distributions.base import Distribution


class Flow(Distribution):
    def __init__(self, transform, distribution, embedding_net=None):
        super().__init__()
        self._transform = transform
        self._distribution = distribution
        self._embedding_net = embedding_net if embedding_net is not None else nn.Identity()

    def _log_prob(self, inputs, context):
        embedded = self._embedding_net(context)
        noise, logabsdet = self._transform(inputs, context=embedded)
        return self._distribution.log_prob(noise, context=embedded) + logabsdet

    def _sample(self, num_samples, context):
        raise NotImplementedError("oracle shim: Flow.sample is not on the LeanNPE path")
