"""nflows.transforms.base subset (restated).  Test infrastructure only.
SURVEY.md Appendix A.1."""
import torch
import torch.nn as nn


class InverseNotAvailable(Exception):
    pass


class Transform(nn.Module):
    def forward(self, inputs, context=None):
        raise NotImplementedError()

    def inverse(self, inputs, context=None):
        raise InverseNotAvailable()


class CompositeTransform(Transform):
    """Applies transforms in order (forward) or each inverse in reversed order (inverse),
    summing logabsdet starting from zeros(batch)."""

    def __init__(self, transforms):
        super().__init__()
        self._transforms = nn.ModuleList(transforms)

    @staticmethod
    def _cascade(inputs, funcs, context):
        outputs = inputs
        total = inputs.new_zeros(inputs.shape[0])
        for func in funcs:
            outputs, logabsdet = func(outputs, context)
            total = total + logabsdet
        return outputs, total

    def forward(self, inputs, context=None):
        return self._cascade(inputs, self._transforms, context)

    def inverse(self, inputs, context=None):
        return self._cascade(inputs, (t.inverse for t in self._transforms[::-1]), context)
