"""nflows.transforms.permutations subset (restated).  SURVEY.md Appendix A.1."""
import torch

from .base import Transform


class Permutation(Transform):
    def __init__(self, permutation, dim=1):
        if permutation.ndimension() != 1:
            raise ValueError("Permutation must be a 1D tensor.")
        super().__init__()
        self._dim = dim
        self.register_buffer("_permutation", permutation)

    @property
    def _inverse_permutation(self):
        return torch.argsort(self._permutation)

    @staticmethod
    def _permute(inputs, permutation, dim):
        outputs = torch.index_select(inputs, dim, permutation)
        return outputs, inputs.new_zeros(inputs.shape[0])

    def forward(self, inputs, context=None):
        return self._permute(inputs, self._permutation, self._dim)

    def inverse(self, inputs, context=None):
        return self._permute(inputs, self._inverse_permutation, self._dim)


class ReversePermutation(Permutation):
    def __init__(self, features, dim=1):
        super().__init__(torch.arange(features - 1, -1, -1), dim)
