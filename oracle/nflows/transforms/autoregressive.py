"""nflows.transforms.autoregressive subset (restated).  SURVEY.md Appendix A.4 / A.5.
The inverse is nflows' generic D-pass fixed point: D full MADE evaluations per layer."""
import numpy as np
import torch
from torch.nn import functional as F

from ..utils import torchutils
from .base import Transform
from . import made as made_module
from .splines import rational_quadratic
from .splines.rational_quadratic import rational_quadratic_spline, unconstrained_rational_quadratic_spline


class AutoregressiveTransform(Transform):
    def __init__(self, autoregressive_net):
        super().__init__()
        self.autoregressive_net = autoregressive_net

    def forward(self, inputs, context=None):
        autoregressive_params = self.autoregressive_net(inputs, context)
        return self._elementwise_forward(inputs, autoregressive_params)

    def inverse(self, inputs, context=None):
        num_inputs = int(np.prod(inputs.shape[1:]))
        outputs = torch.zeros_like(inputs)
        logabsdet = None
        for _ in range(num_inputs):
            autoregressive_params = self.autoregressive_net(outputs, context)
            outputs, logabsdet = self._elementwise_inverse(inputs, autoregressive_params)
        return outputs, logabsdet

    def _output_dim_multiplier(self):
        raise NotImplementedError()

    def _elementwise_forward(self, inputs, autoregressive_params):
        raise NotImplementedError()

    def _elementwise_inverse(self, inputs, autoregressive_params):
        raise NotImplementedError()


class MaskedPiecewiseRationalQuadraticAutoregressiveTransform(AutoregressiveTransform):
    def __init__(self, features, hidden_features, context_features=None, num_bins=10, tails=None,
                 tail_bound=1.0, num_blocks=2, use_residual_blocks=True, random_mask=False,
                 activation=F.relu, dropout_probability=0.0, use_batch_norm=False,
                 min_bin_width=rational_quadratic.DEFAULT_MIN_BIN_WIDTH,
                 min_bin_height=rational_quadratic.DEFAULT_MIN_BIN_HEIGHT,
                 min_derivative=rational_quadratic.DEFAULT_MIN_DERIVATIVE):
        self.num_bins = num_bins
        self.min_bin_width = min_bin_width
        self.min_bin_height = min_bin_height
        self.min_derivative = min_derivative
        self.tails = tails
        self.tail_bound = tail_bound
        autoregressive_net = made_module.MADE(
            features=features, hidden_features=hidden_features, context_features=context_features,
            num_blocks=num_blocks, output_multiplier=self._output_dim_multiplier(),
            use_residual_blocks=use_residual_blocks, random_mask=random_mask, activation=activation,
            dropout_probability=dropout_probability, use_batch_norm=use_batch_norm)
        super().__init__(autoregressive_net)

    def _output_dim_multiplier(self):
        if self.tails == "linear":
            return self.num_bins * 3 - 1
        elif self.tails is None:
            return self.num_bins * 3 + 1
        raise ValueError

    def _elementwise(self, inputs, autoregressive_params, inverse=False):
        batch_size, features = inputs.shape[0], inputs.shape[1]
        transform_params = autoregressive_params.view(batch_size, features, self._output_dim_multiplier())

        unnormalized_widths = transform_params[..., : self.num_bins]
        unnormalized_heights = transform_params[..., self.num_bins: 2 * self.num_bins]
        unnormalized_derivatives = transform_params[..., 2 * self.num_bins:]

        if hasattr(self.autoregressive_net, "hidden_features"):
            unnormalized_widths = unnormalized_widths / np.sqrt(self.autoregressive_net.hidden_features)
            unnormalized_heights = unnormalized_heights / np.sqrt(self.autoregressive_net.hidden_features)

        if self.tails is None:
            spline_fn, spline_kwargs = rational_quadratic_spline, {}
        elif self.tails == "linear":
            spline_fn = unconstrained_rational_quadratic_spline
            spline_kwargs = {"tails": self.tails, "tail_bound": self.tail_bound}
        else:
            raise ValueError

        outputs, logabsdet = spline_fn(
            inputs=inputs, unnormalized_widths=unnormalized_widths,
            unnormalized_heights=unnormalized_heights, unnormalized_derivatives=unnormalized_derivatives,
            inverse=inverse, min_bin_width=self.min_bin_width, min_bin_height=self.min_bin_height,
            min_derivative=self.min_derivative, **spline_kwargs)
        return outputs, torchutils.sum_except_batch(logabsdet)

    def _elementwise_forward(self, inputs, autoregressive_params):
        return self._elementwise(inputs, autoregressive_params)

    def _elementwise_inverse(self, inputs, autoregressive_params):
        return self._elementwise(inputs, autoregressive_params, inverse=True)
