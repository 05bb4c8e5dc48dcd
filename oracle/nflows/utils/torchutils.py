"""nflows.utils.torchutils subset (restated).  Test infrastructure only."""
import torch


def tile(x, n):
    """Repeat every element of the flattened `x` n times, contiguously:
    [a, b] -> [a]*n + [b]*n.  Used for MADE output degrees."""
    if not (isinstance(n, int) and n > 0):
        raise TypeError("Argument 'n' must be a positive integer.")
    return x.reshape(-1).repeat(n).reshape(n, -1).transpose(1, 0).reshape(-1)


def sum_except_batch(x, num_batch_dims=1):
    reduce_dims = list(range(num_batch_dims, x.ndimension()))
    return torch.sum(x, dim=reduce_dims) if reduce_dims else x


def searchsorted(bin_locations, inputs, eps=1e-6):
    """Bin index = (#knots <= x) - 1, with eps added IN PLACE to the last knot so that
    x == right edge falls in the last bin."""
    bin_locations[..., -1] += eps
    return torch.sum(inputs[..., None] >= bin_locations, dim=-1) - 1
