"""CPU: host-side mirror of the reference interface -- state-dict key set, strict
load of an nflows-layout checkpoint, ParamScaler arithmetic against the oracle restatement."""
import torch

import nsf_oracle as O
import posteriflow_b200 as P


def test_state_dict_keys_match_nflows_layout():
    spec = O.FlowSpec()
    ref = O.full_state_dict(spec, O.init_weights(spec))
    flow = P.NSFPosteriorFlow(features=11, context_features=288, hidden_features=256, num_layers=10, num_bins=16,
                              tail_bound=5.0, temperature_scale=1.0, use_masked_context=False)
    sd = flow.state_dict()
    assert set(sd.keys()) == set(ref.keys())
    for k in ref:
        assert tuple(sd[k].shape) == tuple(ref[k].shape), k
    # masks / degrees / permutations are nflows' deterministic ones
    for k in ref:
        if k.endswith((".mask", ".degrees", "_permutation", "_ar_perm", "_ar_inv_perm")):
            assert torch.equal(sd[k].long(), ref[k].long()), k
    flow.load_state_dict(ref, strict=True)
    assert sum(p.numel() for p in flow.parameters()) == 6210611
    w = flow.layer_weights()
    assert len(w) == 10 * len(P.ops.LAYER_KEYS)
    assert w[0].shape == (256, 11) and w[16].shape == (517, 256)


def test_init_distribution():
    flow = P.NSFPosteriorFlow(features=11, context_features=288, num_layers=1)
    net = flow.transform._transforms[1].autoregressive_net
    assert float(net.blocks[0].linear_layers[1].weight.abs().max()) <= 1e-3
    assert float(net.blocks[1].linear_layers[1].bias.abs().max()) <= 1e-3
    assert 0.9 / 16 < float(net.blocks[0].linear_layers[0].weight.abs().max()) <= 1 / 16
    assert float(net.context_layer.weight.abs().max()) <= 1 / 288 ** 0.5


def test_param_scaler_matches_oracle():
    for pre in (False, True):
        a, b = P.ParamScaler(premerger=pre), O.Scaler(premerger=pre)
        assert torch.equal(a.lo, b.lo) and torch.equal(a.hi, b.hi)
        assert torch.equal(a.log_mask, b.log_mask) and torch.equal(a.circ_mask, b.circ_mask)
        y = torch.randn(64, 11) * 1.5
        assert torch.equal(a.wrap(y), b.wrap(y))
        th = a.denormalize(a.wrap(y))
        assert torch.equal(th, b.denormalize(b.wrap(y)))
        assert torch.equal(a.normalize(th), b.normalize(th))


def test_unsupported_branches_raise():
    import pytest
    with pytest.raises(NotImplementedError):
        P.NSFPosteriorFlow(features=11, context_features=288, use_masked_context=True)
    with pytest.raises(NotImplementedError):
        P.NSFPosteriorFlow(features=11, context_features=288, tail_bound={0: 3.0})


def test_autoregressive_order_is_kept_and_travels_with_the_state_dict():
    """flows.py:550-586: set_autoregressive_order stores a permutation + its inverse; forward gathers x[:, perm],
    inverse scatters back (GPU arithmetic: tests/test_gpu_glue.py).  A checkpoint carrying a non-identity order must
    not evaluate as if it were the identity."""
    import pytest
    f = P.NSFPosteriorFlow(features=11, context_features=288, num_layers=1)
    assert f._ar_identity
    order = [2, 0, 1] + list(range(3, 11))
    f.set_autoregressive_order(order)
    assert not f._ar_identity and f._ar_perm.tolist() == order
    assert f._ar_inv_perm.tolist() == torch.argsort(torch.tensor(order)).tolist()
    with pytest.raises(ValueError):
        f.set_autoregressive_order([0, 0, 1] + list(range(3, 11)))
    g = P.NSFPosteriorFlow(features=11, context_features=288, num_layers=1)
    g.load_state_dict(f.state_dict(), strict=True)
    assert not g._ar_identity and g._ar_perm.tolist() == order
    g.set_autoregressive_order(list(range(11)))
    assert g._ar_identity


def test_flow_handles_are_weakly_registered():
    """ops._HANDLES must not keep a FlowHandle alive (the arena, weight images, streams and events of every flow ever
    built would leak): it is a WeakValueDictionary, the module owns the only strong reference."""
    import weakref
    assert isinstance(P.ops._HANDLES, weakref.WeakValueDictionary)
