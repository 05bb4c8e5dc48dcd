"""pf_sample_fused_static: the fused sampler with its scratch in a file-scope __device__ array (no allocation inside the
call).  No counterpart in the reference (include/posteriflow_b200.h).

Collected LAST on purpose (file name).  Called straight through the ctypes binding of
the C-ABI, against the existing pf_sample_fused on the same inputs: same kernels, same projection values, only the
address of the scratch differs -- the bound is the one the suite uses for "same kernel, different staging"
(tests/test_gpu_round2.py::test_row2ctx_equals_the_materialised_context: 1e-6 / 1e-5 absolute)."""
import ctypes as C

import pytest
import torch

from util import load_golden, make_flow

pytestmark = pytest.mark.gpu

STATIC_BYTES = 32 << 20


def _scaler(device):
    import posteriflow_b200 as P
    return P.ParamScaler().to(device)


def _call(fh, sc, entry, ctx, z, seed, n, flags, ws=None):
    """one call of `entry` ("pf_sample_fused" / "pf_sample_fused_static") through ctypes; returns
    (status, (theta, logdet, railed, bytes of device memory that disappeared during the call))"""
    from posteriflow_b200 import _cabi as A
    from posteriflow_b200.ops import _scaler_struct, _stream
    E, D = ctx.shape[0], fh.shape[0]
    theta = torch.full((E, n, D), float("nan"), dtype=torch.float32, device=ctx.device)
    logdet = torch.full((E * n,), float("nan"), dtype=torch.float32, device=ctx.device)
    railed = torch.full((E * n,), 7, dtype=torch.uint8, device=ctx.device)
    st, keep = _scaler_struct(sc.lo, sc.hi, sc.log_mask, sc.circ_mask)
    zp = z.data_ptr() if z is not None else None
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info(fh.index)[0]
    with torch.cuda.device(fh.index):
        if entry == "pf_sample_fused":
            rc = fh.lib.pf_sample_fused(fh.h, ctx.data_ptr(), C.byref(st), zp, seed, E, n, theta.data_ptr(),
                                        logdet.data_ptr(), railed.data_ptr(), flags, ws.data_ptr(), ws.numel(), _stream())
        else:
            rc = fh.lib.pf_sample_fused_static(fh.h, ctx.data_ptr(), C.byref(st), zp, seed, E, n, theta.data_ptr(),
                                               logdet.data_ptr(), railed.data_ptr(), flags, _stream())
    torch.cuda.synchronize()
    grew = free0 - torch.cuda.mem_get_info(fh.index)[0]   # device memory taken while the C call ran
    del keep
    return rc, (theta, logdet, railed, grew)


@pytest.mark.parametrize("tier", ["bf16", "fp32"])
def test_static_scratch_sampler_equals_the_workspace_form_on_two_calls_in_a_row(tier):
    from posteriflow_b200 import _cabi as A
    spec, w, g = load_golden("lean_default")
    flow = make_flow(spec, w, "cuda", tier=tier)
    fh = flow.handle()
    sc = _scaler("cuda")
    flags = A.PF_FLAG_SANITIZE_CTX | A.PF_FLAG_CLAMP | A.PF_FLAG_WRAP | A.PF_FLAG_DENORMALIZE
    gen = torch.Generator().manual_seed(23)
    # two calls in a row of DIFFERENT shape: the second one finds the first one's projections in the shared array
    # (more events first, so stale rows lie beyond the second call's) and must not be affected by them
    shapes = [(9, 300), (4, 515)] if tier == "bf16" else [(5, 40), (2, 77)]
    for k, (E, n) in enumerate(shapes):
        ctx = torch.randn(E, spec.context_features, generator=gen).cuda()
        z = torch.randn(E * n, spec.features, generator=gen).cuda() if k == 0 else None   # given noise, then Philox
        need = fh.lib.pf_flow_workspace_bytes(fh.h, A.PF_OP_SAMPLE, E * n, E)
        ws = torch.empty(need, dtype=torch.uint8, device="cuda")
        rc, ref = _call(fh, sc, "pf_sample_fused", ctx, z, 1234 + k, n, flags, ws)
        assert rc == 0
        rc, got = _call(fh, sc, "pf_sample_fused_static", ctx, z, 1234 + k, n, flags)
        assert rc == 0, fh.lib.pf_last_cuda_error()
        fh.check()
        assert not bool(torch.isnan(got[0]).any()) and not bool(torch.isnan(got[1]).any())
        assert float((got[0] - ref[0]).abs().max()) <= 1e-6
        assert float((got[1] - ref[1]).abs().max()) <= 1e-5
        assert torch.equal(got[2], ref[2]) and int(got[2].max()) <= 1
        # the driver's free-memory figure did not move across the call (the first call on a device may load the module
        # that holds the array: only the second is held to it)
        if k == 1:
            assert got[3] == 0, got[3]


def test_static_scratch_sampler_rejects_sizes_beyond_the_array():
    from posteriflow_b200 import _cabi as A
    spec, w, g = load_golden("lean_default")
    flow = make_flow(spec, w, "cuda", tier="bf16")
    fh = flow.handle()
    sc = _scaler("cuda")
    per_event = spec.num_layers * 3 * 260 * 4            # padded projection row of the tensor-core tier (KTOT = 260)
    E_ok = STATIC_BYTES // per_event - 1                 # 1074 events fit ...
    E_big = STATIC_BYTES // per_event + 2                # ... 1077 do not
    flags = A.PF_FLAG_SANITIZE_CTX | A.PF_FLAG_CLAMP
    gen = torch.Generator().manual_seed(3)
    ctx = torch.randn(E_big, spec.context_features, generator=gen).cuda()
    rc, out = _call(fh, sc, "pf_sample_fused_static", ctx, None, 5, 2, flags)
    assert rc == -4                                      # PF_ERR_WORKSPACE, nothing launched, outputs untouched
    msg = fh.lib.pf_last_cuda_error().decode()
    assert "pf_sample_fused_static" in msg and str(STATIC_BYTES) in msg and str(E_big) in msg
    assert bool(torch.isnan(out[0]).all())
    rc, out = _call(fh, sc, "pf_sample_fused_static", ctx[:E_ok].contiguous(), None, 5, 2, flags)
    assert rc == 0, fh.lib.pf_last_cuda_error()
    fh.check()
    assert not bool(torch.isnan(out[0]).any())
    ws = torch.empty(fh.lib.pf_flow_workspace_bytes(fh.h, A.PF_OP_SAMPLE, E_ok * 2, E_ok), dtype=torch.uint8, device="cuda")
    rc, ref = _call(fh, sc, "pf_sample_fused", ctx[:E_ok].contiguous(), None, 5, 2, flags, ws)
    assert rc == 0
    assert float((out[0] - ref[0]).abs().max()) <= 1e-6 and float((out[1] - ref[1]).abs().max()) <= 1e-5
