"""CPU: the product's spline device functions (pf_spline.cuh are __host__ __device__) compiled for the
host by nvcc and checked against the oracle's rational-quadratic spline -- forward, inverse, tails, bin
edges, and the hand-written VJP against autograd through the oracle."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest
import torch

import nsf_oracle as O

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)


@pytest.fixture(scope="module")
def hlib():
    out = os.path.join(ROOT, "posteriflow_b200", "build", "host_spline_harness.so")
    os.makedirs(os.path.dirname(out), exist_ok=True)
    src = os.path.join(HERE, "host_spline_harness.cu")
    hdr = os.path.join(ROOT, "posteriflow_b200", "csrc", "pf_spline.cuh")
    if not os.path.exists(out) or os.path.getmtime(out) < max(os.path.getmtime(src), os.path.getmtime(hdr)):
        subprocess.run(["/usr/local/cuda/bin/nvcc", "-O1", "-std=c++17", "-shared", "-Xcompiler", "-fPIC",
                        "--fmad=false", "-o", out, src], check=True, capture_output=True)
    return C.CDLL(out)


def fp(a):
    return a.ctypes.data_as(C.c_void_p)


def oracle_apply(nb, params, v, B, inverse, dtype=torch.float32):
    spec = O.FlowSpec(features=1, num_bins=nb, tail_bound=B)
    p = torch.from_numpy(params).to(dtype)
    x = torch.from_numpy(v).to(dtype)[:, None]
    out, lad = O.spline_elementwise(spec, x, p, inverse)
    return out[:, 0], lad


@pytest.mark.parametrize("nb,B", [(16, 5.0), (8, 2.0), (24, 3.0), (32, 5.0)])
@pytest.mark.parametrize("inverse", [0, 1])
def test_spline_matches_oracle(hlib, nb, B, inverse):
    rng = np.random.default_rng(nb * 10 + inverse)
    n, M = 4000, 3 * nb - 1
    params = (rng.standard_normal((n, M)) * 2.0).astype(np.float32)
    v = (rng.standard_normal(n) * B * 0.6).astype(np.float32)
    v[:6] = [B, -B, B + 1.5, -B - 0.25, 0.0, np.nextafter(np.float32(B), np.float32(0))]
    out, lad = np.empty(n, np.float32), np.empty(n, np.float32)
    assert hlib.h_rqs_apply(nb, inverse, fp(params), fp(v), C.c_float(B), n, fp(out), fp(lad)) == 0
    r64, l64 = oracle_apply(nb, params, v, B, bool(inverse), torch.float64)
    r32, l32 = oracle_apply(nb, params, v, B, bool(inverse), torch.float32)
    own = float((r32.double() - r64).abs().max())
    assert float(np.abs(out - r64.numpy()).max()) <= 4 * own + 2e-6
    ownl = float((l32.double() - l64).abs().max())
    assert float(np.abs(lad - l64.numpy()).max()) <= 4 * ownl + 5e-6
    # tails: identity, zero logdet
    assert out[2] == v[2] and out[3] == v[3] and lad[2] == 0.0 and lad[3] == 0.0


def test_points_on_knots(hlib):
    nb, B = 16, 5.0
    rng = np.random.default_rng(5)
    n, M = 512, 3 * nb - 1
    params = rng.standard_normal((n, M)).astype(np.float32)
    cw, _ = O._knots(torch.from_numpy(params[:, :nb]), -B, B, 1e-3)
    v = cw[torch.arange(n), torch.arange(n) % (nb + 1)].numpy().astype(np.float32)
    out, lad = np.empty(n, np.float32), np.empty(n, np.float32)
    hlib.h_rqs_apply(nb, 0, fp(params), fp(v), C.c_float(B), n, fp(out), fp(lad))
    r64, l64 = oracle_apply(nb, params, v, B, False, torch.float64)
    assert float(np.abs(out - r64.numpy()).max()) <= 2e-5  # continuous across knots
    assert np.isfinite(lad).all()


@pytest.mark.parametrize("nb,B", [(16, 5.0), (8, 2.0)])
def test_spline_vjp_matches_autograd(hlib, nb, B):
    rng = np.random.default_rng(nb)
    n, M = 1500, 3 * nb - 1
    params = (rng.standard_normal((n, M)) * 1.5).astype(np.float32)
    x = (rng.standard_normal(n) * B * 0.5).astype(np.float32)
    x[:2] = [B + 1.0, -B - 2.0]
    gy = rng.standard_normal(n).astype(np.float32)
    gl = rng.standard_normal(n).astype(np.float32)
    gp, gx = np.empty((n, M), np.float32), np.empty(n, np.float32)
    assert hlib.h_rqs_backward(nb, fp(params), fp(x), C.c_float(B), fp(gy), fp(gl), n, fp(gp), fp(gx)) == 0
    spec = O.FlowSpec(features=1, num_bins=nb, tail_bound=B)
    p = torch.from_numpy(params).double().requires_grad_(True)
    xx = torch.from_numpy(x).double()[:, None].requires_grad_(True)
    out, lad = O.spline_elementwise(spec, xx, p, False)
    ((out[:, 0] * torch.from_numpy(gy).double()).sum() + (lad * torch.from_numpy(gl).double()).sum()).backward()
    sp = max(1.0, float(p.grad.abs().max()))
    assert float(np.abs(gp - p.grad.numpy()).max()) / sp <= 2e-4
    sx = max(1.0, float(xx.grad.abs().max()))
    assert float(np.abs(gx - xx.grad[:, 0].numpy()).max()) / sx <= 2e-4
