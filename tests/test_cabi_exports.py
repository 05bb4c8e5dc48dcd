"""CPU: the C-ABI library builds, loads, and exports every symbol include/posteriflow_b200.h declares;
the argument-validation paths that need no GPU behave; the product has no CPU fallback."""
import ctypes as C
import os
import re

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    from posteriflow_b200.build import build_extension
    build_extension()
    from posteriflow_b200 import _cabi as A
    return A.load_library()


def header_functions():
    src = open(os.path.join(ROOT, "include", "posteriflow_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(pf_[a-z0-9_]+)\s*\(", src)))


def test_every_declared_symbol_is_exported(lib):
    from posteriflow_b200 import _cabi as A
    names = header_functions()
    assert len(names) >= 15
    for n in names:
        assert hasattr(lib, n), f"{n} declared in the header but not exported"
    assert sorted(A.EXPORTS) == names
    assert lib.pf_abi_version() == 1


def test_status_strings_and_null_handling(lib):
    assert lib.pf_status_str(0) == b"ok"
    for s in range(-6, 0):
        assert len(lib.pf_status_str(s)) > 3
    assert lib.pf_flow_create(None, None) == -1
    assert lib.pf_flow_workspace_bytes(None, 1, 10, 1) == 0
    assert lib.pf_flow_saved_bytes(None, 10, 1) == 0
    assert lib.pf_flow_last_launch_count(None) == 0
    lib.pf_flow_destroy(None)


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU behaviour")
def test_no_cpu_fallback(lib):
    from posteriflow_b200 import _cabi as A
    import posteriflow_b200 as P
    desc = A.FlowDesc(11, 288, 256, 10, 16, 2, 5.0, 3.0, 0, 0)
    h = C.c_void_p()
    assert lib.pf_flow_create(C.byref(desc), C.byref(h)) == -2  # PF_ERR_CUDA: no device
    assert b"(" in lib.pf_last_cuda_error()
    flow = P.NSFPosteriorFlow(features=11, context_features=288, num_layers=2)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        flow.inverse(torch.zeros(2, 11), torch.zeros(2, 288))
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        flow.forward(torch.zeros(2, 11), torch.zeros(2, 288))


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "posteriflow_b200")
    for dp, _, fs in os.walk(pkg):
        for f in fs:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                s = open(os.path.join(dp, f)).read()
                assert "nsf_oracle" not in s and "import oracle" not in s and "from oracle" not in s, f
