"""CPU: the C-ABI library loads, exports every symbol include/b200spmv.h declares,
and refuses to compute without a GPU (no silent CPU fallback)."""
import ctypes as C
import os
import re

import pytest
import torch

import spmv_b200
from spmv_b200._lib import SIGNATURES

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    text = open(os.path.join(ROOT, "include", "b200spmv.h")).read()
    return sorted(set(re.findall(r"B200_API\s+(?:const char\*|int)\s+(b200_\w+)\s*\(", text)))


def test_header_symbols_exported():
    L = spmv_b200.lib()
    names = _declared()
    assert len(names) >= 40
    for n in names:
        assert hasattr(L, n), f"{n} declared in include/b200spmv.h but not exported"


def test_bindings_cover_header():
    assert set(SIGNATURES) | {"b200_last_error"} == set(_declared())


def test_abi_version():
    assert spmv_b200.lib().b200_abi_version() == 1


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU behaviour")
def test_no_cpu_fallback():
    L = spmv_b200.lib()
    plan = C.c_void_p()
    rc = L.b200_csr_plan_create(C.byref(plan), 4, 4, 0, None, 0, None)
    assert rc != 0 and b"no CPU path" in L.b200_last_error()
    with pytest.raises(spmv_b200.B200Error):
        spmv_b200.check(rc)
    rc = L.b200_gather_f64(4, C.c_void_p(8), C.c_void_p(8), C.c_void_p(8), None)
    assert rc != 0


def test_argument_checks():
    L = spmv_b200.lib()
    assert L.b200_csr_plan_create(None, 1, 1, 0, None, 0, None) == -1
    assert L.b200_gather_f64(-1, None, None, None, None) == -1
