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


def test_register_budget_of_the_tuned_kernels():
    """ptxas' allocation for the pipeline kernels decides how many CTAs an SM holds, and it is fragile: unrelated
    edits to the kernel once took the 27-point walk from 96 to 120 registers (one resident CTA instead of two,
    0.85 -> 1.42 ms).  Budgets = 65536 registers / (threads per CTA x CTAs per SM of plan_default_shape)."""
    import os
    import re
    import subprocess
    log = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "build", "kernels", "spmv_csr.ptxas.log")
    if not os.path.exists(log):
        pytest.skip("no ptxas log (libraries were prebuilt elsewhere)")
    blocks = re.split(r"ptxas info\s+: Compiling entry function '([^']+)' for 'sm_100a'", open(log).read())
    regs = {}
    for i in range(1, len(blocks), 2):
        name = subprocess.run(["c++filt", blocks[i]], capture_output=True, text=True).stdout.strip()
        m = re.search(r"Used (\d+) registers", blocks[i + 1])
        regs[re.sub(r"\(.*", "", name).replace("void b200::", "")] = int(m.group(1))
    budget = {  # kernel: (threads per CTA, CTAs per SM)
        "spmv_pipe_kernel<double, false, false, false, 8, 0>": (288, 2),   # general 27-point (16-entry walk)
        "spmv_pipe_kernel<double, false, true, false, 8, 0>": (288, 2),    # ... with the p.Ap epilogue
        "spmv_pipe_kernel<double, false, false, false, 8, 4>": (288, 4),   # general 7-point (8-entry single chunk)
        "spmv_pipe_kernel<double, true, false, false, 16, 5>": (544, 2),   # symmetric 7-point (4-entry single chunk)
        "spmv_pipe_kernel<double, true, false, false, 8, 2>": (288, 3),    # symmetric 27-point (8-entry walk)
        "spmv_irr_kernel<double, false, false, false, 16, 8>": (544, 2),   # split stream walk (power-law rows)
    }
    for k, (threads, ctas) in budget.items():
        assert k in regs, k
        alloc = (regs[k] + 7) // 8 * 8  # registers are allocated in units of 8 per thread
        assert alloc * threads * ctas <= 65536, (k, regs[k])
