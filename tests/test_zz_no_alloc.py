"""b200_cg_dot_static_f64: the dot product whose reduction scratch is a static device array.

No counterpart in the reference.  Collected last
on purpose (file name).  Called through the C-ABI
(spmv_b200.kernels.dot_static -> ctypes) and compared with the existing entry point
b200_cg_dot_f64 -- same kernel, grid and summation order, so the bar is bit-identity -- on two
calls in a row, which also shows that what the first call left in the shared array does no harm.
"""
import ctypes as C

import numpy as np
import pytest

torch = pytest.importorskip("torch")


def _need_gpu():
    if not torch.cuda.is_available():
        pytest.skip("needs a GPU")


def test_symbol_exported_and_limit_declared():
    import os
    import re
    from spmv_b200 import _lib
    assert hasattr(_lib.lib(), "b200_cg_dot_static_f64")
    hdr = open(os.path.join(os.path.dirname(__file__), "..", "include", "b200spmv.h")).read()
    assert re.search(r"#define\s+B200_CG_DOT_STATIC_MAX_N\s+\(1 << 28\)", hdr)


@pytest.mark.gpu
@pytest.mark.parametrize("n", [1, 2, 31, 4097, 1_000_003, 16_777_216])
def test_static_dot_matches_caller_scratch_dot_twice(n):
    _need_gpu()
    from spmv_b200 import kernels as K
    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    g = torch.Generator(device="cpu").manual_seed(1234 + n)
    scratch = K.new_dot_scratch(dev)
    for call in range(2):  # two calls in a row on the one shared array, different data each time
        a = (torch.rand(n, generator=g, dtype=torch.float64) - 0.5).to(dev)
        b = (torch.rand(n, generator=g, dtype=torch.float64) * 4.0 - 2.0).to(dev)
        want = torch.full((1,), float("nan"), dtype=torch.float64, device=dev)
        got = torch.full((1,), float("nan"), dtype=torch.float64, device=dev)
        K.dot(a, b, want, scratch)
        K.dot_static(a, b, got)
        torch.cuda.synchronize()
        w, h = want.item(), got.item()
        assert np.isfinite(w) and np.float64(w).tobytes() == np.float64(h).tobytes(), (n, call, w, h)
        # and against plain fp64 arithmetic, within the tolerance of tests/test_kernels_gpu.py::test_dot_deterministic
        ah, bh = a.cpu().numpy(), b.cpu().numpy()
        assert abs(h - float(np.dot(ah, bh))) <= 1e-13 * max(1.0, float(np.abs(ah * bh).sum()))


@pytest.mark.gpu
def test_static_dot_unaligned_and_empty():
    _need_gpu()
    from spmv_b200 import kernels as K
    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    base = torch.arange(1, 10_002, dtype=torch.float64, device=dev) / 1000.0
    a, b = base[1:], base[:-1]  # a is 8-byte aligned only: the scalar-load kernel
    scratch = K.new_dot_scratch(dev)
    for _ in range(2):
        want = torch.zeros(1, dtype=torch.float64, device=dev)
        got = torch.zeros(1, dtype=torch.float64, device=dev)
        K.dot(a, b, want, scratch)
        K.dot_static(a, b, got)
        torch.cuda.synchronize()
        assert want.item() == got.item() and want.item() > 0
    # n = 0: the sum of nothing
    out = torch.full((1,), 7.0, dtype=torch.float64, device=dev)
    K.dot_static(a[:0], b[:0], out)
    torch.cuda.synchronize()
    assert out.item() == 0.0


@pytest.mark.gpu
def test_static_dot_allocates_nothing_and_rejects_long_vectors():
    _need_gpu()
    from spmv_b200 import _lib
    from spmv_b200 import kernels as K
    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    n = 1 << 20
    a = torch.ones(n, dtype=torch.float64, device=dev)
    out = torch.zeros(1, dtype=torch.float64, device=dev)
    K.dot_static(a, a, out)  # first call on this device resolves the symbol address
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    torch_alloc0 = torch.cuda.memory_allocated()
    for _ in range(50):
        K.dot_static(a, a, out)
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info()
    assert out.item() == float(n)
    assert free1 >= free0 and torch.cuda.memory_allocated() == torch_alloc0
    # beyond the stated maximum: refused before anything is launched, with the limit in the message
    lib = _lib.lib()
    rc = lib.b200_cg_dot_static_f64(C.c_int32((1 << 28) + 1), C.c_void_p(a.data_ptr()), C.c_void_p(a.data_ptr()),
                                    C.c_void_p(out.data_ptr()), None)
    assert rc != 0
    msg = lib.b200_last_error()
    msg = msg.decode() if isinstance(msg, bytes) else str(msg)
    assert "B200_CG_DOT_STATIC_MAX_N" in msg
    torch.cuda.synchronize()
    assert out.item() == float(n)  # untouched
