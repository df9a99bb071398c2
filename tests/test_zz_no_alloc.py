"""pdb200_l2norm_multi_static: the squared-L2-norm entry point that keeps its scratch inside the library.

Collected last on purpose (file name): it must pass both on its own and after every other GPU test has used the
library, because the scratch array is per-device state that earlier calls leave dirty.  The entry point is driven
through the C ABI (ctypes) exactly as a non-Python host would, and compared with pdb200_l2norm_multi."""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu


def _dll():
    from paddle_b200 import lib
    return lib.kernels.require()  # the raw ctypes.CDLL; raises if the library was not built


def _both(dll, tensors, out_static, out_ref, accumulate):
    from paddle_b200.ops.multi_tensor import grads_table
    tab = grads_table(tensors)
    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    fn = dll.pdb200_l2norm_multi_static
    fn.restype = ctypes.c_int
    fn.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p]
    rc = fn(tab.table_ptr, tab.chunks_ptr, tab.n_chunks, out_static.data_ptr(), int(accumulate), st)
    assert rc == 0, _err(dll)
    ref = dll.pdb200_l2norm_multi
    ref.restype = ctypes.c_int
    ref.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_int64, ctypes.c_void_p,
                    ctypes.c_int, ctypes.c_void_p]
    scratch = torch.empty(max(tab.n_chunks, 1), dtype=torch.float32, device="cuda")
    rc = ref(tab.table_ptr, tab.chunks_ptr, tab.n_chunks, scratch.data_ptr(), scratch.numel(), out_ref.data_ptr(),
             int(accumulate), st)
    assert rc == 0, _err(dll)
    torch.cuda.synchronize()
    return tab


def _err(dll):
    dll.pdb200_last_error.restype = ctypes.c_char_p
    return (dll.pdb200_last_error() or b"").decode()


def test_static_scratch_matches_existing_entry_point_twice():
    dll = _dll()
    torch.manual_seed(7)
    big = [torch.randn(3_000_001, device="cuda", dtype=torch.bfloat16), torch.randn(70_000, device="cuda"),
           torch.randn(129, 513, device="cuda", dtype=torch.float16)]
    small = [torch.randn(1000, device="cuda")]  # fewer chunks than the first call: stale partials must not leak in
    for tensors in (big, small):
        a = torch.full((1,), 123.0, device="cuda")
        b = torch.full((1,), 123.0, device="cuda")
        _both(dll, tensors, a, b, accumulate=False)
        want = sum(t.float().pow(2).sum() for t in tensors)
        torch.testing.assert_close(a, b, rtol=0, atol=0)  # same kernels, same partial order: bit-identical
        torch.testing.assert_close(a[0], want, rtol=2e-4, atol=1e-3)
    # accumulate adds to what is already in out
    a = torch.full((1,), 5.0, device="cuda")
    b = torch.full((1,), 5.0, device="cuda")
    _both(dll, small, a, b, accumulate=True)
    torch.testing.assert_close(a, b, rtol=0, atol=0)
    torch.testing.assert_close(a[0], 5.0 + small[0].pow(2).sum(), rtol=2e-4, atol=1e-3)


def test_static_scratch_rejects_oversize():
    dll = _dll()
    dll.pdb200_l2norm_multi_static_max_chunks.restype = ctypes.c_int
    cap = dll.pdb200_l2norm_multi_static_max_chunks()
    assert cap >= 1 << 16
    out = torch.zeros(1, device="cuda")
    fn = dll.pdb200_l2norm_multi_static
    fn.restype = ctypes.c_int
    fn.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p]
    rc = fn(None, None, cap + 1, out.data_ptr(), 0, None)
    assert rc != 0
    assert "at most" in _err(dll)
    torch.cuda.synchronize()
    assert float(out) == 0.0  # nothing was launched
