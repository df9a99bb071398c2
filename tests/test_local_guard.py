"""The local-memory guard of the tensor-core launchers: operands that overlap a registered mapping of
another GPU's memory are refused before any tensor map is built or any kernel is launched.  Host-only logic, so the
refusal path runs on the CPU box too."""
import ctypes

import pytest

from paddle_b200 import lib

pytestmark = pytest.mark.skipif(not lib.have_kernels(), reason="native kernel library not built")


def test_foreign_range_is_refused_and_released():
    k = lib.kernels.require()
    base, nbytes = 0x7F0000000000, 1 << 30
    assert k.pdb200_check_local(ctypes.c_void_p(base + 4096), 1024) == 0
    assert k.pdb200_register_foreign_range(base, nbytes) == 0
    try:
        assert k.pdb200_check_local(ctypes.c_void_p(base + 4096), 1024) != 0
        assert b"another GPU" in k.pdb200_last_error()
        # a buffer that merely ends inside the range, and one that starts at its last byte
        assert k.pdb200_check_local(ctypes.c_void_p(base - 512), 1024) != 0
        assert k.pdb200_check_local(ctypes.c_void_p(base + nbytes - 1), 16) != 0
        # neighbours on both sides are fine
        assert k.pdb200_check_local(ctypes.c_void_p(base - 1024), 1024) == 0
        assert k.pdb200_check_local(ctypes.c_void_p(base + nbytes), 1024) == 0
    finally:
        k.pdb200_unregister_foreign_range(base)
    assert k.pdb200_check_local(ctypes.c_void_p(base + 4096), 1024) == 0


def test_gemm_refuses_foreign_operand_before_touching_the_gpu():
    """Both GEMM entry points return an error for an A / B / C pointer inside a foreign range; nothing is launched
    (this test has no GPU)."""
    k = lib.kernels.require()
    base, nbytes = 0x7E0000000000, 1 << 28
    k.pdb200_register_foreign_range(base, nbytes)
    try:
        for fn in (k.pdb200_gemm_bf16, k.pdb200_gemm_bf16_pair):
            rc = fn(ctypes.c_void_p(base + 256), ctypes.c_void_p(base + (1 << 20)), ctypes.c_void_p(base + (2 << 20)), None,
                    256, 256, 64, 0, 0, 1, 0.0, 0, None)
            assert rc != 0
            assert b"another GPU" in k.pdb200_last_error()
    finally:
        k.pdb200_unregister_foreign_range(base)
