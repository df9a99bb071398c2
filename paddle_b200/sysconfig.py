"""paddle.sysconfig (reference: python/paddle/sysconfig.py): where the C headers and the native libraries live.
Here: ``include/`` holds the two C-ABI headers (pdb200_kernels.h, pdb200_runtime.h), ``lib/`` the built .so files."""
import os

__all__ = ["get_include", "get_lib"]


def get_include():
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    packaged = os.path.join(os.path.dirname(os.path.abspath(__file__)), "include")
    return packaged if os.path.isdir(packaged) else os.path.join(root, "include")


def get_lib():
    return os.path.join(os.path.dirname(os.path.abspath(__file__)), "lib")
