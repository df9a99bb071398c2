"""NCCL version checks (reference: python/paddle/distributed/utils/nccl_utils.py:33-43)."""
from paddle_b200.base import core


def get_nccl_version_str(ver):
    return f"{ver // 1000}.{(ver // 100) % 10}.{ver % 100}"


def check_nccl_version_for_p2p():
    v = core.nccl_version()
    assert v == 0 or v >= 2804, f"The version of NCCL is required to be at least v2.8.4 while the version is v{get_nccl_version_str(v)}"


def check_nccl_version_for_bf16():
    return core.nccl_version() >= 21000
