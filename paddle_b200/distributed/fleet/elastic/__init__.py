"""Elastic training (reference: python/paddle/distributed/fleet/elastic/manager.py:43-634, collective.py:28-67).
The reference keeps membership in etcd; this build keeps it in a shared-directory / TCP-store backed KV so it runs
without external services.  See manager.py."""
from .manager import ElasticManager, ElasticStatus, LauncherInterface  # noqa: F401


def enable_elastic(args, distribute_mode=None):
    return bool(getattr(args, "elastic_server", None) or str(getattr(args, "nnodes", "1")).find(":") > 0)


def launch_elastic(args, distribute_mode=None):
    from paddle_b200.distributed.launch.main import launch
    return launch()
