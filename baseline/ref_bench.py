"""Reference arm of bench.py: the reference's own public API (paddle.distributed.fleet / group_sharded_parallel, its
own Llama-shaped layers built from paddle.nn) on the same config.  Only reachable when baseline/_ref holds an
importable build of the reference; on this image the install fails (see build_log_excerpt.txt), so bench.py prints
the `unavailable` line instead and this module is never imported."""
import json


def main(args):
    print(json.dumps({"impl": "reference", "unavailable": "baseline/_ref imports but no runnable GPU build was produced offline"}))
    return 0
