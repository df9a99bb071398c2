#!/usr/bin/env bash
# Test tiers: the same split the driver uses.
#   tools/run_tests.sh cpu            everything that needs no GPU (multi-process paths over gloo included)
#   tools/run_tests.sh gpu            single-GPU kernel / model tests (on a B200:  gpurun -- tools/run_tests.sh gpu)
#   tools/run_tests.sh multigpu N     fused NVLink paths against NCCL on N >= 2 GPUs
#   tools/run_tests.sh sanitize       TSAN + ASAN builds of the host runtime (CPU), see tools/sanitize.sh for the GPU tools
#   tools/run_tests.sh perf FILE      perf-regression gate of a fresh bench line against profiles/
set -eu
cd "$(dirname "$0")/.."
case "${1:-cpu}" in
  cpu) python -m pytest tests/ -x -q -m "not gpu" ;;
  gpu) python -m pytest tests/ -x -q -m gpu ;;
  multigpu) python -m pytest tests/test_nvl_gpu.py -x -q ;;
  sanitize) tools/sanitize.sh tsan && tools/sanitize.sh asan ;;
  perf) shift; python tools/perf_check.py "$@" --baseline profiles/bench_n1.json --baseline profiles/bench_n8.json ;;
  *) echo "usage: $0 cpu|gpu|multigpu|sanitize|perf"; exit 2 ;;
esac
