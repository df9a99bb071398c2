#!/usr/bin/env bash
# Race / memory checking of the native code (the reference's knob is cmake/flags.cmake:129-164 SANITIZER_TYPE).
#
#   tools/sanitize.sh gpu [pytest -k expression]   single-GPU kernel tests under compute-sanitizer: memcheck, then
#                                                  racecheck and synccheck on a small subset (they are ~50x slower).
#                                                  Run on a GPU box, e.g.  gpurun -- tools/sanitize.sh gpu
#   tools/sanitize.sh tsan                         builds csrc/runtime with -fsanitize=thread into build/tsan/ and runs
#                                                  the store / watchdog / bucket-assignment tests against that library
#   tools/sanitize.sh asan                         same with -fsanitize=address,undefined
# Reports go to gpurun_out/sanitize_*.log; the exit code is non-zero when a tool reports an error.
set -u
cd "$(dirname "$0")/.."
mkdir -p gpurun_out build
MODE="${1:-gpu}"
rc=0
case "$MODE" in
  gpu)
    EXPR="${2:-norm or swiglu or rope or adamw or l2norm or cross_entropy or topk}"
    SMALL="${3:-norm_fwd or swiglu}"
    export PDB200_SANITIZE=1
    compute-sanitizer --tool memcheck --error-exitcode 9 --log-file gpurun_out/sanitize_memcheck.log \
        python -m pytest tests/test_kernels_gpu.py -x -q -k "$EXPR" > gpurun_out/sanitize_memcheck.out 2>&1 || rc=1
    tail -3 gpurun_out/sanitize_memcheck.out; grep -c "ERROR SUMMARY: 0 errors" gpurun_out/sanitize_memcheck.log || rc=1
    for tool in racecheck synccheck; do
      compute-sanitizer --tool $tool --error-exitcode 9 --log-file gpurun_out/sanitize_$tool.log \
          python -m pytest tests/test_kernels_gpu.py -x -q -k "$SMALL" > gpurun_out/sanitize_$tool.out 2>&1 || rc=1
      tail -2 gpurun_out/sanitize_$tool.out; tail -2 gpurun_out/sanitize_$tool.log
    done
    ;;
  tsan|asan)
    FLAGS="-fsanitize=thread"; LIBSAN=$(gcc -print-file-name=libtsan.so)
    if [ "$MODE" = asan ]; then FLAGS="-fsanitize=address,undefined"; LIBSAN=$(gcc -print-file-name=libasan.so); fi
    OUT=build/$MODE; mkdir -p $OUT
    CUDA_HOME="${CUDA_HOME:-/usr/local/cuda}"
    g++ -O1 -g -std=c++17 -fPIC -pthread $FLAGS -Iinclude -I$CUDA_HOME/include csrc/runtime/*.cc -shared \
        -o $OUT/libpdb200_runtime.so -L$CUDA_HOME/lib64 -lcudart -ldl || exit 2
    # the tests load the library through PDB200_RUNTIME_LIB; the sanitizer runtime has to be preloaded into python
    LD_PRELOAD="$LIBSAN" PDB200_RUNTIME_LIB="$PWD/$OUT/libpdb200_runtime.so" TSAN_OPTIONS="report_signal_unsafe=0 exitcode=66" \
    ASAN_OPTIONS="detect_leaks=0" python -m pytest tests/test_runtime_native.py -x -q -m "not gpu" > gpurun_out/sanitize_$MODE.out 2>&1 || rc=1
    tail -5 gpurun_out/sanitize_$MODE.out
    grep -c "WARNING: ThreadSanitizer\|ERROR: AddressSanitizer" gpurun_out/sanitize_$MODE.out && rc=1
    ;;
  *) echo "usage: $0 gpu|tsan|asan"; exit 2;;
esac
exit $rc
