#!/usr/bin/env bash
# Reference arm: try to configure (and, if that works, build) the unmodified reference for sm_100 from a copy under
# /tmp (the source tree is read-only).  Everything is offline; the log goes to baseline/build.log.
# Usage: baseline/build_ref.sh [configure|build]      (default: configure)
set -u
MODE="${1:-configure}"
HERE="$(cd "$(dirname "$0")" && pwd)"
LOG="$HERE/build.log"
SRC=/tmp/pdb200_ref_src
BLD=/tmp/pdb200_ref_build
{
  echo "== $(date -u) build_ref.sh $MODE"
  echo "== pip install attempt (as the task statement prescribes)"
  rm -rf /tmp/pdb200_ref_pip && cp -r /root/reference /tmp/pdb200_ref_pip 2>/dev/null
  ( cd /tmp && timeout 600 python -m pip install --no-index --no-build-isolation --find-links /opt/wheelhouse \
      --target "$HERE/_ref" /tmp/pdb200_ref_pip 2>&1 | tail -25 )
  echo "== pip --no-deps attempt"
  ( cd /tmp && timeout 600 python -m pip install --no-index --no-build-isolation --no-deps --find-links /opt/wheelhouse \
      --target "$HERE/_ref" /tmp/pdb200_ref_pip 2>&1 | tail -25 )
  rm -rf /tmp/pdb200_ref_pip
  echo "== cmake configure attempt"
  rm -rf "$SRC" "$BLD" && cp -r /root/reference "$SRC" && mkdir -p "$BLD" && cd "$BLD" && \
  timeout 1200 cmake "$SRC" -GNinja -DCMAKE_BUILD_TYPE=Release -DWITH_GPU=ON -DWITH_DISTRIBUTE=ON -DWITH_NCCL=ON \
      -DCUDA_ARCH_NAME=Manual -DCUDA_ARCH_BIN=100 -DWITH_TESTING=OFF -DWITH_MKL=OFF -DWITH_MKLDNN=OFF \
      -DWITH_PYTHON=ON -DPY_VERSION=3.12 -DON_INFER=OFF -DWITH_CINN=OFF -DWITH_XBYAK=OFF -DWITH_PSCORE=OFF 2>&1 | tail -60
  RC=$?
  echo "== cmake configure rc=$RC"
  if [ "$MODE" = build ] && [ $RC -eq 0 ]; then
    timeout 36000 ninja -j"$(nproc)" 2>&1 | tail -40
    echo "== ninja rc=$?"
  fi
} > "$LOG" 2>&1
tail -5 "$LOG"
