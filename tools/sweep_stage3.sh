#!/usr/bin/env bash
# Stage-3 transport knob sweep at N GPUs (run under gpurun --gpus N).  Each line of output is one bench JSON line
# prefixed by the knob setting.  Usage: tools/sweep_stage3.sh N "VAR=val VAR=val" "VAR=val" ...
N=$1; shift
mkdir -p gpurun_out
for cfg in "$@"; do
  out=$(env $cfg timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 \
        --master-port 29501 bench.py --gpus $N --steps 4 --warmup 3 --no-e2e 2>gpurun_out/sweep_last.err | tail -1)
  echo "$cfg :: $(echo "$out" | python -c 'import sys,json
try:
    d=json.loads(sys.stdin.read()); print(d["ms_per_step"], "exposed", d.get("exposed_comm_ms"), "skipped", d.get("ms_per_step_comm_skipped"), "mhz", d["clocks"]["sm_mhz"])
except Exception as e: print("FAILED", e)')"
done
