#!/bin/bash
# A/B of one environment switch inside ONE gpurun call (boxes differ by +-5 %, so off/on must alternate on the same box):
#   tools/ab.sh VAR [off_value] [on_value] [rounds] [extra batch_sweep args]
# prints ms_per_step of the resident config-2 step for each run
VAR=$1; OFF=${2:-0}; ON=${3:-1}; ROUNDS=${4:-3}; shift 4 2>/dev/null
for i in $(seq 1 "$ROUNDS"); do
  for v in "$OFF" "$ON"; do
    r=$(env "$VAR=$v" python tools/batch_sweep.py --batches 64 --iters 30 "$@" 2>/dev/null | tail -1 | python -c 'import json,sys; d=json.loads(sys.stdin.read()); print("%.3f ms  gemm %.0f TF/s" % (d["ms_per_step"], d["gemm_tflops"]))')
    echo "$VAR=$v  $r"
  done
done
