#!/bin/bash
# A/B of two builds of the library inside ONE gpurun call: mio_tts_cpp_b200/libmio_tts_b200_prev.so (built from the previous
# commit: git stash; make; cp; git stash pop; make) against the current one, alternated ROUNDS times.
ROUNDS=${1:-3}; shift
for i in $(seq 1 "$ROUNDS"); do
  for v in prev cur; do
    if [ $v = prev ]; then export MIO_TTS_B200_LIB=$PWD/mio_tts_cpp_b200/libmio_tts_b200_prev.so; else unset MIO_TTS_B200_LIB; fi
    r=$(python tools/batch_sweep.py --batches 64 --iters 30 "$@" 2>/dev/null | tail -1 | python -c 'import json,sys; d=json.loads(sys.stdin.read()); print("%.3f ms  gemm %.0f TF/s" % (d["ms_per_step"], d["gemm_tflops"]))')
    echo "$v  $r"
  done
done
