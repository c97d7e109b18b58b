#!/bin/bash
# ncu --set full capture (with source) of ONE launch of a non-GEMM kernel of a config-2 step, summarised on the box:
#   tools/prof_one_kernel.sh <kernel name regex> <launches to skip> <out prefix under gpurun_out/>
K=${1:-attention_band}; SKIP=${2:-49}; OUT=${3:-kern}
ncu --set full --import-source on --clock-control none -k "regex:$K" --launch-skip "$SKIP" --launch-count 1 -f -o /tmp/$OUT \
    python tools/batch_sweep.py --batches 64 --iters 1 > gpurun_out/$OUT.ncu.log 2>&1
python tools/ncu_hot.py /tmp/$OUT.ncu-rep 60 > gpurun_out/$OUT.hot.txt 2>&1
python tools/summarize_ncu.py full /tmp/$OUT.ncu-rep gpurun_out/$OUT.full.txt
