#!/bin/bash
# ncu source-page capture of one gemm_ln_kernel launch (O-projection = even launches, down-projection = odd)
SKIP=${1:-4}; OUT=${2:-r02_gemm_ln}
ncu --set full --import-source on --clock-control none -k "regex:gemm_ln_kernel" --launch-skip "$SKIP" --launch-count 1 -f -o /tmp/$OUT \
    python tools/batch_sweep.py --batches 64 --iters 1 > gpurun_out/$OUT.ncu.log 2>&1
python tools/ncu_hot.py /tmp/$OUT.ncu-rep 40 > gpurun_out/$OUT.hot.txt 2>&1
python tools/ncu_hot.py /tmp/$OUT.ncu-rep window 520 800 > gpurun_out/$OUT.window.txt 2>&1
python tools/summarize_ncu.py full /tmp/$OUT.ncu-rep gpurun_out/$OUT.full.txt
