#!/bin/bash
# ncu --set full capture (with source) of ONE GEMM launch of a config-2 step, summarised on the box:
#   tools/prof_one_gemm.sh <epilogue 0..3> <launches of that epilogue to skip> <out prefix under gpurun_out/>
# epilogue: 0 STORE32, 1 STORE16, 2 SWIGLU16, 3 ROPE16 (QKV).  Per step: 40 / 1 / 14 / 14 launches; 3 warm-up steps.
EPI=${1:-3}; SKIP=${2:-49}; OUT=${3:-gemm_epi$EPI}
ncu --set full --import-source on --clock-control none --kernel-name-base demangled \
    -k "regex:gemm_tcgen05_kernel<\(int\)256, \(int\)$EPI" --launch-skip "$SKIP" --launch-count 1 -f -o /tmp/$OUT \
    python tools/batch_sweep.py --batches 64 --iters 1 > gpurun_out/$OUT.ncu.log 2>&1
python tools/ncu_hot.py /tmp/$OUT.ncu-rep 60 > gpurun_out/$OUT.hot.txt 2>&1
python tools/summarize_ncu.py full /tmp/$OUT.ncu-rep gpurun_out/$OUT.full.txt
ncu -i /tmp/$OUT.ncu-rep --page details --csv 2>/dev/null | cut -d, -f12- | grep -i -E "stall|Issue Slot|Hit Rate|Eligible|Occupancy|Busy|Bandwidth|IPC|Throughput|Duration|Cycles" | cut -c1-200 > gpurun_out/$OUT.details.txt
