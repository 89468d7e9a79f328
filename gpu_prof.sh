#!/bin/bash
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --no-cpu ${WL:+--workload $WL}"
$CMD > gpurun_out/p_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k1_march -s 3 -c 1 -f -o gpurun_out/${OUT:-prof_k1d} $CMD > gpurun_out/p_ncu.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/p_ncu.log
