#!/bin/bash
# ncu --set full of k1_march for the library $LIB (a variant .so), after a plain run exited 0
mkdir -p gpurun_out
export GAMER_B200_LIB=$PWD/$LIB
CMD="python bench.py --steps 2 --warmup 3 --no-cpu --no-extras ${WL:+--workload $WL}"
$CMD > gpurun_out/p_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k1_march -s 3 -c 1 -f -o gpurun_out/${OUT:-prof_variant} $CMD > gpurun_out/p_ncu.log 2>&1
echo "ncu rc=$?"; tail -2 gpurun_out/p_ncu.log
