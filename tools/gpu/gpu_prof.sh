#!/bin/bash
# one ncu --set full capture of kernel $KERN (default k1_march) from the bench command, after a plain run exited 0
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --no-cpu --no-extras ${WL:+--workload $WL}"
$CMD > gpurun_out/p_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:${KERN:-k1_march} -s ${SKIP:-3} -c 1 -f -o gpurun_out/${OUT:-prof_k1} $CMD > gpurun_out/p_ncu.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/p_ncu.log
