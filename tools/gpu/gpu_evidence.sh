#!/bin/bash
# the round's ncu evidence for the final library: launch list of the bench command, --set full captures of K1 (both views) and K2
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --no-cpu --no-extras"
$CMD > gpurun_out/e_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/e_plain.log; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/e_launches.csv $CMD > gpurun_out/e_ncu0.log 2>&1; echo "launch list rc=$?"
ncu --set full --clock-control none --import-source on -k regex:k1_march -s 3 -c 1 -f -o gpurun_out/prof_r2_k1 $CMD > gpurun_out/e_ncu1.log 2>&1; echo "k1 rc=$?"
ncu --set full --clock-control none --import-source on -k regex:k2_tonemap -s 4 -c 1 -f -o gpurun_out/prof_r2_k2 $CMD > gpurun_out/e_ncu2.log 2>&1; echo "k2 rc=$?"
CMD2="$CMD --workload spiral_2048_cli_view"
$CMD2 > gpurun_out/e_plain_cli.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k1_march -s 3 -c 1 -f -o gpurun_out/prof_r2_k1_cli $CMD2 > gpurun_out/e_ncu3.log 2>&1; echo "k1 cli rc=$?"
