#!/bin/bash
# K2 iteration pass: the tonemap tests, the bench line's roofline_k2, one ncu --set full capture of k2_tonemap_v4
mkdir -p gpurun_out
python -m pytest tests -m gpu -q -x -k "tonemap or assemble" > gpurun_out/k2_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/k2_pytest.log
CMD="python bench.py --steps 4 --warmup 3 --no-cpu --no-extras"
$CMD > gpurun_out/k2_bench.json 2> gpurun_out/k2_bench.err || { echo "bench failed"; tail -5 gpurun_out/k2_bench.err; exit 1; }
python - <<PY
import json
d=json.load(open("gpurun_out/k2_bench.json")); print(json.dumps(d["roofline_k2"])); print(d["kernel_ms"])
PY
ncu --set full --clock-control none --import-source on -k regex:k2_tonemap -s 4 -c 1 -f -o gpurun_out/prof_k2 $CMD > gpurun_out/k2_ncu.log 2>&1; echo "ncu rc=$?"
