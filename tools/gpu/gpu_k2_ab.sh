#!/bin/bash
# K2 A/B on one box: the tonemap tests on the base library, then roofline_k2 of the bench line for the base library and
# every gamer_b200/lib/variants/lib_k2_*.so
mkdir -p gpurun_out
python -m pytest tests -m gpu -q -x -k "tonemap or assemble" > gpurun_out/k2_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/k2_pytest.log
for f in "" gamer_b200/lib/variants/lib_k2_*.so; do
  if [ -z "$f" ]; then unset GAMER_B200_LIB; name=base; else [ -f "$f" ] || continue; export GAMER_B200_LIB=$PWD/$f; name=$(basename $f .so); fi
  python bench.py --steps 3 --warmup 3 --no-cpu --no-extras --e2e-contexts 1 > gpurun_out/k2ab_$name.json 2>> gpurun_out/k2ab.err
  python - <<PY
import json
d=json.load(open("gpurun_out/k2ab_$name.json")); r=d["roofline_k2"]
print("$name", "k2 cold us", round(r["launch_ms"]*1e3,2), "frac", round(r["frac"],3), "in-step us", round(d["kernel_ms"]["k2_avg_in_step"]*1e3,2), "k1", round(d["kernel_ms"]["k1_avg"],3))
PY
done
tail -3 gpurun_out/k2ab.err
