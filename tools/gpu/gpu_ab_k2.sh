#!/bin/bash
# A/B of K2 (post-process) across library builds: cold-L2 launch time at 2048^2 and 4096^2
mkdir -p gpurun_out
python -m pytest tests -m gpu -q -x -k "tonemap or rasterizer_api or fixture_cli" 2>&1 | tail -2
for f in "" gamer_b200/lib/variants/lib_*.so; do
  if [ -z "$f" ]; then unset GAMER_B200_LIB; name=base; else export GAMER_B200_LIB=$PWD/$f; name=$(basename $f .so); fi
  [ -n "$f" ] && [ ! -f "$f" ] && continue
  python bench.py --steps 6 --warmup 3 --no-cpu > gpurun_out/k2_$name.json 2>> gpurun_out/ab.err
  python bench.py --steps 3 --warmup 3 --no-cpu --workload sombrero_4096_edge > gpurun_out/k2_4k_$name.json 2>> gpurun_out/ab.err
  python - <<PY
import json
a=json.load(open("gpurun_out/k2_$name.json")); b=json.load(open("gpurun_out/k2_4k_$name.json"))
print("$name", "2048: k2 cold us", round(1e3*a["kernel_ms"]["k2_cold_l2"],2), "in-step", round(1e3*a["kernel_ms"]["k2_avg_in_step"],2), "frac", round(a["roofline_k2"]["frac"],3),
      "| 4096: cold us", round(1e3*b["kernel_ms"]["k2_cold_l2"],2), "frac", round(b["roofline_k2"]["frac"],3))
PY
done
tail -3 gpurun_out/ab.err
