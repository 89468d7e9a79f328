#!/bin/bash
# A/B of library builds: base (gamer_b200/lib/libgamer_b200.so) + every gamer_b200/lib/variants/lib_*.so
mkdir -p gpurun_out
python -m pytest tests -m gpu -q -x 2>&1 | tail -3
for f in "" gamer_b200/lib/variants/lib_*.so; do
  if [ -z "$f" ]; then unset GAMER_B200_LIB; name=base; else export GAMER_B200_LIB=$PWD/$f; name=$(basename $f .so); fi
  [ -n "$f" ] && [ ! -f "$f" ] && continue
  python bench.py --steps 4 --warmup 3 --no-cpu --no-extras > gpurun_out/ab_$name.json 2>> gpurun_out/ab.err
  python bench.py --steps 2 --warmup 3 --no-cpu --no-extras --workload spiral_2048_cli_view > gpurun_out/ab_cli_$name.json 2>> gpurun_out/ab.err
  python - <<PY
import json
a=json.load(open("gpurun_out/ab_$name.json")); b=json.load(open("gpurun_out/ab_cli_$name.json"))
print("$name", "oblique k1 ms", round(a["kernel_ms"]["k1_avg"],3), "cli k1 ms", round(b["kernel_ms"]["k1_avg"],3), "chk", a["e2e"]["checksum"])
PY
done
tail -5 gpurun_out/ab.err
