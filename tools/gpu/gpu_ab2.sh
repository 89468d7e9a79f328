#!/bin/bash
# same-box A/B of K1 builds, interleaved twice: base + every gamer_b200/lib/variants/lib_*.so, oblique and in-disk views
mkdir -p gpurun_out
for rep in 1 2; do
for f in "" gamer_b200/lib/variants/lib_*.so; do
  if [ -z "$f" ]; then unset GAMER_B200_LIB; name=base; else export GAMER_B200_LIB=$PWD/$f; name=$(basename $f .so); fi
  [ -n "$f" ] && [ ! -f "$f" ] && continue
  python bench.py --steps 6 --warmup 3 --no-cpu --no-extras > gpurun_out/ab_$name.json 2>> gpurun_out/ab.err
  python bench.py --steps 2 --warmup 3 --no-cpu --no-extras --workload spiral_2048_cli_view > gpurun_out/ab_cli_$name.json 2>> gpurun_out/ab.err
  python - <<PY
import json
a=json.load(open("gpurun_out/ab_$name.json")); b=json.load(open("gpurun_out/ab_cli_$name.json"))
print("$name", "oblique k1 ms", round(a["kernel_ms"]["k1_avg"],3), "cli k1 ms", round(b["kernel_ms"]["k1_avg"],3), "chk", a["e2e"]["checksum"], "clk", a["clocks"]["sm_mhz"])
PY
done
done
tail -3 gpurun_out/ab.err
