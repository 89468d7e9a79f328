for v in "" m3 m3u2 m2u2 m4; do
  if [ -z "$v" ]; then unset GAMER_B200_LIB; name=base; else export GAMER_B200_LIB=$PWD/gamer_b200/lib/variants/lib_$v.so; name=$v; fi
  python bench.py --steps 4 --warmup 3 --no-cpu > gpurun_out/ab_$name.json 2>> gpurun_out/ab.err
  python bench.py --steps 2 --warmup 3 --no-cpu --workload spiral_2048_cli_view > gpurun_out/ab_cli_$name.json 2>> gpurun_out/ab.err
  python - <<PY
import json
a=json.load(open("gpurun_out/ab_$name.json")); b=json.load(open("gpurun_out/ab_cli_$name.json"))
print("$name", "oblique k1 ms", round(a["kernel_ms"]["k1_avg"],3), "cli k1 ms", round(b["kernel_ms"]["k1_avg"],3))
PY
done
