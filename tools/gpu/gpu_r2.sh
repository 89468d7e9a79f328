#!/bin/bash
# round-2 iteration pass: the no-alloc test alone in a fresh process, the whole GPU suite, smoke(), then the K1 A/B
# (base library + every gamer_b200/lib/variants/lib_*.so) on both bench views
mkdir -p gpurun_out
python -m pytest tests/test_zz_no_alloc.py -m gpu -q > gpurun_out/r2_noalloc_alone.log 2>&1; echo "noalloc alone rc=$?"; tail -2 gpurun_out/r2_noalloc_alone.log
python -m pytest tests -m gpu -q -x > gpurun_out/r2_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/r2_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2_smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/r2_smoke.log
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
