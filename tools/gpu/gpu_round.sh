#!/bin/bash
# full GPU pass: tests, benches (both arms), then ONE ncu launch list of the bench command
mkdir -p gpurun_out
python -m pytest tests -m gpu -q > gpurun_out/r_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r_pytest.log; tail -4 gpurun_out/r_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r_smoke.log 2>&1; echo "smoke rc=$?"
python bench.py --steps 10 --warmup 3 > gpurun_out/r_bench_n1.json 2> gpurun_out/r_bench.err; echo "bench rc=$?"
python bench.py --steps 5 --warmup 3 --no-cpu --workload spiral_2048_cli_view > gpurun_out/r_bench_cli.json 2>> gpurun_out/r_bench.err
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r_bench_ref.json 2>> gpurun_out/r_bench.err
CMD="python bench.py --steps 2 --warmup 3 --no-cpu"
$CMD > gpurun_out/r_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/r_launches.csv $CMD > gpurun_out/r_ncu1.log 2>&1
echo "ncu rc=$?"
python - <<'PY'
import json
for f in ("r_bench_n1","r_bench_cli"):
    d=json.loads(open(f"gpurun_out/{f}.json").read().strip().splitlines()[-1])
    print(f, "ms/frame", round(d["ms_per_step"],3), "Gs/s", round(d["value"]/1e9,2), "frac", round(d["roofline"]["frac"],4), "e2e ms", round(d["e2e"]["ms_per_step"],3), d.get("cpu_baseline",{}).get("value"))
PY
