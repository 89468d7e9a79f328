python -m pytest tests -m gpu -q 2>&1 | tail -5
python bench.py --steps 5 --warmup 3 --no-cpu > gpurun_out/it_oblique.json 2> gpurun_out/it.err
python bench.py --steps 3 --warmup 3 --no-cpu --workload spiral_2048_cli_view > gpurun_out/it_cli.json 2>> gpurun_out/it.err
python - <<'PY'
import json
for f in ("it_oblique","it_cli"):
    d=json.load(open(f"gpurun_out/{f}.json"))
    print(f, "ms/frame", round(d["ms_per_step"],3), "k1", round(d["kernel_ms"]["k1_avg"],3), "Gsamples/s", round(d["value"]/1e9,2), "fp32 frac", round(d["roofline"]["frac"],4), "e2e ms", round(d["e2e"]["ms_per_step"],3))
PY
