#!/bin/bash
# the driver's round-end sequence on one GPU: reference arm, then our arm (default flags), wall-clocked
mkdir -p gpurun_out
( time python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/full_ref.json 2> gpurun_out/full_ref.err ) 2>&1 | grep real
( time python bench.py > gpurun_out/full_n1.json 2> gpurun_out/full_n1.err ) 2>&1 | grep real
tail -3 gpurun_out/full_n1.err
python - <<'PY'
import json
d=json.loads(open("gpurun_out/full_n1.json").read().strip().splitlines()[-1])
r=json.loads(open("gpurun_out/full_ref.json").read().strip().splitlines()[-1])
print("ms/frame", round(d["ms_per_step"],3), "Gs/s", round(d["value"]/1e9,2), "frac", round(d["roofline"]["frac"],4), "k1", round(d["kernel_ms"]["k1_avg"],3))
print("e2e", round(d["e2e"]["ms_per_step"],3), "ctx", d["e2e"]["contexts"], "| serial", round(d["e2e_serial"]["ms_per_step"],3))
print("k2", d["roofline_k2"]["launch_ms"], round(d["roofline_k2"]["frac"],3))
print("cpu", d["cpu_baseline"]["value"], d["cpu_baseline"]["cores"], d["cpu_baseline"]["kind"], "port", d["cpu_baseline"].get("oracle_port"))
print("parity", d["parity"])
for k,v in d["also"].items(): print(k, {a:(round(b,3) if isinstance(b,float) else b) for a,b in v.items() if a in ("value","unit","ms_per_step","ms_total","rank_ms_max","rank_ms_mean","n_units")}, "e2e", (v.get("e2e") or {}).get("ms_per_step", (v.get("e2e") or {}).get("ms_total")))
print("ref arm", r["value"], r["cpu_baseline"]["cores"], r["product_library_loaded"], "same config", r["config"]==d["config"])
PY
