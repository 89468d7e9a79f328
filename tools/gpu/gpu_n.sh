#!/bin/bash
# multi-GPU pass on one box: bit-exactness tests, then the default bench line at each N given (reference arm at the first N)
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,name --format=csv,noheader > gpurun_out/mg_gpus.txt
python -m pytest tests/test_gpu_multi.py -m gpu -q 2>&1 | tail -3
first=1
for n in "$@"; do
  if [ $first = 1 ]; then
    python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29541 bench.py --impl reference --gpus $n --steps 2 --warmup 1 > gpurun_out/n${n}_ref.json 2> gpurun_out/n${n}_ref.err; first=0
  fi
  ( time timeout 900 python bench.py --gpus $n > gpurun_out/n${n}.json 2> gpurun_out/n${n}.err ) 2>&1 | grep real
  python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/n${n}.json").read().strip().splitlines()[-1])
except Exception as e:
    print("n=$n: no line:", e); print(open("gpurun_out/n${n}.err").read()[-1500:]); raise SystemExit
print("N=$n ms/step", round(d["ms_per_step"],3), "Gs/s", round(d["value"]/1e9,2), "e2e Gs/s", round(d["e2e"]["value"]/1e9,2), "serial", round(d.get("e2e_serial",d["e2e"])["value"]/1e9,2), "launches", d["gpu_launches"])
if "parity" in d: print("  parity", {k:v for k,v in d["parity"].items() if k in ("max_err_over_bound","samples_equal","eight_bit_equal_frac","eight_bit_max_diff","rows")})
for k,v in d.get("also",{}).items():
    e=v.get("e2e") or {}
    print("  ",k, {a:(round(b,3) if isinstance(b,float) else b) for a,b in v.items() if a in ("value","unit","ms_per_step","ms_total","rank_ms_max","rank_ms_mean","n_units")}, "e2e", round(e.get("ms_per_step", e.get("ms_total",0)),3), e.get("how","")[:40], e.get("host_buffers_page_locked"))
PY
done
python - <<'PY'
import json,glob
for f in sorted(glob.glob("gpurun_out/n*_ref.json")):
    r=json.loads(open(f).read().strip().splitlines()[-1]); print(f, "ref arm Ms/s", round(r["value"]/1e6,2), "cores", r["cpu_baseline"]["cores"], "lib loaded", r["product_library_loaded"])
PY
