#!/bin/bash
# multi-GPU run: parity tests, then frames (weak) and tiles (strong) benches at N in "$@"
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/mg_gpus.txt
timeout 900 python -m pytest tests/test_gpu_multi.py -x -q -m gpu > gpurun_out/mg_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/mg_pytest.log
tail -3 gpurun_out/mg_pytest.log
for n in "$@"; do
  timeout 600 python bench.py --gpus $n --steps 10 --warmup 3 --no-cpu > gpurun_out/mg_frames_n$n.json 2> gpurun_out/mg_frames_n$n.err; echo "frames n=$n rc=$?"
  timeout 600 python bench.py --gpus $n --steps 6 --warmup 3 --no-cpu --mode tiles --workload sombrero_4096_edge > gpurun_out/mg_tiles_n$n.json 2> gpurun_out/mg_tiles_n$n.err; echo "tiles n=$n rc=$?"
done
python - "$@" <<'PY'
import json, sys
for n in sys.argv[1:]:
    for m in ("frames", "tiles"):
        try:
            d = json.loads(open(f"gpurun_out/mg_{m}_n{n}.json").read().strip().splitlines()[-1])
            print(m, "n", n, "Gs/s", round(d["value"]/1e9, 2), "ms/step", round(d["ms_per_step"], 2), "e2e Gs/s", round(d["e2e"]["value"]/1e9, 2))
        except Exception as e:
            print(m, n, "failed", e)
PY
