#!/bin/bash
# multi-GPU run: parity tests, then frames (weak) and tiles (strong) benches at N = $1
N=${1:-2}
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/mg_gpus.txt
timeout 900 python -m pytest tests/test_gpu_multi.py -x -q -m gpu > gpurun_out/mg_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/mg_pytest.log
tail -5 gpurun_out/mg_pytest.log
for n in 1 $N; do
  timeout 600 python bench.py --gpus $n --steps 10 --warmup 3 --no-cpu > gpurun_out/mg_frames_n$n.json 2> gpurun_out/mg_frames_n$n.err; echo "frames n=$n rc=$?"
  timeout 600 python bench.py --gpus $n --steps 6 --warmup 3 --no-cpu --mode tiles --workload sombrero_4096_edge > gpurun_out/mg_tiles_n$n.json 2> gpurun_out/mg_tiles_n$n.err; echo "tiles n=$n rc=$?"
done
tail -c 600 gpurun_out/mg_frames_n$N.json; tail -c 400 gpurun_out/mg_frames_n$N.err
tail -c 600 gpurun_out/mg_tiles_n$N.json; tail -c 400 gpurun_out/mg_tiles_n$N.err
