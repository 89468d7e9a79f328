#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -q -x > gpurun_out/n8_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/n8_pytest.log
python tests/tools/eight_bit_report.py > gpurun_out/eight_bit.txt 2>&1; tail -4 gpurun_out/eight_bit.txt
bash tools/gpu/gpu_n.sh "$@"
