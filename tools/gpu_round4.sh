#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/pytest_gpu.log
bash tools/bench_all.sh > gpurun_out/bench_all.txt 2>&1; cat gpurun_out/bench_all.txt
