#!/bin/bash
# 1 GPU: full GPU test tier (incl. the emulated-rank collective tests) + bench
set -u
mkdir -p gpurun_out
python -m huggingface_sagemaker_tensorflow_distributed_b200.build > gpurun_out/c2_build.log 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/c2_tests.log 2>&1
echo "tests rc=$?" >> gpurun_out/c2_tests.log
tail -15 gpurun_out/c2_tests.log
timeout 200 python bench.py --steps 100 --warmup 5 > gpurun_out/c2_bench.json 2> gpurun_out/c2_bench.err
cat gpurun_out/c2_bench.json
