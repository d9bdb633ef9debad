#!/bin/bash
# Round-2 GPU call 1 (1 GPU): validate + time the experimental attention kernels, SDPA library timing, baseline bench.
set -u
mkdir -p gpurun_out
PKG=huggingface_sagemaker_tensorflow_distributed_b200
python -m $PKG.build --experimental --watchdog > gpurun_out/c1_build.log 2>&1
timeout 400 python -m pytest tests/test_experimental_gpu.py -q > gpurun_out/c1_exp_tests.log 2>&1
echo "exp tests rc=$?" >> gpurun_out/c1_exp_tests.log
python -m $PKG.build --experimental > gpurun_out/c1_build2.log 2>&1
timeout 300 python tools/attn_bench_vs_sdpa.py > gpurun_out/c1_attn_bench.log 2>&1
echo "bench rc=$?" >> gpurun_out/c1_attn_bench.log
python -m $PKG.build > gpurun_out/c1_build3.log 2>&1
timeout 200 python bench.py --steps 100 --warmup 5 > gpurun_out/c1_bench.json 2> gpurun_out/c1_bench.err
tail -5 gpurun_out/c1_exp_tests.log; cat gpurun_out/c1_attn_bench.log; cat gpurun_out/c1_bench.json
