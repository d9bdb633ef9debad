#!/bin/bash
set -u
mkdir -p gpurun_out
python -m huggingface_sagemaker_tensorflow_distributed_b200.build > gpurun_out/c4_build.log 2>&1
timeout 600 python -m pytest tests/test_allreduce_emulated_gpu.py tests/test_zz_no_alloc.py -x -q > gpurun_out/c4_tests.log 2>&1
echo "tests rc=$?" >> gpurun_out/c4_tests.log
tail -15 gpurun_out/c4_tests.log
