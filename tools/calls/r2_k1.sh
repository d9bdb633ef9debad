#!/bin/bash
# 1 GPU kernel iteration: kernel tests + engine tests + bench (+ optional env A/B)
set -u
mkdir -p gpurun_out
TAG=${1:-k1}
python -m huggingface_sagemaker_tensorflow_distributed_b200.build > gpurun_out/${TAG}_build.log 2>&1
timeout 600 python -m pytest tests/test_kernels_gpu.py tests/test_engine_gpu.py tests/test_gemm_gpu.py tests/test_attention_gpu.py -x -q > gpurun_out/${TAG}_tests.log 2>&1
echo "tests rc=$?" >> gpurun_out/${TAG}_tests.log
tail -6 gpurun_out/${TAG}_tests.log
timeout 200 python bench.py --steps 100 --warmup 5 > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err
python -c "
import json;d=json.loads(open('gpurun_out/${TAG}_bench.json').read().strip().splitlines()[-1]);print('bench', d['ms_per_step'],'ms', d['value'],'samples/s e2e',d['e2e']['value'],'launches/step',d['gpu_launches']/d['steps'])" || tail -5 gpurun_out/${TAG}_bench.err
for v in ${ENVS:-}; do
  env $v timeout 200 python bench.py --steps 100 --warmup 5 2>/dev/null | python -c "
import json,sys;d=json.loads(sys.stdin.read().strip().splitlines()[-1]);print('$v', d['ms_per_step'],'ms', d['value'])"
done
