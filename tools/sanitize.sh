#!/bin/bash
# compute-sanitizer targets (SURVEY.md §5 "Race detection / sanitizers") — run on the GPU box, ONE tool per
# gpurun call, every command under `timeout`:
#     gpurun --timeout 400 -- 'bash tools/sanitize.sh memcheck'      out-of-bounds / misaligned accesses
#     gpurun --timeout 400 -- 'bash tools/sanitize.sh racecheck'     shared-memory hazards between warps
#     gpurun --timeout 400 -- 'bash tools/sanitize.sh synccheck'     invalid barrier usage
#     gpurun --gpus 2 --timeout 400 -- 'bash tools/sanitize.sh p2p'  memcheck of the fused all-reduce, 2 ranks
# Targets are the smallest shapes that still walk every code path of a kernel (all operand layouts,
# partial tiles, split-K, S=128/256 attention with padding, dropout on).  racecheck knows nothing about the
# async proxy (TMA / tcgen05 reads of smem): hazards it reports between a generic-proxy store and a later
# mbarrier-ordered bulk operation are expected and listed as such in the summary.
# NOTE (round 1): compute-sanitizer is closed on this GPU pool ("runs under it have left GPUs needing a
# reset"); the script is kept for pools where it is available.  The substitutes used here: every kernel is
# compared with a PyTorch fp32 statement of the same op on small and odd shapes (partial tiles, padding),
# outputs are pre-filled with NaN / sentinels so unwritten or over-written regions show up, and the peer-flag
# spins of the all-reduce carry a clock watchdog that turns a missing peer into an error word.
set -u
TOOL=${1:-memcheck}
OUT=gpurun_out/sanitize_${TOOL}.log
mkdir -p gpurun_out
K_SMALL='test_all_operand_layouts and 128-256-64 or test_split_k_wgrad and 2 or test_ln or test_colsum or test_head or test_embed'
case "$TOOL" in
  memcheck|racecheck|synccheck)
    timeout 300 compute-sanitizer --tool "$TOOL" --target-processes all --error-exitcode 9 \
      python -m pytest tests/test_gemm_gpu.py tests/test_kernels_gpu.py -m gpu -x -q -k "$K_SMALL" > "$OUT" 2>&1
    echo "rc(kernels)=$?" >> "$OUT"
    timeout 300 compute-sanitizer --tool "$TOOL" --target-processes all --error-exitcode 9 \
      python -m pytest tests/test_attention_gpu.py -m gpu -x -q -k "128 or 256 or dropout" >> "$OUT" 2>&1
    echo "rc(attention)=$?" >> "$OUT"
    ;;
  p2p)
    timeout 300 compute-sanitizer --tool memcheck --target-processes all --error-exitcode 9 \
      python -m pytest tests/test_multigpu.py -m gpu -x -q > "$OUT" 2>&1
    echo "rc(p2p)=$?" >> "$OUT"
    ;;
  *) echo "usage: $0 memcheck|racecheck|synccheck|p2p"; exit 2 ;;
esac
grep -E "ERROR SUMMARY|rc\(|passed|failed|Invalid|hazard" "$OUT" | tail -20
