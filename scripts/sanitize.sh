#!/bin/bash
# compute-sanitizer recipe (ONE tool per gpurun call, smallest case; see /opt/skills/guides/B200_PROFILING.md).
#   gpurun --timeout 900 -- 'bash scripts/sanitize.sh memcheck'
# racecheck only sees intra-GPU shared-memory hazards; the cross-GPU flag protocol of csrc/p2p.cu is covered by
# tests/test_gpu_multi.py (bit-exact sums against NCCL, 20 back-to-back calls without a host sync).
TOOL=${1:-memcheck}
mkdir -p gpurun_out
compute-sanitizer --tool "$TOOL" --log-file gpurun_out/sanitizer_"$TOOL".log \
    python -m pytest tests/test_gpu_ops.py -m gpu -q -x -k "conv_fprop and 64-64-16-3 or bn_forward and 64 or pools" -p no:cacheprovider
tail -5 gpurun_out/sanitizer_"$TOOL".log
