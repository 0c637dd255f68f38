#!/bin/bash
# compute-sanitizer recipe (ONE tool per gpurun call, smallest case; see /opt/skills/guides/B200_PROFILING.md).
#   gpurun --timeout 900 -- 'bash scripts/sanitize.sh memcheck'        (then racecheck, synccheck in calls of their own)
# The case (benchmarks/experiments/sanitize_case.py) runs every kernel family once and checks its results; it is run
# plainly first - a sanitizer is only pointed at a program that exits 0 without it.
# racecheck only sees intra-GPU shared-memory hazards; the cross-GPU flag / switch barriers of csrc/p2p.cu are covered
# by tests/test_gpu_multi.py (bit-exact sums against NCCL, the 1000-collective stress test, the dead-peer test).
TOOL=${1:-memcheck}
mkdir -p gpurun_out
timeout 300 python benchmarks/experiments/sanitize_case.py > gpurun_out/sanitize_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/sanitize_plain.log; exit 1; }
timeout 1500 compute-sanitizer --tool "$TOOL" --log-file gpurun_out/sanitizer_"$TOOL".log \
    python benchmarks/experiments/sanitize_case.py > gpurun_out/sanitize_"$TOOL"_stdout.log 2>&1
echo "sanitizer exit $?"
tail -3 gpurun_out/sanitize_"$TOOL"_stdout.log
grep -c "=========" gpurun_out/sanitizer_"$TOOL".log
tail -6 gpurun_out/sanitizer_"$TOOL".log
