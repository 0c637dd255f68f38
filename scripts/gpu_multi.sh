#!/bin/bash
# Multi-GPU check of the communication runtime on an N-GPU box: what the box exposes (multicast / NVLS), the
# multi-GPU tests, the all-reduce sweep and a short bench.  Every step under its own timeout, logs in gpurun_out/.
#   scripts/gpu_multi.sh N [tests|sweep|bench|probe ...]
N=${1:-2}; shift
STEPS=${@:-probe tests sweep bench}
mkdir -p gpurun_out
export PTDB200_P2P_TIMEOUT_S=${PTDB200_P2P_TIMEOUT_S:-60}
run() { echo "=== $*"; "$@"; echo "--- exit $?"; }
for s in $STEPS; do
case $s in
probe)
  nvidia-smi topo -m > gpurun_out/topo_n$N.txt 2>&1
  NCCL_DEBUG=INFO NCCL_DEBUG_SUBSYS=INIT,ENV timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N \
    --master-addr 127.0.0.1 --master-port 29511 benchmarks/experiments/nvls_probe.py > gpurun_out/nvls_probe_n$N.log 2>&1
  echo "probe exit $?"
  grep -i -E "nvls|multicast|symm|MULTICAST|FABRIC|POSIX|pidfd|VIRTUAL" gpurun_out/nvls_probe_n$N.log | head -30
  ;;
tests)
  timeout 1500 python -m pytest tests/test_gpu_multi.py -m gpu -q -p no:cacheprovider --timeout 400 -rs \
    > gpurun_out/pytest_multi_n$N.log 2>&1
  echo "pytest exit $?"; tail -40 gpurun_out/pytest_multi_n$N.log
  ;;
sweep)
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 \
    benchmarks/allreduce_sweep.py --max-mb 256 ${SWEEP_ARGS:-} --out gpurun_out/allreduce_sweep_n$N.json > gpurun_out/allreduce_sweep_n$N.log 2>&1
  echo "sweep exit $?"; tail -16 gpurun_out/allreduce_sweep_n$N.log
  ;;
bench)
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 \
    bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/bench_n$N.log 2>&1
  echo "bench exit $?"; tail -3 gpurun_out/bench_n$N.log
  ;;
esac
done
