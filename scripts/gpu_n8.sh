#!/bin/bash
# The 8-GPU evidence run: all-reduce sweep (all arms, CTA sweep of the switch kernel), the 8-rank tests, the headline
# bench with its control runs (which term owns the 1 -> 8 step-time growth), single-process DataParallel both arms.
mkdir -p gpurun_out
N=${1:-8}
export PTDB200_P2P_TIMEOUT_S=60
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 400 $TR --master-port 29512 benchmarks/allreduce_sweep.py --max-mb 256 --mc-blocks 8,16,32,64,128 \
  --out gpurun_out/allreduce_sweep_n$N.json > gpurun_out/allreduce_sweep_n$N.log 2>&1
echo "sweep exit $?"; grep " B  ours" gpurun_out/allreduce_sweep_n$N.log | cut -c1-330
bench() {  # name, env...
  name=$1; shift
  env "$@" timeout 240 $TR --master-port 29513 bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/bench_n${N}_$name.log 2>&1
  echo "bench $name exit $?: $(grep -o '"value": [0-9.]*, "ms_per_step": [0-9.]*' gpurun_out/bench_n${N}_$name.log) e2e $(grep -o '"e2e": {"value": [0-9.]*, "unit": "images/sec", "ms_per_step": [0-9.]*' gpurun_out/bench_n${N}_$name.log | grep -o 'ms_per_step.*')"
}
bench default X=1
bench no_allreduce PTDB200_DDP_SKIP_ALLREDUCE=1
bench no_bcast PTDB200_DDP_BCAST=0
bench two_shot PTDB200_AR_ALGO=two_shot
bench wire_bf16 PTDB200_DDP_WIRE=bf16
bench default2 X=1
timeout 300 python benchmarks/dp_bench.py --out gpurun_out/dp_ours_n$N.json > gpurun_out/dp_ours_n$N.log 2>&1; echo "dp ours exit $?"; tail -1 gpurun_out/dp_ours_n$N.log | cut -c1-260
timeout 300 python benchmarks/dp_bench.py --impl reference --out gpurun_out/dp_ref_n$N.json > gpurun_out/dp_ref_n$N.log 2>&1; echo "dp ref exit $?"; tail -1 gpurun_out/dp_ref_n$N.log | cut -c1-260
timeout 600 python -m pytest tests/test_gpu_multi.py -m gpu -q -p no:cacheprovider --timeout 300 -rs -k "$N-nvls or $N-ipc or model and $N" \
  > gpurun_out/pytest_multi_n$N.log 2>&1
echo "pytest exit $?"; tail -8 gpurun_out/pytest_multi_n$N.log
