#!/bin/bash
# all-reduce variants at N GPUs: weak vs .sys accesses, CTA count
N=${1:-2}
for weak in 0 1; do for blocks in 64 96 128; do
  echo "== weak=$weak blocks=$blocks"
  PTDB200_AR_WEAK=$weak PTDB200_AR_BLOCKS=$blocks timeout 120 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2956$weak benchmarks/allreduce_sweep.py --max-mb 256 2>&1 | grep -E "(16777216|67108864|268435456) B" | cut -c1-150
done; done
