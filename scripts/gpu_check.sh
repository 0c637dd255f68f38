#!/bin/bash
# First-contact GPU run: every test group in its own process (a faulting kernel poisons only its own CUDA
# context), each under a timeout, logs under gpurun_out/.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total,clocks.sm,clocks.max.sm --format=csv > gpurun_out/gpu.txt 2>&1
python - > gpurun_out/env.txt 2>&1 <<'PY'
import torch
print(torch.__version__, torch.cuda.is_available(), torch.cuda.device_count(), torch.cuda.get_device_name(0))
print(torch.cuda.get_device_properties(0))
PY
TGROUPS="test_library_loaded test_layout_roundtrip test_linear test_conv_fprop test_conv_dgrad test_conv_wgrad test_conv_large test_stem test_bn_forward test_conv_bn_act test_pools test_cross_entropy test_small_linear"
for g in $TGROUPS; do
  echo "=== $g" >> gpurun_out/tests.log
  timeout 300 python -m pytest tests/test_gpu_ops.py -m gpu -q -k "$g" -x --tb=short -p no:cacheprovider >> gpurun_out/tests.log 2>&1
  echo "exit $?" >> gpurun_out/tests.log
done
grep -E "^===|passed|failed|error|exit|Error|rel|assert" gpurun_out/tests.log | head -150
