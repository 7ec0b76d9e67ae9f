#!/bin/bash
# run every -m gpu test file in its own process (a trap in one kernel must not poison the rest)
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
for f in tests/test_gpu_kernels.py tests/test_gpu_mlp.py tests/test_gpu_render.py tests/test_gpu_train.py tests/test_gpu_pose.py tests/test_gpu_checkpoint.py; do
  b=$(basename $f .py)
  timeout 600 python -m pytest $f -m gpu -q --tb=short -s > gpurun_out/$b.log 2>&1
  echo "$b exit $?" | tee -a gpurun_out/summary.txt
  grep -E "passed|failed" gpurun_out/$b.log | tail -2
done
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1
echo "smoke exit $?" | tee -a gpurun_out/summary.txt
for s in $EXTRA_SCRIPTS; do
  b=$(basename $s .py)
  timeout 600 python $s > gpurun_out/$b.log 2>&1
  echo "$b exit $?" | tee -a gpurun_out/summary.txt
  tail -30 gpurun_out/$b.log
done
