#!/bin/bash
# run every -m gpu test file in its own process (a trap in one kernel must not poison the rest)
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
for f in tests/test_gpu_kernels.py tests/test_gpu_mlp.py tests/test_gpu_render.py; do
  b=$(basename $f .py)
  timeout 600 python -m pytest $f -m gpu -q -x --tb=short -s > gpurun_out/$b.log 2>&1
  echo "$b exit $?" | tee -a gpurun_out/summary.txt
  tail -5 gpurun_out/$b.log
done
timeout 300 python -m pytest tests/test_gpu_mlp.py -m gpu -q --tb=short -s > gpurun_out/test_gpu_mlp_all.log 2>&1
echo "mlp_all exit $?" | tee -a gpurun_out/summary.txt
timeout 300 python -m pytest tests/test_gpu_kernels.py -m gpu -q --tb=short > gpurun_out/test_gpu_kernels_all.log 2>&1
echo "kernels_all exit $?" | tee -a gpurun_out/summary.txt
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1
echo "smoke exit $?" | tee -a gpurun_out/summary.txt
timeout 300 python tools/time_fwd.py > gpurun_out/time_fwd.log 2>&1
echo "time_fwd exit $?" | tee -a gpurun_out/summary.txt
cat gpurun_out/time_fwd.log
