#!/bin/bash
mkdir -p gpurun_out
for t in mlp render train; do
timeout 300 python -m pytest tests/test_gpu_$t.py -m gpu -q --tb=short -x > gpurun_out/t_$t.log 2>&1; echo "$t exit $?"; tail -4 gpurun_out/t_$t.log
done
timeout 300 python tools/time_fwd.py > gpurun_out/time_fwd.log 2>&1; echo "time exit $?"; tail -12 gpurun_out/time_fwd.log
