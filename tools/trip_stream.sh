#!/bin/bash
mkdir -p gpurun_out
for t in kernels render; do
timeout 600 python -m pytest tests/test_gpu_$t.py -m gpu -q --tb=short -x > gpurun_out/t_$t.log 2>&1; echo "$t tests exit $?"; tail -2 gpurun_out/t_$t.log
done
timeout 300 python tools/bench_stream.py 262144 > gpurun_out/bench_stream.log 2>&1; echo "stream exit $?"; cat gpurun_out/bench_stream.log | tail -14
