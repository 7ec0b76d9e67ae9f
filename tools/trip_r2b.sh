#!/bin/bash
# round-2 trip b: new parity / drop-in tests, static + pipelined step equivalence, sphere training, bench in three modes
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
python -m pytest tests/test_gpu_parity_v2.py tests/test_gpu_dropin.py -m gpu -q -s > gpurun_out/r2b_new_tests.log 2>&1
python -m pytest tests/test_gpu_train.py tests/test_gpu_kernels.py -m gpu -q -x -s -k "static or composite_train or graphed or two_branch" > gpurun_out/r2b_static.log 2>&1
python tests/golden/train_on_gpu.py 3000 > gpurun_out/r2b_train.log 2>&1
for m in autograd static pipeline; do
  NB_STEP=$m python bench.py --no-cpu-baseline --no-render --steps 50 > gpurun_out/r2b_bench_$m.json 2> gpurun_out/r2b_bench_$m.err
done
tail -3 gpurun_out/r2b_static.log
