#!/bin/bash
cd "$(dirname "$0")/.."
for w in none pack fwd; do
  NB_COARSE_BWD_AFTER=$w NB_STEP=static python bench.py --no-cpu-baseline --no-render --steps 50 > gpurun_out/r2d_bench_static_$w.json 2> gpurun_out/r2d_bench_static_$w.err
done
NB_STEP=autograd python bench.py --no-cpu-baseline --no-render --steps 50 > gpurun_out/r2d_bench_autograd.json 2>/dev/null
NB_COARSE_BWD_AFTER=fwd NB_STEP=static python tools/timeline.py > gpurun_out/r2d_timeline_static_fwd.txt 2>/dev/null
python -m pytest tests/test_gpu_train.py -m gpu -q -x 2>&1 | tail -3
