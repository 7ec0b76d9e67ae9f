#!/bin/bash
mkdir -p gpurun_out
timeout 400 python -m pytest tests/test_gpu_train.py -m gpu -q -x 2>&1 | tail -3
for w in "1,1" "1.1,1.0" "1.1,1.1" "1.0,1.15"; do
  echo "== NB_WGRAD_W=$w"; NB_WGRAD_W=$w timeout 200 python tools/wgrad_skew.py 1024 192 2>&1 | grep -A3 "kernel span" | cut -c1-120
  NB_WGRAD_W=$w timeout 200 python tools/wgrad_skew.py 1024 64 2>&1 | grep "kernel span"
done
NB_WGRAD_W=1,1 timeout 200 python tools/wgrad_skew.py 1024 192 2>&1 | tail -37
