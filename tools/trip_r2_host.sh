#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_train.py tests/test_gpu_dropin.py tests/test_gpu_pose.py -m gpu -q -x 2>&1 | tail -3
timeout 300 python tools/prof_eager.py 2>&1 | grep "host ms/step\|300 steps"
timeout 300 python tools/gaps_eager.py 2>&1 | grep -A12 "steps: span"
