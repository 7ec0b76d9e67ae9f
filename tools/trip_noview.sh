#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_render.py tests/test_gpu_mlp.py -x -q -m gpu > gpurun_out/noview_pytest.log 2>&1; echo "pytest exit $?"; tail -25 gpurun_out/noview_pytest.log
