#!/bin/bash
# what the driver runs at round end, in its order (+ the step timeline)
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/ -x -q -m gpu > gpurun_out/driver_pytest.log 2>&1; echo "pytest -m gpu exit $?"; tail -3 gpurun_out/driver_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/driver_smoke.log 2>&1; echo "smoke exit $?"; tail -2 gpurun_out/driver_smoke.log
timeout 900 python bench.py --impl reference > gpurun_out/driver_bench_ref.json 2>gpurun_out/driver_bench_ref.err; echo "ref exit $?"; tail -c 300 gpurun_out/driver_bench_ref.json
timeout 900 python bench.py > gpurun_out/driver_bench.json 2>gpurun_out/driver_bench.err; echo "bench exit $?"; tail -n 1 gpurun_out/driver_bench.json | cut -c1-300
timeout 300 python tools/timeline.py > gpurun_out/timeline.txt 2>gpurun_out/timeline.err; echo "timeline exit $?"; tail -1 gpurun_out/timeline.txt
