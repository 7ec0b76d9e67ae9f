#!/bin/bash
# streamed backward: correctness, then the 1,024-ray step with and without
mkdir -p gpurun_out
timeout 400 python -m pytest tests/test_gpu_train.py -m gpu -q -x -k "streamed" > gpurun_out/stream_test.log 2>&1
echo "rc=$?" >> gpurun_out/stream_test.log
tail -5 gpurun_out/stream_test.log
for v in 0 1 50000; do
  NB_STREAM_BWD=$v timeout 400 python bench.py --no-cpu-baseline --no-render --no-sustained --steps 50 --warmup 5 > gpurun_out/stream_bench_$v.json 2> gpurun_out/stream_bench_$v.err
  echo "NB_STREAM_BWD=$v rc=$?"; python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/stream_bench_$v.json").read().strip().splitlines()[-1])
    print(d["value"], d["ms_per_step"], d["e2e"]["value"])
except Exception as e: print("no line", e)
PY
done
for dg in 30 44; do
  wg=$((148-2*dg))
  NB_STREAM_BWD=1 NB_STREAM_DGRAD=$dg NB_STREAM_WGRAD=$wg timeout 400 python bench.py --no-cpu-baseline --no-render --no-sustained --steps 50 --warmup 5 > gpurun_out/stream_bench_d$dg.json 2> gpurun_out/stream_bench_d$dg.err
  echo "dgrad=$dg wgrad=$wg rc=$?"; tail -c 600 gpurun_out/stream_bench_d$dg.json | grep -o '"value": [0-9.]*' | head -1
done
