#!/bin/bash
# final round evidence: all -m gpu tests + smoke, ncu captures, the default bench line and the reference arm
mkdir -p gpurun_out
rm -f gpurun_out/summary.txt
bash tools_gpu_trip.sh
bash tools/trip_prof.sh
timeout 900 python bench.py 2>gpurun_out/bench.err | tail -n 1 > gpurun_out/bench.json; echo "bench exit $?"
timeout 900 python bench.py --impl reference --steps 3 --warmup 1 2>gpurun_out/bench_ref.err | tail -n 1 > gpurun_out/bench_ref.json; echo "bench ref exit $?"
python -c "
import json;d=json.load(open('gpurun_out/bench.json'));print(d['value'],d['ms_per_step'],d['e2e']['value'],d['roofline']['frac'],d.get('cpu_baseline'));print(open('gpurun_out/bench_ref.json').read()[:600])"
