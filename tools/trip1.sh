#!/bin/bash
# GPU trip: tests + smoke, bench, kernel timings, ncu launch list + full capture of the MLP kernels
mkdir -p gpurun_out
bash tools_gpu_trip.sh
timeout 600 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?" | tee -a gpurun_out/summary.txt
tail -c 3000 gpurun_out/bench.json
timeout 300 python tools/time_fwd.py > gpurun_out/time_fwd.log 2>&1; tail -12 gpurun_out/time_fwd.log
timeout 120 python tools/prof_kernels.py > gpurun_out/prof_plain.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:mlp_ -s 4 -c 5 -f -o gpurun_out/prof_mlp python tools/prof_kernels.py > gpurun_out/ncu_full.log 2>&1
echo "ncu full exit $?" | tee -a gpurun_out/summary.txt
timeout 300 python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-render > gpurun_out/bench_short.json 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-render > gpurun_out/ncu_launches.log 2>&1
echo "ncu launches exit $?" | tee -a gpurun_out/summary.txt
