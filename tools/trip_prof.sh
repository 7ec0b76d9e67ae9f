#!/bin/bash
# ncu evidence trip: full capture of the MLP kernels and of the streaming kernels, launch list of the bench step
mkdir -p gpurun_out
timeout 120 python tools/prof_kernels.py > gpurun_out/prof_plain.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:mlp_ -s 4 -c 5 -f -o gpurun_out/prof_mlp python tools/prof_kernels.py > gpurun_out/ncu_full.log 2>&1
echo "ncu mlp exit $?"
timeout 120 python tools/bench_stream.py 65536 > gpurun_out/bench_stream_small.log 2>&1 && \
NB_STREAM_REPS=1 NB_STREAM_WARMUP=0 timeout 600 ncu --set full --clock-control none --import-source on -k regex:'composite_|sample_pdf_merge|sample_coarse|ray_points' -s 3 -c 12 -f -o gpurun_out/prof_stream python tools/bench_stream.py 262144 > gpurun_out/ncu_stream.log 2>&1
echo "ncu stream exit $?"
timeout 300 python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-render > gpurun_out/bench_short.json 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-render > gpurun_out/ncu_launches.log 2>&1
echo "ncu launches exit $?"
