#!/bin/bash
# final round-2 evidence trip (one GPU): full GPU test suite, smoke, the bench as the driver runs it, the reference arm,
# ncu launch list of the bench command, full ncu capture of the MLP kernels, streaming-kernel bench, wgrad per-CTA trace
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
python -m pytest tests -m gpu -q > gpurun_out/r2_gputests.log 2>&1; tail -2 gpurun_out/r2_gputests.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2_smoke.log 2>&1; tail -2 gpurun_out/r2_smoke.log
python bench.py > gpurun_out/r2_bench_1gpu.json 2> gpurun_out/r2_bench_1gpu.err; echo "bench rc $?"
python bench.py --impl reference > gpurun_out/r2_bench_reference.json 2> gpurun_out/r2_bench_reference.err; echo "reference rc $?"
timeout 120 python tools/prof_kernels.py > gpurun_out/prof_plain.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:mlp_ -s 4 -c 5 -f -o gpurun_out/r2_prof_mlp python tools/prof_kernels.py > gpurun_out/ncu_full.log 2>&1
echo "ncu mlp exit $?"
ncu -i gpurun_out/r2_prof_mlp.ncu-rep --page raw --csv > gpurun_out/r2_ncu_mlp_raw.csv 2>/dev/null
timeout 300 python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-render --no-sustained > gpurun_out/bench_short.json 2>/dev/null && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/r2_launches_bench.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-render --no-sustained > gpurun_out/ncu_launches.log 2>&1
echo "ncu launches exit $?"
timeout 300 python tools/bench_stream.py > gpurun_out/r2_bench_stream.txt 2>&1; echo "stream rc $?"
timeout 100 python tools/wgrad_skew.py 1024 192 > gpurun_out/r2_wgrad_skew_fine.txt 2>&1
timeout 100 python tools/wgrad_skew.py 1024 64 > gpurun_out/r2_wgrad_skew_coarse.txt 2>&1
rm -f gpurun_out/r2_prof_mlp.ncu-rep.tmp
