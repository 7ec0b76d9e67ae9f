#!/bin/bash
# final 8-GPU evidence: N-rank == 1-rank check, the full bench, one GPU of the same box right after (warm) for the ratio
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
N=${1:-8}
tr() { timeout $1 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29631 "${@:2}"; }
tr 300 tests/dist_check.py > gpurun_out/r2_dist_check_${N}gpu.log 2>&1; echo "dist_check rc $?"; grep "gradient\|replays\|DIST_CHECK" gpurun_out/r2_dist_check_${N}gpu.log
tr 900 bench.py --gpus $N --steps 50 --warmup 5 > gpurun_out/r2_bench_${N}gpu.json 2> gpurun_out/r2_bench_${N}gpu.err; echo "bench rc $?"
python bench.py --steps 50 --warmup 5 --no-render --no-sustained --no-cpu-baseline > gpurun_out/r2_bench_${N}gpu_single.json 2>/dev/null; echo "single rc $?"
