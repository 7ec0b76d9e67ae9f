#!/bin/bash
# 2-GPU check: N-rank == 1-rank equivalence test, then the bench line at N=2 with and without the split all-reduce
mkdir -p gpurun_out
N=${1:-2}
timeout 600 python -m pytest tests/test_gpu_dist.py -m gpu -q --tb=short -x > gpurun_out/t_dist.log 2>&1; echo "dist tests exit $?"; tail -3 gpurun_out/t_dist.log
for sp in 0 1 0 1; do
NB_SPLIT_AR=$sp timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2965$sp bench.py --gpus $N --steps 40 --warmup 5 --no-render 2>gpurun_out/bench${N}_sp$sp.err | tail -n 1 > gpurun_out/bench_${N}gpu_sp$sp.json; echo "bench$N split=$sp exit $?"
python -c "
import json;d=json.load(open('gpurun_out/bench_${N}gpu_sp$sp.json'));print('split=$sp',d['value'],d['ms_per_step'],d['e2e']['value'])" || tail -30 gpurun_out/bench${N}_sp$sp.err
done
