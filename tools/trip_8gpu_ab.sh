#!/bin/bash
mkdir -p gpurun_out
N=8
for sp in 0 1; do
NB_SPLIT_AR=$sp timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2966$sp bench.py --gpus $N --steps 40 --warmup 5 --no-render 2>gpurun_out/bench${N}_sp$sp.err | tail -n 1 > gpurun_out/bench_${N}gpu_sp$sp.json; echo "bench$N split=$sp exit $?"
python -c "
import json;d=json.load(open('gpurun_out/bench_${N}gpu_sp$sp.json'));print('split=$sp',d['value'],d['ms_per_step'],d['e2e']['value'])" || tail -30 gpurun_out/bench${N}_sp$sp.err
done
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29677 bench.py --gpus $N --steps 30 --warmup 5 2>gpurun_out/bench$N.err | tail -n 1 > gpurun_out/bench_${N}gpu.json; echo "bench$N exit $?"
python -c "
import json;d=json.load(open('gpurun_out/bench_${N}gpu.json'));print(d['value'],d['ms_per_step'],d['e2e']['value'],d.get('render'),d.get('large_batch'))" || tail -30 gpurun_out/bench$N.err
