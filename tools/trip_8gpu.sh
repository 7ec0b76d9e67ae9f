#!/bin/bash
mkdir -p gpurun_out
N=${1:-8}
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29677 bench.py --gpus $N --steps 30 --warmup 5 2>gpurun_out/bench$N.err | tail -n 1 > gpurun_out/bench_${N}gpu.json; echo "bench$N exit $?"
python -c "
import json;d=json.load(open('gpurun_out/bench_${N}gpu.json'));print(d['value'],d['ms_per_step'],d['e2e']['value'],d.get('render'),d.get('large_batch'))" || tail -30 gpurun_out/bench$N.err
