#!/bin/bash
# 2-GPU check: N-rank == 1-rank equivalence test, then the bench line at N=2
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_dist.py -m gpu -q --tb=short -x > gpurun_out/t_dist.log 2>&1; echo "dist tests exit $?"; tail -3 gpurun_out/t_dist.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29655 bench.py --gpus 2 --steps 30 --warmup 5 2>gpurun_out/bench2.err | tail -n 1 > gpurun_out/bench_2gpu.json; echo "bench2 exit $?"
python -c "
import json;d=json.load(open('gpurun_out/bench_2gpu.json'));print(d['value'],d['ms_per_step'],d['e2e']['value'],d.get('render'),d.get('large_batch'))" || tail -30 gpurun_out/bench2.err
