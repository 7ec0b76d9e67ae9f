#!/bin/bash
# A/B of the graph concurrency switches (NB_OVERLAP: two-branch training graph, NB_FORK_HEADS: head gradients on a
# forked stream) + the MLP / train tests
mkdir -p gpurun_out
for t in mlp train; do
timeout 600 python -m pytest tests/test_gpu_$t.py -m gpu -q --tb=short -x > gpurun_out/t_$t.log 2>&1; echo "$t tests exit $?"; tail -3 gpurun_out/t_$t.log
done
for cfg in "1 0" "1 1" "1 0" "1 1" "0 1"; do
  set -- $cfg
  tag=ov$1_fk$2
  NB_OVERLAP=$1 NB_FORK_HEADS=$2 timeout 600 python bench.py --no-cpu-baseline --no-render --steps 40 2>gpurun_out/bench_$tag.err | tail -n 1 > gpurun_out/bench_$tag.json; echo "bench $tag exit $?"
  python -c "
import json;d=json.load(open('gpurun_out/bench_$tag.json'));print('$tag', d['value'],d['ms_per_step'],d['e2e']['value'], d['e2e']['eager_main_py_loop']['value'], d['gpu_launches'])" || tail -20 gpurun_out/bench_$tag.err
done
