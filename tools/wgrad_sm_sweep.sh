#!/bin/bash
# does the HBM-read-bound wgrad kernel run faster per SM when it owns fewer SMs? (kernel timed alone)
mkdir -p gpurun_out
for g in 148 111 74 56 37; do
  NB_WGRAD_GRID=$g timeout 300 python bench.py --no-cpu-baseline --no-render --steps 3 --warmup 3 2>/dev/null | tail -n 1 | python -c "
import json,sys;d=json.loads(sys.stdin.read());k=d['roofline']['per_kernel'];print('wgrad grid $g:', k['mlp_wgrad+small(fine)']['ms'], 'ms; dgrad', k['mlp_dgrad(fine)']['ms'])"
done
