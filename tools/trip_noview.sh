#!/bin/bash
mkdir -p gpurun_out
timeout 600 python tools/fit_curve.py --impl ours --steps 600 > gpurun_out/fit_ours.txt 2>gpurun_out/fit_ours.err; echo "fit exit $?"; cat gpurun_out/fit_ours.txt; tail -3 gpurun_out/fit_ours.err
