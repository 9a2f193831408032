#!/bin/bash
# ncu evidence of the default bench (run only after the plain command has exited 0): launch list + one full capture of the quad kernel
set -x
cd ${GRAFT_REPO_ROOT:-.}
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,clocks.mem,power.limit --format=csv > gpurun_out/smi.txt
DNR_ADAPT_SPLIT=0 timeout 300 python bench.py --steps 2 --warmup 3 --no-extra --no-cpu-baseline > gpurun_out/bench_fixed.json 2> gpurun_out/bench_fixed.err || exit 1
DNR_ADAPT_SPLIT=0 timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_eval.csv python bench.py --steps 2 --warmup 3 --no-extra --no-cpu-baseline > gpurun_out/ncu_list.log 2>&1
DNR_ADAPT_SPLIT=0 timeout 600 ncu --set full --clock-control none --import-source on -k regex:mlp_quad -s 4 -c 1 -o gpurun_out/quad_full python bench.py --steps 2 --warmup 3 --no-extra --no-cpu-baseline > gpurun_out/ncu_quad.log 2>&1
nvidia-smi --query-gpu=clocks.sm,clocks.max.sm,clocks_event_reasons.active --format=csv >> gpurun_out/smi.txt
