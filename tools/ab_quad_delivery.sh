#!/bin/bash
# A/B of the quad kernel's operand delivery on ONE box: plain bulk copies + forwarding hop (DNR_QUAD_TMA2=0) vs tensor-map
# TMA with cta_group::2 completion on the even CTA's barrier (default), alternating, driver-style step counts.
cd ${GRAFT_REPO_ROOT:-.}; mkdir -p gpurun_out
for i in 1 2 3; do
for v in 0 1; do DNR_QUAD_TMA2=$v timeout 200 python bench.py --steps 20 --warmup 5 --no-extra --no-cpu-baseline 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('tma2=$v frame_ms', round(d['ms_per_step'],3), 'quad_ms', round(d['breakdown_ms']['dnr_mlp_quad_fwd']['ms_per_step'],3), 'sm_mhz', d['clocks']['sm_mhz'], 'frac', round(d['roofline']['frac'],3), 'rays/s', round(d['value']))"; done; done > gpurun_out/ab_quad_delivery.txt 2>&1
cat gpurun_out/ab_quad_delivery.txt
