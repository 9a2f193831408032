#!/bin/bash
# usage (under gpurun --gpus N): bash tools/run_gpu_scale_workloads.sh N   -- full Multiface (1 M rays) and DTU frames sharded over N GPUs
set -x
N=${1:-8}
cd ${GRAFT_REPO_ROOT:-.}
mkdir -p gpurun_out
for w in multiface dtu; do
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus $N --steps 5 --warmup 3 --workload $w --no-extra > gpurun_out/scale${N}_$w.json 2> gpurun_out/scale${N}_$w.err
done
tail -c 300 gpurun_out/scale${N}_multiface.err
