#!/bin/bash
# usage (under gpurun --gpus N): bash tools/run_gpu_scale.sh N   -- strong-scaling bench + reference arm as the driver launches them
set -x
N=${1:-2}
cd ${GRAFT_REPO_ROOT:-.}
mkdir -p gpurun_out
python -m pytest tests/test_gpu_parity_r2.py -m gpu -q --tb=short -k "second_device" 2>&1 | tail -8 > gpurun_out/pytest_second_device.log
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/scale$N.json 2> gpurun_out/scale$N.err
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 2 --warmup 0 --impl reference > gpurun_out/scale${N}_ref.json 2> gpurun_out/scale${N}_ref.err
tail -c 400 gpurun_out/scale$N.err
