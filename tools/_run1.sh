set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.limit --format=csv > gpurun_out/smi.txt
python -m pytest tests -m gpu -q -x --deselect tests/test_gpu_parity_r2.py 2>&1 | tail -15 > gpurun_out/pytest_old.log
python -m pytest tests/test_gpu_parity_r2.py -m gpu -q 2>&1 | tail -80 > gpurun_out/pytest_new.log
python tools/parity_probe.py > gpurun_out/parity_probe.json 2> gpurun_out/parity_probe.err
python bench.py --steps 10 --warmup 3 > gpurun_out/bench1.json 2> gpurun_out/bench1.err
tail -3 gpurun_out/pytest_old.log gpurun_out/pytest_new.log
