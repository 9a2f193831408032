set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
python -m pytest tests -m gpu -q --tb=short 2>&1 | tail -40 > gpurun_out/pytest5.log
for v in 2 4 8; do
for w in facescape dtu; do DNR_ENCODE_KERNEL=$v python tools/encode_bench.py $w > gpurun_out/enc${v}_$w.json 2>>gpurun_out/enc.err; done
DNR_ENCODE_KERNEL=$v python tools/encode_bench.py multiface 262144 > gpurun_out/enc${v}_multiface.json 2>>gpurun_out/enc.err
done
cat gpurun_out/enc?_*.json; tail -4 gpurun_out/pytest5.log
