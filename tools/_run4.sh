set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
python -m pytest tests -m gpu -q --tb=short 2>&1 | tail -60 > gpurun_out/pytest4.log
for v in 2 4; do
DNR_ENCODE_KERNEL=$v python tools/encode_bench.py facescape > gpurun_out/enc${v}_facescape.json 2>gpurun_out/enc.err
DNR_ENCODE_KERNEL=$v python tools/encode_bench.py multiface 262144 > gpurun_out/enc${v}_multiface.json 2>>gpurun_out/enc.err
DNR_ENCODE_KERNEL=$v python tools/encode_bench.py dtu > gpurun_out/enc${v}_dtu.json 2>>gpurun_out/enc.err
done
ncu --set full --clock-control none --import-source on -k regex:encode_fwd -s 2 -c 1 -o gpurun_out/enc2_facescape python tools/encode_bench.py facescape 0 1 > /dev/null 2> gpurun_out/ncu_enc.err
cat gpurun_out/enc?_*.json; tail -4 gpurun_out/pytest4.log
