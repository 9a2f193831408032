set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
python -m pytest tests/test_gpu_novel.py tests/test_gpu_parity_r2.py -m gpu -q --tb=short 2>&1 | tail -150 > gpurun_out/pytest3.log
python -m pytest tests -m gpu -q -x --deselect tests/test_gpu_novel.py --deselect tests/test_gpu_parity_r2.py 2>&1 | tail -5 > gpurun_out/pytest3_old.log
python tools/encode_bench.py facescape > gpurun_out/enc_facescape.json 2>gpurun_out/enc.err
python tools/encode_bench.py multiface 262144 > gpurun_out/enc_multiface.json 2>>gpurun_out/enc.err
ncu --set full --clock-control none --import-source on -k regex:encode_fwd -s 2 -c 1 -o gpurun_out/enc_facescape python tools/encode_bench.py facescape 0 1 > /dev/null 2> gpurun_out/ncu_enc.err
ncu --set full --clock-control none --import-source on -k regex:encode_fwd -s 2 -c 1 -o gpurun_out/enc_multiface python tools/encode_bench.py multiface 262144 1 > /dev/null 2>> gpurun_out/ncu_enc.err
cat gpurun_out/enc_*.json
