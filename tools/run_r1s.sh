set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
python -m pytest tests/test_parity_gpu.py -m gpu -x -q -k "arap" 2>&1 | tail -5 > gpurun_out/r1s_pytest.txt
cat gpurun_out/r1s_pytest.txt
python bench.py --workload arap_1m --no-cpu 2>gpurun_out/r1s_bench_arap.err | tail -1 > gpurun_out/r1s_bench_arap_1m.json
OPTB200_GATHER_UNFUSED=1 python bench.py --workload arap_1m --no-cpu 2>/dev/null | tail -1 > gpurun_out/r1s_bench_arap_1m_unfused.json
