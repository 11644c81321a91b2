set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
python -m pytest tests/test_parity_gpu.py -m gpu -x -q -k "poisson" 2>&1 | tail -5 > gpurun_out/r1r_pytest.txt
cat gpurun_out/r1r_pytest.txt
python bench.py --no-cpu 2>gpurun_out/r1r_bench_poisson.err | tail -1 > gpurun_out/r1r_bench_poisson_4096.json
python bench.py --no-cpu --kind LMGPU 2>/dev/null | tail -1 > gpurun_out/r1r_bench_poisson_4096_kindLMGPU.json
python bench.py --no-cpu --double 2>/dev/null | tail -1 > gpurun_out/r1r_bench_poisson_4096_double.json
