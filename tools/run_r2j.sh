# r2j: ARAP gather on AoS solver vectors; graph benches
set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_parity_gpu.py tests/test_csr_gpu.py tests/test_ref_driver.py -m gpu -q -x -k "arap" 2>&1 | tail -5 > gpurun_out/r2j_pytest.txt
cat gpurun_out/r2j_pytest.txt
timeout 600 python bench.py --workload arap_1m --steps 5 --warmup 3 --no-cpu 2> gpurun_out/r2j_arap.err | tail -1 > gpurun_out/r2j_bench_arap_1m.json
python -c "import json;d=json.load(open('gpurun_out/r2j_bench_arap_1m.json'));print(d['value'],d['roofline']['frac'],d['roofline']['kernel_ms'],d['clocks'])"
