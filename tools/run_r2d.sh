# r2d: fused kernel with one release site per step; reference-driver tests; ncu of the fused kernel; 512^3 default bench
set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_parity_gpu.py tests/test_ref_driver.py -m gpu -q -k "volumetric or ref" 2>&1 | tail -25 > gpurun_out/r2d_pytest.txt
cat gpurun_out/r2d_pytest.txt
CMD="python bench.py --workload volumetric_256 --steps 3 --warmup 3 --no-cpu"
$CMD 2> gpurun_out/r2d_vol256.err > gpurun_out/r2d_vol256.out &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_volumetric_apply_packed -s 5 -c 1 -f -o gpurun_out/r2d_vol256 $CMD > gpurun_out/r2d_ncu.log 2>&1
tail -1 gpurun_out/r2d_vol256.out > gpurun_out/r2d_bench_volumetric_256.json
python -c "import json;d=json.load(open('gpurun_out/r2d_bench_volumetric_256.json'));print(d['value'],d['roofline']['kernel_ms'],d['clocks'])"
timeout 900 python bench.py --steps 5 --warmup 3 2> gpurun_out/r2d_default.err | tail -1 > gpurun_out/r2d_bench_volumetric_512.json
python -c "import json;d=json.load(open('gpurun_out/r2d_bench_volumetric_512.json'));print(d['value'],d['e2e']['value'],d['roofline']['frac'],d['roofline']['frac_of_nominal_8TBs'],d['roofline']['kernel_ms'],d['clocks'],d['parity']['ok'])"
