# r2c: fused volumetric kernel with 16-byte cp.async staging of p and z
set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_parity_gpu.py -m gpu -x -q -k "volumetric" 2>&1 | tail -5 > gpurun_out/r2c_pytest.txt
cat gpurun_out/r2c_pytest.txt
for v in fused nofuse; do
  if [ $v = nofuse ]; then export OPTB200_NO_FUSE=1; else unset OPTB200_NO_FUSE; fi
  timeout 600 python bench.py --workload volumetric_256 --steps 3 --warmup 3 --no-cpu 2> gpurun_out/r2c_vol256_$v.err | tail -1 > gpurun_out/r2c_bench_volumetric_256_$v.json
  python -c "import json;d=json.load(open('gpurun_out/r2c_bench_volumetric_256_$v.json'));print('$v',d['value'],d['roofline']['kernel_ms'],d['clocks'])"
done
