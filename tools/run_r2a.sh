# r2a: first GPU call of round 2 -- the fused volumetric kernel (PCGStep3 folded into the packed J^T J kernel), the
# mask-change fix, the full-size parity tests and the reworked bench (default = volumetric 512^3)
set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total,clocks.sm,clocks.max.sm --format=csv
timeout 900 python -m pytest tests -m gpu -x -q --ignore=tests/test_fullsize_gpu.py 2>&1 | tail -15 > gpurun_out/r2a_pytest.txt
cat gpurun_out/r2a_pytest.txt
timeout 900 python -m pytest tests/test_fullsize_gpu.py -m gpu -q 2>&1 | tail -30 > gpurun_out/r2a_pytest_fullsize.txt
cat gpurun_out/r2a_pytest_fullsize.txt
for v in fused nofuse; do
  if [ $v = nofuse ]; then export OPTB200_NO_FUSE=1; else unset OPTB200_NO_FUSE; fi
  timeout 600 python bench.py --workload volumetric_256 --steps 3 --warmup 3 --no-cpu 2> gpurun_out/r2a_vol256_$v.err | tail -1 > gpurun_out/r2a_bench_volumetric_256_$v.json
  cat gpurun_out/r2a_bench_volumetric_256_$v.json
done
unset OPTB200_NO_FUSE
timeout 900 python bench.py --steps 3 --warmup 3 2> gpurun_out/r2a_default.err | tail -1 > gpurun_out/r2a_bench_volumetric_512.json
cat gpurun_out/r2a_bench_volumetric_512.json
tail -5 gpurun_out/r2a_default.err
