# r2z: shape_from_shading PCGComputeCtC through diagJTJ (diagonal only) vs the generic evalJTF kernel
set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_parity_gpu.py -m gpu -q -x -k "sfs" 2>&1 | tail -3 > gpurun_out/r2z_pytest.txt
cat gpurun_out/r2z_pytest.txt
for v in diag generic; do
  if [ $v = generic ]; then export OPTB200_GENERIC_CTC=1; else unset OPTB200_GENERIC_CTC; fi
  for dbl in "" "--double"; do
  timeout 600 python bench.py --workload sfs_2048 --kind LMGPU $dbl --steps 5 --warmup 3 --no-cpu 2> gpurun_out/r2z_$v.err | tail -1 > gpurun_out/r2z_bench_sfs_2048_kindLMGPU${dbl#--}_$v.json
  python -c "import json;d=json.load(open('gpurun_out/r2z_bench_sfs_2048_kindLMGPU${dbl#--}_$v.json'));print('$v $dbl', round(d['value'],1), d['final_cost'], d['pcg_iterations_per_step'], d.get('parity'))"
  done
done
unset OPTB200_GENERIC_CTC
for v in diag generic; do
python -c "import json;d=json.load(open('gpurun_out/r2z_bench_sfs_2048_kindLMGPU_$v.json'));print('$v', {k:round(x,4) for k,x in d['roofline']['kernel_ms'].items() if x})"
done
