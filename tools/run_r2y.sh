# r2y: shape_from_shading computeModelCost on the marching pipeline
set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_parity_gpu.py -m gpu -q -x -k "sfs" 2>&1 | tail -3 > gpurun_out/r2y_pytest.txt
cat gpurun_out/r2y_pytest.txt
for v in march generic; do
  if [ $v = generic ]; then export OPTB200_GENERIC_MODEL_COST=1; else unset OPTB200_GENERIC_MODEL_COST; fi
  for dbl in "" "--double"; do
  timeout 600 python bench.py --workload sfs_2048 --kind LMGPU $dbl --steps 5 --warmup 3 --no-cpu 2> gpurun_out/r2y_$v.err | tail -1 > gpurun_out/r2y_bench_sfs_2048_kindLMGPU${dbl#--}_$v.json
  python -c "import json;d=json.load(open('gpurun_out/r2y_bench_sfs_2048_kindLMGPU${dbl#--}_$v.json'));print('$v $dbl', round(d['value'],1), d['final_cost'], d['pcg_iterations_per_step'])"
  done
done
