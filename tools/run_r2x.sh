# r2x: LM exit test folded into the tail of PCGStep2: parity tests, then A/B on the LM configurations
set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_parity_gpu.py tests/test_csr_gpu.py tests/test_c_driver.py -m gpu -q -x -k "LM or lm" 2>&1 | tail -3 > gpurun_out/r2x_pytest.txt
cat gpurun_out/r2x_pytest.txt
for w in sfs_2048 optical_flow_2048 poisson_4096; do
  for v in fold separate; do
    if [ $v = separate ]; then export OPTB200_LM_SEPARATE_TEST=1; else unset OPTB200_LM_SEPARATE_TEST; fi
    timeout 600 python bench.py --workload $w --kind LMGPU --steps 5 --warmup 3 --no-cpu 2> gpurun_out/r2x_${w}_$v.err | tail -1 > gpurun_out/r2x_bench_${w}_kindLMGPU_$v.json
    python -c "import json;d=json.load(open('gpurun_out/r2x_bench_${w}_kindLMGPU_$v.json'));print('$w $v', round(d['value'],1), d['final_cost'], d['pcg_iterations_per_step'], d['gpu_launches'])"
  done
done
