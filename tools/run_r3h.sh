# r3h: remaining bench workloads on the final tree (no CPU leg)
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
for w in volumetric_256 image_warping_512 embedded_1m cotangent_1m robust_1m; do
  timeout 100 python bench.py --workload $w --steps 5 --warmup 3 --no-cpu 2> gpurun_out/r3h_$w.err | tail -1 > gpurun_out/r3h_bench_$w.json
  python -c "import json;d=json.load(open('gpurun_out/r3h_bench_$w.json'));print('$w',round(d['value'],1),round(d['e2e']['value'],1),d['final_cost'],d['clocks']['sm_mhz'],d['clocks']['reasons'])"
done
for w in poisson_4096; do
  timeout 100 python bench.py --workload $w --kind LMGPU --steps 5 --warmup 3 --no-cpu 2> gpurun_out/r3h_${w}_lm.err | tail -1 > gpurun_out/r3h_bench_${w}_kindLMGPU.json
  python -c "import json;d=json.load(open('gpurun_out/r3h_bench_${w}_kindLMGPU.json'));print('$w LM',round(d['value'],1),round(d['e2e']['value'],1),d['final_cost'],d['clocks']['sm_mhz'])"
  timeout 100 python bench.py --workload $w --double --steps 5 --warmup 3 --no-cpu 2> gpurun_out/r3h_${w}_f64.err | tail -1 > gpurun_out/r3h_bench_${w}_double.json
  python -c "import json;d=json.load(open('gpurun_out/r3h_bench_${w}_double.json'));print('$w f64',round(d['value'],1),round(d['e2e']['value'],1),d['final_cost'],d['clocks']['sm_mhz'])"
done
