# r3g: the other BASELINE configurations on the final tree (no CPU leg), one line each
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
for w in poisson_4096 optical_flow_2048 sfs_2048 arap_1m; do
  timeout 100 python bench.py --workload $w --steps 5 --warmup 3 --no-cpu 2> gpurun_out/r3g_$w.err | tail -1 > gpurun_out/r3g_bench_$w.json
  python -c "import json;d=json.load(open('gpurun_out/r3g_bench_$w.json'));print('$w',round(d['value'],1),round(d['e2e']['value'],1),d['final_cost'],d['clocks']['sm_mhz'],d['clocks']['reasons'])"
done
for w in optical_flow_2048 sfs_2048; do
  timeout 100 python bench.py --workload $w --kind LMGPU --steps 5 --warmup 3 --no-cpu 2> gpurun_out/r3g_${w}_lm.err | tail -1 > gpurun_out/r3g_bench_${w}_kindLMGPU.json
  python -c "import json;d=json.load(open('gpurun_out/r3g_bench_${w}_kindLMGPU.json'));print('$w LM',round(d['value'],1),round(d['e2e']['value'],1),d['final_cost'],d['clocks']['sm_mhz'])"
done
