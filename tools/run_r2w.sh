# r2w: mbarrier try_wait suspend hint in the fused volumetric kernel (A/B on one box); deferred-delta bit-identity tests
set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_parity_gpu.py -m gpu -q -x -k "deferred or volumetric" 2>&1 | tail -3 > gpurun_out/r2w_pytest.txt
cat gpurun_out/r2w_pytest.txt
for rep in 1 2; do
for h in 0 2000 20000; do
  OPTB200_MBAR_HINT_NS=$h timeout 600 python bench.py --workload volumetric_256 --steps 5 --warmup 3 --no-cpu 2> gpurun_out/r2w_vol256_h$h.err | tail -1 > gpurun_out/r2w_bench_volumetric_256_hint${h}_$rep.json
  python -c "import json;d=json.load(open('gpurun_out/r2w_bench_volumetric_256_hint${h}_$rep.json'));print('hint $h rep $rep', round(d['value'],1), d['final_cost'], round(d['roofline']['kernel_ms']['PCGStep3+1'],4), d['clocks']['sm_mhz'], d['clocks']['reasons'])"
done
done
for h in 0 2000; do
  OPTB200_MBAR_HINT_NS=$h timeout 900 python bench.py --steps 3 --warmup 2 --no-cpu 2> gpurun_out/r2w_vol512_h$h.err | tail -1 > gpurun_out/r2w_bench_volumetric_512_hint$h.json
  python -c "import json;d=json.load(open('gpurun_out/r2w_bench_volumetric_512_hint$h.json'));print('512 hint $h', round(d['value'],1), d['final_cost'], round(d['roofline']['kernel_ms']['PCGStep3+1'],4), d['clocks']['sm_mhz'], d['clocks']['reasons'])"
done
