# r2s: deferred delta update in PCGStep2 (GN + fused kernel): full suite, then A/B on the big grids
set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q --ignore=tests/test_fullsize_gpu.py 2>&1 | tail -6 > gpurun_out/r2s_pytest.txt
cat gpurun_out/r2s_pytest.txt
line() { python -c "import json;d=json.load(open('$1'));r=d['roofline'] or {};print('$2',round(d['value'],1),d['final_cost'],{k:round(v,4) for k,v in (r.get('kernel_ms') or {}).items() if v and k in ('PCGStep3+1','PCGStep2')},d['clocks']['sm_mhz'],d['clocks']['reasons'])"; }
for w in poisson_4096 optical_flow_2048 sfs_2048 volumetric_256 volumetric_512; do
  for v in lazy eager; do
    if [ $v = eager ]; then export OPTB200_NO_LAZY_DELTA=1; else unset OPTB200_NO_LAZY_DELTA; fi
    timeout 900 python bench.py --workload $w --steps 5 --warmup 3 --no-cpu 2> gpurun_out/r2s_${w}_$v.err | tail -1 > gpurun_out/r2s_bench_${w}_$v.json
    line gpurun_out/r2s_bench_${w}_$v.json ${w}_$v
  done
done
unset OPTB200_NO_LAZY_DELTA
timeout 600 python -m pytest tests/test_fullsize_gpu.py -m gpu -q -x 2>&1 | tail -4 > gpurun_out/r2s_pytest_fullsize.txt
cat gpurun_out/r2s_pytest_fullsize.txt
