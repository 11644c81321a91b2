# r2r: graph energies with coalesced partial rows (+ atomics A/B), SFS band-height sweep, parity lines with the 12-iteration check
set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_parity_gpu.py -m gpu -q -x -k "embedded or cotangent or robust" 2>&1 | tail -4 > gpurun_out/r2r_pytest.txt
cat gpurun_out/r2r_pytest.txt
line() { python -c "import json;d=json.load(open('$1'));r=d['roofline'] or {};p=d.get('parity') or {};print('$2',round(d['value'],1),{k:round(v,4) for k,v in (r.get('kernel_ms') or {}).items() if v},'parity',p.get('ok'),p.get('rel_err_cost'),p.get('max_rel_err_unknowns'))"; }
for w in embedded_1m cotangent_1m robust_1m; do
  timeout 600 python bench.py --workload $w --steps 5 --warmup 3 --cpu-budget 10 2> gpurun_out/r2r_$w.err | tail -1 > gpurun_out/r2r_bench_$w.json
  line gpurun_out/r2r_bench_$w.json $w
  OPTB200_GRAPH_ATOMICS=1 timeout 600 python bench.py --workload $w --steps 5 --warmup 3 --no-cpu 2> gpurun_out/r2r_${w}_atomics.err | tail -1 > gpurun_out/r2r_bench_${w}_atomics.json
  line gpurun_out/r2r_bench_${w}_atomics.json ${w}_atomics
done
for R in 0 12 24 32 48; do
  if [ $R = 0 ]; then unset OPTB200_SFS_BAND; else export OPTB200_SFS_BAND=$R; fi
  timeout 600 python bench.py --workload sfs_2048 --steps 5 --warmup 3 --no-cpu 2> gpurun_out/r2r_sfs_band$R.err | tail -1 > gpurun_out/r2r_bench_sfs_2048_band$R.json
  line gpurun_out/r2r_bench_sfs_2048_band$R.json sfs_band$R
done
unset OPTB200_SFS_BAND
for v in "" "--double" "--kind LMGPU"; do
  n=sfs_2048$(echo $v | tr -d ' -')
  timeout 900 python bench.py --workload sfs_2048 $v --steps 5 --warmup 3 --cpu-budget 10 2> gpurun_out/r2r_$n.err | tail -1 > gpurun_out/r2r_bench_$n.json
  line gpurun_out/r2r_bench_$n.json $n
done
