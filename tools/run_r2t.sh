# r2t: segmented graph scatter with the partial rows stored at their vertex-sorted positions
set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_parity_gpu.py -m gpu -q -x -k "embedded or cotangent or robust" 2>&1 | tail -4 > gpurun_out/r2t_pytest.txt
cat gpurun_out/r2t_pytest.txt
for w in embedded_1m cotangent_1m robust_1m; do
  timeout 600 python bench.py --workload $w --steps 5 --warmup 3 --cpu-budget 10 2> gpurun_out/r2t_$w.err | tail -1 > gpurun_out/r2t_bench_$w.json
  python -c "import json;d=json.load(open('gpurun_out/r2t_bench_$w.json'));r=d['roofline'];p=d['parity'];print('$w',round(d['value'],1),{k:round(v,4) for k,v in r['kernel_ms'].items() if v},'parity',p['ok'],p['rel_err_cost'],p['max_rel_err_unknowns'])"
done
