# r2v: scalar read-back through mapped pinned memory instead of the D2H copy engine: e2e A/B; ncu of k_pcg_step2 (three DM forms)
set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_parity_gpu.py tests/test_capi.py tests/test_c_driver.py -m gpu -q -x 2>&1 | tail -3 > gpurun_out/r2v_pytest.txt
cat gpurun_out/r2v_pytest.txt
for w in poisson_4096 volumetric_512; do
  for v in mapped memcpy; do
    if [ $v = memcpy ]; then export OPTB200_SCALAR_MEMCPY=1; else unset OPTB200_SCALAR_MEMCPY; fi
    timeout 900 python bench.py --workload $w --steps 5 --warmup 3 --no-cpu 2> gpurun_out/r2v_${w}_$v.err | tail -1 > gpurun_out/r2v_bench_${w}_$v.json
    python -c "import json;d=json.load(open('gpurun_out/r2v_bench_${w}_$v.json'));e=d['e2e'];print('$w $v', round(d['value'],1), 'e2e', round(e['value'],1), 'ms/step', round(d['ms_per_step'],1), round(e['ms_per_step'],1), 'cold', round(e['cold_solve_ms'],1), d['clocks']['reasons'])"
  done
done
unset OPTB200_SCALAR_MEMCPY
CMD="python bench.py --workload volumetric_256 --steps 1 --warmup 1 --no-cpu"
$CMD > gpurun_out/r2v_plain.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_pcg_step2 -s 10 -c 3 -f -o gpurun_out/r2v_step2_full $CMD > gpurun_out/r2v_ncu.log 2>&1
tail -2 gpurun_out/r2v_ncu.log
