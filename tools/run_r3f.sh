# r3f: final single-GPU validation: whole -m gpu suite, BASELINE-size checks, smoke, image_warping rotation table A/B,
# the default bench line, and the ncu launch list of the default command
set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q --ignore=tests/test_fullsize_gpu.py 2>&1 | tail -15 > gpurun_out/r3f_pytest.txt
cat gpurun_out/r3f_pytest.txt
timeout 400 python -m pytest tests/test_fullsize_gpu.py -m gpu -q 2>&1 | tail -6 > gpurun_out/r3f_pytest_fullsize.txt
cat gpurun_out/r3f_pytest_fullsize.txt
timeout 200 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -3 > gpurun_out/r3f_smoke.txt
cat gpurun_out/r3f_smoke.txt
for v in table sincos; do
  if [ $v = sincos ]; then export OPTB200_IW_SINCOS=1; else unset OPTB200_IW_SINCOS; fi
  timeout 200 python bench.py --workload image_warping_512 --steps 5 --warmup 3 --no-cpu 2> gpurun_out/r3f_iw_$v.err | tail -1 > gpurun_out/r3f_bench_image_warping_512_$v.json
  python -c "import json;d=json.load(open('gpurun_out/r3f_bench_image_warping_512_$v.json'));print('$v',round(d['value'],1),d['final_cost'],{k:round(x,5) for k,x in d['roofline']['kernel_ms'].items() if x})"
done
unset OPTB200_IW_SINCOS
timeout 400 python bench.py --steps 5 --warmup 3 2> gpurun_out/r3f_default.err | tail -1 > gpurun_out/r3f_bench_volumetric_512.json
python -c "import json;d=json.load(open('gpurun_out/r3f_bench_volumetric_512.json'));print(round(d['value'],1),round(d['e2e']['value'],1),d['final_cost'],d['parity']['ok'],d['roofline']['frac'],d['clocks'])"
timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none -s 300 -c 300 --csv --log-file gpurun_out/r3f_launches_volumetric512.csv python bench.py --steps 2 --warmup 1 --no-cpu > gpurun_out/r3f_ncu.log 2>&1
tail -2 gpurun_out/r3f_ncu.log | cut -c1-200
