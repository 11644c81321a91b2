# r3j: the default bench line once more on another box (r3f's box was power-capped to 1672 MHz)
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 110 python bench.py --steps 5 --warmup 3 2> gpurun_out/r3j_default.err | tail -1 > gpurun_out/r3j_bench_volumetric_512.json
python -c "import json;d=json.load(open('gpurun_out/r3j_bench_volumetric_512.json'));print(round(d['value'],1),round(d['e2e']['value'],1),d['final_cost'],d['parity']['ok'],d['roofline']['frac'],d['clocks'])"
