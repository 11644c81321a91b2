# r3e (8 GPUs): slab parity at world 2/4/8 with the final engine, then the default bench at N = 8 and N = 4
set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_multigpu.py -m gpu -x -q -s 2>&1 | grep -E "mgpu|passed|failed|skipped|Error|error" | cut -c1-150 > gpurun_out/r3e_pytest_mgpu.txt
grep -c " OK  vs_oracle= OK " gpurun_out/r3e_pytest_mgpu.txt; grep -E "MISMATCH|passed|failed" gpurun_out/r3e_pytest_mgpu.txt
for N in 8 4; do
  L="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29601"
  timeout 600 $L bench.py --gpus $N --steps 5 --warmup 3 2> gpurun_out/r3e_vol512_n$N.err | tail -1 > gpurun_out/r3e_scale_volumetric512_n$N.json
  python -c "import json;d=json.load(open('gpurun_out/r3e_scale_volumetric512_n$N.json'));print($N,d['value'],d['e2e']['value'],d['e2e']['host_cpus_bound'],d['final_cost'],d['iteration_ms']['total'],d['iteration_ms']['kernels'],d['parity']['ok'],d['clocks']['sm_mhz'])"
done
