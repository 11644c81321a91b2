# r2u (2 GPUs): slab parity against the oracle with the deferred delta update; N=1,2 strong scaling of volumetric 256^3
set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_multigpu.py -m gpu -x -q -s 2>&1 | grep -E "mgpu|passed|failed|skipped|Error|error" | cut -c1-150 > gpurun_out/r2u_pytest_mgpu.txt
grep -c " OK  vs_oracle= OK " gpurun_out/r2u_pytest_mgpu.txt; grep -E "MISMATCH|passed|failed" gpurun_out/r2u_pytest_mgpu.txt
for N in 1 2; do
  if [ $N = 1 ]; then L="python"; else L="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29601"; fi
  timeout 600 $L bench.py --gpus $N --workload volumetric_256 --steps 3 --warmup 3 --no-cpu 2> gpurun_out/r2u_vol256_n$N.err | tail -1 > gpurun_out/r2u_scale_volumetric256_n$N.json
  python -c "import json;d=json.load(open('gpurun_out/r2u_scale_volumetric256_n$N.json'));print($N,d['value'],d['final_cost'],d['iteration_ms']['total'],d['iteration_ms']['kernels'])"
done
