# r2g (8 GPUs): slab parity against the oracle at 4 and 8 ranks (peer memory and NCCL only), then the strong-scaling curve of the
# default workload (volumetric 512^3) at N = 1, 2, 4, 8 on the same box
set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
nvidia-smi -L | wc -l
timeout 1500 python -m pytest tests/test_multigpu.py -m gpu -q -s -k "4 or 8" 2>&1 | grep -E "mgpu|passed|failed|skipped|Error|error" | cut -c1-330 > gpurun_out/r2g_pytest_mgpu.txt
cat gpurun_out/r2g_pytest_mgpu.txt | cut -c1-120
for N in 8 4 2 1; do
  if [ $N = 1 ]; then L="python"; else L="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2960$N"; fi
  timeout 900 $L bench.py --gpus $N --steps 3 --warmup 2 --no-cpu 2> gpurun_out/r2g_vol512_n$N.err | tail -1 > gpurun_out/r2g_scale_volumetric512_n$N.json
  python -c "import json;d=json.load(open('gpurun_out/r2g_scale_volumetric512_n$N.json'));print($N,d['value'],d['e2e']['value'],d['final_cost'],d['iteration_ms']['total'],d['iteration_ms']['kernels'],d['clocks'])"
done
