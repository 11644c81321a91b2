# r2e (2 GPUs): slab paths against the oracle (peer memory and NCCL only), fused volumetric slab kernel, strong scaling of 256^3 at N=2
set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
nvidia-smi -L
timeout 1200 python -m pytest tests/test_multigpu.py -m gpu -x -q -s 2>&1 | grep -E "mgpu|passed|failed|skipped|Error|error" | cut -c1-400 > gpurun_out/r2e_pytest_mgpu.txt
cat gpurun_out/r2e_pytest_mgpu.txt
for N in 1 2; do
  if [ $N = 1 ]; then L="python"; else L="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29601"; fi
  timeout 600 $L bench.py --gpus $N --workload volumetric_256 --steps 3 --warmup 3 --no-cpu 2> gpurun_out/r2e_vol256_n$N.err | tail -1 > gpurun_out/r2e_scale_volumetric256_n$N.json
  python -c "import json;d=json.load(open('gpurun_out/r2e_scale_volumetric256_n$N.json'));print($N,d['value'],d['final_cost'],d['iteration_ms'],d['roofline']['kernel_ms'] if d['roofline'] else None)"
done
