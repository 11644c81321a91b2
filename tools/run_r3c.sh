# r3c (2 GPUs): slab parity with the guarded C layer; default bench (volumetric 512^3 strong) at N=2 with ranks bound next to their GPU
set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_multigpu.py -m gpu -x -q -s 2>&1 | grep -E "mgpu|passed|failed|skipped|Error|error" | cut -c1-150 > gpurun_out/r3c_pytest_mgpu.txt
grep -c " OK  vs_oracle= OK " gpurun_out/r3c_pytest_mgpu.txt; grep -E "MISMATCH|passed|failed" gpurun_out/r3c_pytest_mgpu.txt
L="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29601"
timeout 900 $L bench.py --gpus 2 --steps 3 --warmup 3 2> gpurun_out/r3c_vol512_n2.err | tail -1 > gpurun_out/r3c_scale_volumetric512_n2.json
python -c "import json;d=json.load(open('gpurun_out/r3c_scale_volumetric512_n2.json'));print(2,d['value'],d['e2e'],d['final_cost'],d['iteration_ms']['total'],d['iteration_ms']['kernels'],d.get('parity'))"
tail -3 gpurun_out/r3c_vol512_n2.err
