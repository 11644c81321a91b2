set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_multigpu.py -m gpu -x -q -s 2>&1 | grep -v Warning | tail -16 | cut -c1-150 > gpurun_out/r1u_pytest.txt
cat gpurun_out/r1u_pytest.txt
for v in p2p nccl; do
  if [ $v = nccl ]; then export OPTB200_NO_P2P=1; fi
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --workload volumetric_256 --steps 3 --warmup 3 --no-cpu 2>gpurun_out/r1u_n2_$v.err | tail -1 > gpurun_out/r1u_scale_volumetric256_n2_$v.json
  python - <<PY
import json
d=json.load(open('gpurun_out/r1u_scale_volumetric256_n2_$v.json'))
print("$v", d['value'], d['ms_per_step'], d['config']['final_cost'], d['roofline']['kernel_ms'])
PY
done
