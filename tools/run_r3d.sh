# r3d (2 GPUs): default bench at N=2 with the slab-mode operator parity in the line
set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
L="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29601"
timeout 900 $L bench.py --gpus 2 --steps 3 --warmup 3 2> gpurun_out/r3d_vol512_n2.err | tail -1 > gpurun_out/r3d_scale_volumetric512_n2.json
python -c "import json;d=json.load(open('gpurun_out/r3d_scale_volumetric512_n2.json'));print(2,d['value'],d['e2e']['value'],d['final_cost'],d.get('parity'))"
tail -5 gpurun_out/r3d_vol512_n2.err
