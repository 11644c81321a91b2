# r2l: ARAP gather: software-pipelined (5 CTAs/SM) vs simple loop (8 CTAs/SM)
set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
for v in pipe simple; do
  if [ $v = simple ]; then export OPTB200_GATHER_SIMPLE=1; else unset OPTB200_GATHER_SIMPLE; fi
  timeout 600 python bench.py --workload arap_1m --steps 5 --warmup 3 --no-cpu 2> gpurun_out/r2l_arap_$v.err | tail -1 > gpurun_out/r2l_bench_arap_1m_$v.json
  python -c "import json;d=json.load(open('gpurun_out/r2l_bench_arap_1m_$v.json'));print('$v',d['value'],d['final_cost'],d['roofline']['kernel_ms']['PCGStep1'])"
done
