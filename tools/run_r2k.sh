# r2k: ARAP gather on AoS vectors: unfused vs fused with PCGStep3; ncu of the gather
set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
for v in plain fused; do
  if [ $v = fused ]; then export OPTB200_GATHER_FUSED=1; else unset OPTB200_GATHER_FUSED; fi
  timeout 600 python bench.py --workload arap_1m --steps 5 --warmup 3 --no-cpu 2> gpurun_out/r2k_arap_$v.err | tail -1 > gpurun_out/r2k_bench_arap_1m_$v.json
  python -c "import json;d=json.load(open('gpurun_out/r2k_bench_arap_1m_$v.json'));print('$v',d['value'],d['final_cost'],d['roofline']['kernel_ms'])"
done
unset OPTB200_GATHER_FUSED
CMD="python bench.py --workload arap_1m --steps 1 --warmup 1 --no-cpu"
$CMD > gpurun_out/r2k_plain.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_gather_jtj -s 20 -c 1 -f -o gpurun_out/r2k_arap $CMD > gpurun_out/r2k_ncu.log 2>&1
tail -2 gpurun_out/r2k_ncu.log
