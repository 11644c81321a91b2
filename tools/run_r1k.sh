# r1k: LM inner loop with the exit test on the device -- GPU tests, LM bench lines with both loops, timing tables
set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/r1k_pytest.txt
for w in sfs_2048 optical_flow_2048; do
  python bench.py --workload $w --no-cpu --kind LMGPU 2>/dev/null | tail -1 > gpurun_out/r1k_bench_${w}_kindLMGPU.json
  OPTB200_LM_HOSTLOOP=1 python bench.py --workload $w --no-cpu --kind LMGPU 2>/dev/null | tail -1 > gpurun_out/r1k_bench_${w}_kindLMGPU_hostloop.json
done
python tools/timing_table.py sfs_2048 --kind LMGPU > gpurun_out/r1k_timing_sfs_lm.txt 2>&1
python tools/timing_table.py optical_flow_2048 --kind LMGPU > gpurun_out/r1k_timing_flow_lm.txt 2>&1
cat gpurun_out/r1k_pytest.txt
