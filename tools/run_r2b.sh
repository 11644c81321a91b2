# r2b: ncu --set full of the fused volumetric kernel at 256^3 (why is PCGStep3+1 slower than PCGStep1 + PCGStep3?)
set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
CMD="python bench.py --workload volumetric_256 --steps 1 --warmup 1 --no-cpu"
$CMD > gpurun_out/r2b_plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_volumetric_apply_packed -s 5 -c 2 -f -o gpurun_out/r2b_vol256 $CMD > gpurun_out/r2b_ncu.log 2>&1
tail -3 gpurun_out/r2b_ncu.log
ls -la gpurun_out | grep r2b
