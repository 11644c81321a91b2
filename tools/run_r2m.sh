# r2m: ncu --set full of the fused volumetric kernel at 512^3 (the default bench) and the launch list of one solve
set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 1 --no-cpu"
$CMD > gpurun_out/r2m_plain.log 2>&1 &&
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:k_volumetric_apply_packed -s 70 -c 1 -f -o gpurun_out/r2m_vol512_full $CMD > gpurun_out/r2m_ncu.log 2>&1
tail -2 gpurun_out/r2m_ncu.log
$CMD > gpurun_out/r2m_plain2.log 2>&1 &&
timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none -s 300 -c 300 --csv --log-file gpurun_out/r2m_launches_volumetric512.csv $CMD > /dev/null 2>&1
tail -3 gpurun_out/r2m_launches_volumetric512.csv | cut -c1-300
