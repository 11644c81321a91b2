set -x
cd $GRAFT_REPO_ROOT
for w in optical_flow_2048 sfs_2048 arap_1m volumetric_256 image_warping_512 volumetric_512; do
  python bench.py --workload $w --no-cpu 2> gpurun_out/r1j_bench_$w.err | tail -1 > gpurun_out/r1j_bench_$w.json
done
python bench.py --workload sfs_2048 --no-cpu --kind LMGPU 2>/dev/null | tail -1 > gpurun_out/r1j_bench_sfs_2048_kindLMGPU.json
python bench.py --workload sfs_2048 --no-cpu --double 2>/dev/null | tail -1 > gpurun_out/r1j_bench_sfs_2048_double.json
python bench.py 2> gpurun_out/r1j_bench_poisson_4096.err | tail -1 > gpurun_out/r1j_bench_poisson_4096.json
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_sfs_march -s 25 -c 1 -o gpurun_out/r1j_sfs_full -f python bench.py --workload sfs_2048 --steps 1 --warmup 1 --no-cpu > /dev/null 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_gather_jtj -s 25 -c 1 -o gpurun_out/r1j_arap_full -f python bench.py --workload arap_1m --steps 1 --warmup 1 --no-cpu > /dev/null 2>&1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file gpurun_out/r1j_launches_sfs2048.csv python bench.py --workload sfs_2048 --steps 1 --warmup 1 --no-cpu > /dev/null 2>&1
ls -la gpurun_out | grep r1j
