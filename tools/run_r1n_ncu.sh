set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
python bench.py --steps 1 --warmup 1 --no-cpu > gpurun_out/r1n_plain.log 2>&1 || exit 1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_poisson_march -s 20 -c 1 -o gpurun_out/r1n_poisson_full -f python bench.py --steps 1 --warmup 1 --no-cpu > gpurun_out/r1n_ncu_full.log 2>&1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -s 300 -c 330 --csv --log-file gpurun_out/r1n_launches_poisson4096.csv python bench.py --steps 1 --warmup 1 --no-cpu > /dev/null 2>&1
python bench.py > gpurun_out/r1n_bench_default.log 2> gpurun_out/r1n_bench_default.err
tail -1 gpurun_out/r1n_bench_default.log > gpurun_out/r1n_bench_poisson_4096.json
python bench.py --impl reference --steps 2 --warmup 1 2>/dev/null | tail -1 > gpurun_out/r1n_bench_reference_arm.json
ls -la gpurun_out | grep r1n
