# final validation of the round: full GPU suite (2 GPUs visible so the slab tests run), default bench with CPU baseline,
# reference arm, ncu capture + launch list of the default bench
set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -6 > gpurun_out/r1t_pytest.txt
cat gpurun_out/r1t_pytest.txt
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r1t_smoke.txt 2>&1; tail -2 gpurun_out/r1t_smoke.txt
export CUDA_VISIBLE_DEVICES=0
python bench.py > gpurun_out/r1t_bench_default.log 2> gpurun_out/r1t_bench_default.err
tail -1 gpurun_out/r1t_bench_default.log > gpurun_out/r1t_bench_poisson_4096.json
python bench.py --impl reference --steps 2 --warmup 1 2>/dev/null | tail -1 > gpurun_out/r1t_bench_reference_arm.json
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_poisson_march -s 20 -c 1 -o gpurun_out/r1t_poisson_full -f python bench.py --steps 1 --warmup 1 --no-cpu > gpurun_out/r1t_ncu_full.log 2>&1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -s 300 -c 330 --csv --log-file gpurun_out/r1t_launches_poisson4096.csv python bench.py --steps 1 --warmup 1 --no-cpu > /dev/null 2>&1
ls -la gpurun_out | grep r1t
