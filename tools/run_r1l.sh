set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
python -m pytest tests/test_drivers_gpu.py -m gpu -q 2>&1 | tail -30 > gpurun_out/r1l_pytest_drivers.txt
python tools/bench_driver_ops.py 4096 > gpurun_out/r1l_driver_ops.jsonl 2> gpurun_out/r1l_driver_ops.err
cat gpurun_out/r1l_pytest_drivers.txt gpurun_out/r1l_driver_ops.jsonl; tail -5 gpurun_out/r1l_driver_ops.err
