set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
python -m pytest tests/test_drivers_gpu.py tests/test_parity_gpu.py -m gpu -q -k "drivers or long_run or optical_flow_driver or gaussian or prewitt or raster or skin or flow_lines" 2>&1 | tail -30 > gpurun_out/r1m_pytest.txt
python tools/bench_driver_ops.py 4096 > gpurun_out/r1m_driver_ops.jsonl 2> gpurun_out/r1m_driver_ops.err
cat gpurun_out/r1m_pytest.txt gpurun_out/r1m_driver_ops.jsonl; tail -5 gpurun_out/r1m_driver_ops.err
