# r3a: error handler (OptB200_SetErrorHandler), Engine::init split, well-conditioned intrinsic / cotangent parity cases: whole suite
set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q --ignore=tests/test_fullsize_gpu.py 2>&1 | tail -8 > gpurun_out/r3a_pytest.txt
cat gpurun_out/r3a_pytest.txt
timeout 600 python -m pytest tests/test_fullsize_gpu.py -m gpu -q -x 2>&1 | tail -4 > gpurun_out/r3a_pytest_fullsize.txt
cat gpurun_out/r3a_pytest_fullsize.txt
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -3 > gpurun_out/r3a_smoke.txt
cat gpurun_out/r3a_smoke.txt
