# r3b: error handler tests + the rest of the suite after them
set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q --ignore=tests/test_fullsize_gpu.py 2>&1 | tail -12 > gpurun_out/r3b_pytest.txt
cat gpurun_out/r3b_pytest.txt
