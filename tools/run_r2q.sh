# r2q: OptB200_FilterGaussianNoAlloc -- its test alone in a fresh process, then the whole GPU suite with it collected last
set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
python -m pytest tests/test_zz_no_alloc.py -m gpu -q 2>&1 | tail -4 > gpurun_out/r2q_pytest_alone.txt
cat gpurun_out/r2q_pytest_alone.txt
python -m pytest tests -m gpu -x -q --ignore=tests/test_fullsize_gpu.py 2>&1 | tail -8 > gpurun_out/r2q_pytest_all.txt
cat gpurun_out/r2q_pytest_all.txt
