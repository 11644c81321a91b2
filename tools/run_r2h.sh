# r2h: full GPU suite (segmented graph scatter, driver ops against oracle/_ref, materialized flags through the reference's OptSolver)
set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q --ignore=tests/test_fullsize_gpu.py 2>&1 | tail -30 > gpurun_out/r2h_pytest.txt
cat gpurun_out/r2h_pytest.txt
