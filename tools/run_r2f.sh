# r2f: CSR known-answer test + the three solver modes
set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_csr_gpu.py -m gpu -q -x 2>&1 | tail -30 > gpurun_out/r2f_pytest_csr.txt
cat gpurun_out/r2f_pytest_csr.txt
