# r3i: the driver's own round-end command, verbatim, in one process
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 170 python -m pytest tests/ -x -q -m gpu 2>&1 | tail -8 > gpurun_out/r3i_pytest_driver_command.txt
cat gpurun_out/r3i_pytest_driver_command.txt
