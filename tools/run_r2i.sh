# r2i: compute-sanitizer over the small single-GPU cases (one tool per call, B200_PROFILING.md): TOOL=memcheck|racecheck
set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
TOOL=${TOOL:-memcheck}
python tools/sanitize_cases.py > gpurun_out/r2i_plain_$TOOL.log 2>&1 &&
timeout 1500 compute-sanitizer --tool $TOOL --error-exitcode 9 python tools/sanitize_cases.py > gpurun_out/r2i_sanitizer_$TOOL.log 2>&1
echo "exit code $?" >> gpurun_out/r2i_sanitizer_$TOOL.log
grep -E "sanitize |ERROR SUMMARY|RACECHECK SUMMARY|exit code|=========.*(Error|Race|Hazard)" gpurun_out/r2i_sanitizer_$TOOL.log | head -60
