mkdir -p gpurun_out
timeout 200 python scripts/sanitize_r2.py > gpurun_out/r2s_plain.log 2>&1; echo "plain rc=$?"; tail -2 gpurun_out/r2s_plain.log
timeout 1500 compute-sanitizer --tool memcheck --error-exitcode 7 python scripts/sanitize_r2.py > gpurun_out/r2s_memcheck.log 2>&1; echo "memcheck rc=$?"; grep -E "ERROR SUMMARY|Invalid|out of bounds|sanitize_r2" gpurun_out/r2s_memcheck.log | head -10
timeout 1500 compute-sanitizer --tool racecheck --error-exitcode 7 python scripts/sanitize_r2.py > gpurun_out/r2s_racecheck.log 2>&1; echo "racecheck rc=$?"; grep -E "RACECHECK SUMMARY|hazard|sanitize_r2" gpurun_out/r2s_racecheck.log | head -10
