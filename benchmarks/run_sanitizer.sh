# one compute-sanitizer tool per gpurun call:  bash benchmarks/run_sanitizer.sh memcheck|racecheck|synccheck
TOOL=${1:-memcheck}
export PYTHONPATH=.
mkdir -p gpurun_out
timeout 300 python benchmarks/sanitize_small.py > gpurun_out/sanitize_plain.log 2>&1 || { tail -5 gpurun_out/sanitize_plain.log; echo "plain run failed: no sanitizer run"; exit 1; }
tail -1 gpurun_out/sanitize_plain.log
timeout 1500 compute-sanitizer --tool $TOOL --print-limit 20 python benchmarks/sanitize_small.py > gpurun_out/sanitizer_$TOOL.log 2>&1
echo "exit $?"
grep -E "ERROR SUMMARY|RACECHECK SUMMARY|sanitize workload ok|Error|error" gpurun_out/sanitizer_$TOOL.log | head -20
