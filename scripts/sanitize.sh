#!/bin/bash
# compute-sanitizer over the small cases of scripts/sanitize_cases.py.  ONE tool per gpurun call (B200_PROFILING.md):
#   gpurun --timeout 900 -- 'scripts/sanitize.sh racecheck'      (then memcheck, synccheck, initcheck in calls of their own)
# The plain run comes first: a program that faults must not be put under the tool.  Log -> gpurun_out/sanitize_<tool>.log;
# the summary lines are copied to profiles/ by hand.
set -u
TOOL=${1:-racecheck}
CASES=${2:-ABC}
mkdir -p gpurun_out
python scripts/sanitize_cases.py "$CASES" > gpurun_out/sanitize_plain.log 2>&1 || { echo "plain run failed"; tail -20 gpurun_out/sanitize_plain.log; exit 1; }
cat gpurun_out/sanitize_plain.log
EXTRA=""
[ "$TOOL" = racecheck ] && EXTRA="--racecheck-report all"
timeout 800 compute-sanitizer --tool "$TOOL" $EXTRA --print-limit 50 python scripts/sanitize_cases.py "$CASES" > gpurun_out/sanitize_${TOOL}.log 2>&1
echo "compute-sanitizer --tool $TOOL rc=$?"
grep -E "ERROR SUMMARY|RACECHECK SUMMARY|hazard|Error|^[ABC]:" gpurun_out/sanitize_${TOOL}.log | head -40
