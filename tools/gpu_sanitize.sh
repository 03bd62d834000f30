#!/bin/bash
# compute-sanitizer over tools/sanitize.py: memcheck, racecheck, initcheck (summaries under gpurun_out/<tag>/)
# NOTE: this GPU pool refuses compute-sanitizer (it exits 86 with a notice); the script is for boxes that allow it.
tag=${1:-sanitize}; out=gpurun_out/$tag; mkdir -p $out
timeout 300 python tools/sanitize.py > $out/plain.log 2>&1; echo "plain exit $?"; tail -2 $out/plain.log
for tool in ${SAN_TOOLS:-memcheck racecheck initcheck}; do
  timeout ${SAN_TIMEOUT:-600} compute-sanitizer --tool $tool --print-limit 30 --error-exitcode 9 python tools/sanitize.py ${SAN_ONLY:-} > $out/$tool.log 2>&1
  echo "$tool exit $?"; grep -c "^ok " $out/$tool.log; grep -E "ERROR SUMMARY|RACECHECK SUMMARY|=========     at |Invalid|Uninitialized|hazard" $out/$tool.log | sort | uniq -c | sort -rn | head -20
done
