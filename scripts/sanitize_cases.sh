#!/bin/bash
# compute-sanitizer over four small launches (plain and pipelined schedule, cfg3 and cfg5); one tool per call.
# usage: scripts/sanitize_cases.sh TOOL OUTFILE
tool=$1; out=$2; : > "$out"
for c in "3 300 600 -1" "3 300 600 1" "5 150 600 -1" "5 150 600 1"; do
  echo "=== $tool: ncu_case.py $c" >> "$out"
  timeout -s KILL 900 compute-sanitizer --tool "$tool" --print-limit 20 python scripts/ncu_case.py $c >> "$out" 2>&1
  echo "exit $?" >> "$out"
done
grep -E "^=== |ERROR SUMMARY|RACECHECK SUMMARY|exit |hazard" "$out" | head -40
