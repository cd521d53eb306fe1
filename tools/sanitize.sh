#!/bin/bash
# compute-sanitizer over the kernels: memcheck, racecheck, synccheck.
#   tools/sanitize.sh <n_devices> <out_dir>      (run on a GPU box: gpurun -- tools/sanitize.sh 1 gpurun_out/sanitize)
# One log per tool; the summary lines of each ("ERROR SUMMARY" / "RACECHECK SUMMARY") are what profiles/ keeps.
set -u
NDEV=${1:-1}
OUT=${2:-gpurun_out/sanitize}
mkdir -p "$OUT"
CS=$(command -v compute-sanitizer || echo /usr/local/cuda/bin/compute-sanitizer)
rc=0
for tool in memcheck racecheck synccheck; do
  log="$OUT/${tool}_g${NDEV}.log"
  extra=""
  [ "$tool" = racecheck ] && extra="--racecheck-report all"
  timeout 700 "$CS" --tool "$tool" $extra --print-limit 40 --launch-timeout 300 \
      python tools/sanitize_cases.py "$NDEV" all > "$log" 2>&1
  code=$?
  echo "== $tool g$NDEV exit $code" | tee -a "$OUT/summary_g${NDEV}.txt"
  grep -E "SUMMARY|sanitize_cases OK|Error|hazard" "$log" | sort | uniq -c | sort -rn | head -20 | tee -a "$OUT/summary_g${NDEV}.txt"
  [ $code -ne 0 ] && rc=$code
done
exit $rc
