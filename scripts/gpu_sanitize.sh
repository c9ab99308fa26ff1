#!/bin/bash
# compute-sanitizer logs of the small all-kernels case (scripts/sanitize_case.py) -> gpurun_out/<tag>_sanitizer_*.log
tag=${1:-r03}
mkdir -p gpurun_out
CS=/usr/local/cuda/bin/compute-sanitizer
python scripts/sanitize_case.py > gpurun_out/${tag}_sanitize_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/${tag}_sanitize_plain.log; exit 1; }
for tool in memcheck racecheck synccheck initcheck; do
  timeout 900 $CS --tool $tool --print-limit 20 python scripts/sanitize_case.py > gpurun_out/${tag}_sanitizer_${tool}.log 2>&1
  echo "$tool exit $?" >> gpurun_out/${tag}_sanitizer_${tool}.log
  tail -4 gpurun_out/${tag}_sanitizer_${tool}.log
done
