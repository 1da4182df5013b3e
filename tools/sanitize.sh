#!/bin/bash
# Runs ON THE GPU BOX (gpurun): compute-sanitizer memcheck and racecheck over small solves of every kernel instantiation
# (SURVEY.md section 5).  The Riccati kernel aliases shared memory across phases (QS / HI) and lets cp.async.bulk land in a
# record stage while the other stage is read: racecheck is the tool for exactly that.  Logs -> gpurun_out/<tag>_sanitizer_*.log
TAG=${1:-r2}
mkdir -p gpurun_out
for tool in memcheck racecheck; do
  SAN_B=6 SAN_N=8 timeout 1500 compute-sanitizer --tool $tool --error-exitcode 3 python tools/sanitize_case.py > gpurun_out/${TAG}_sanitizer_$tool.log 2>&1
  echo "$tool rc=$?"
  tail -4 gpurun_out/${TAG}_sanitizer_$tool.log
done
