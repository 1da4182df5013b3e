#!/bin/bash
# Runs ON THE GPU BOX (via gpurun): one `ncu --set full` capture of the solver kernels of the 2nd FDDP iteration
# (B=4096 FF, N=50) and CSV exports of it; the .ncu-rep is dropped if it would blow gpurun's 64 MiB return limit.
#   tools/ncu_capture.sh <tag> [kernel-regex] [skip] [count]
set -u
TAG=${1:-r1}
RE=${2:-"k_backward|k_forward|k_node_diff"}
SKIP=${3:-5}
CNT=${4:-5}
mkdir -p gpurun_out
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"$RE" --launch-skip $SKIP -c $CNT \
  -f -o gpurun_out/full_$TAG python tools/gpu_check.py prof > gpurun_out/ncu_full_$TAG.log 2>&1
echo "ncu rc=$?"
ncu -i gpurun_out/full_$TAG.ncu-rep --page raw --csv > gpurun_out/full_${TAG}_raw.csv 2>/dev/null
ncu -i gpurun_out/full_$TAG.ncu-rep --page details --csv > gpurun_out/full_${TAG}_details.csv 2>/dev/null
ncu -i gpurun_out/full_$TAG.ncu-rep --page source --csv > gpurun_out/full_${TAG}_source.csv 2>/dev/null
SZ=$(stat -c %s gpurun_out/full_$TAG.ncu-rep)
if [ "$SZ" -gt 30000000 ]; then rm -f gpurun_out/full_$TAG.ncu-rep; echo "rep dropped ($SZ bytes)"; fi
gzip -f gpurun_out/full_${TAG}_source.csv
ls -la gpurun_out
