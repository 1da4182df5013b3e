#!/bin/bash
# Runs ON THE GPU BOX: `ncu --set full` with the source page of the B = 1 rollout and sweep kernels (latency profile:
# one warp per SM, so the stall samples per instruction ARE the critical path).   tools/ncu_lat.sh <tag>
TAG=${1:-lat}
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"k_forward_coop|k_backward" --launch-skip 9 -c 6 \
  -f -o gpurun_out/lat_$TAG python tools/gpu_check.py lat > gpurun_out/ncu_lat_$TAG.log 2>&1
echo "ncu rc=$?"
ncu -i gpurun_out/lat_$TAG.ncu-rep --page raw --csv > gpurun_out/lat_${TAG}_raw.csv 2>/dev/null
ncu -i gpurun_out/lat_$TAG.ncu-rep --page source --csv > gpurun_out/lat_${TAG}_source.csv 2>/dev/null
gzip -f gpurun_out/lat_${TAG}_source.csv
SZ=$(stat -c %s gpurun_out/lat_$TAG.ncu-rep); if [ "$SZ" -gt 30000000 ]; then rm -f gpurun_out/lat_$TAG.ncu-rep; fi
ls -la gpurun_out | grep lat_
