#!/bin/bash
# Runs ON THE GPU BOX: A/B of engine builds (tools/ab/*.so) on the bench workload; prints one line per build.
for lib in tools/ab/*.so; do
  FFMPC_LIB=$PWD/$lib timeout 120 python bench.py --steps 3 --warmup 2 --no-cpu-baseline --no-latency "$@" 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); k=d['roofline']['kernel_ms_per_step']
print('$lib', round(d['value']), 'solves/s', {a: round(b,1) for a,b in k.items()})"
done
