#!/bin/bash
# A/B builds of the engine with extra -D flags (compile-time tuning knobs), into tools/ab/ (git-ignored; travels with gpurun).
#   tools/build_variant.sh <name> [-DFLAG=VALUE ...]      ->  tools/ab/libffmpc_<name>.so   (select with FFMPC_LIB=<path>)
set -e
NAME=$1; shift
CSRC=franka_force_feedback_mpc_b200/csrc
OUT=tools/ab/obj_$NAME
mkdir -p $OUT
for u in ffmpc_engine ffmpc_launch_bwd ffmpc_launch_coop_node ffmpc_launch_coop_fwd ffmpc_peak; do
  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC "$@" -Xptxas -v -c $CSRC/$u.cu -o $OUT/$u.o 2> $OUT/$u.ptxas.log &
done
wait
nvcc -gencode arch=compute_100a,code=sm_100a -shared -cudart shared -o tools/ab/libffmpc_$NAME.so $OUT/*.o
grep -A2 "k_backwardILb1\|k_node_diff_coopILi1ELb1\|k_forward_coopILi1ELb1" $OUT/*.ptxas.log | grep -E "registers|spill" | head -12
