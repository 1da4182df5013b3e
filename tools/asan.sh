#!/bin/bash
# AddressSanitizer + UndefinedBehaviorSanitizer over the kernel sources (fiber emulator build) and the oracle, on the CPU.
set -e
CSRC=franka_force_feedback_mpc_b200/csrc
FL="-O1 -g -std=c++17 -fPIC -pthread -fsanitize=address,undefined -fno-omit-frame-pointer"
g++ $FL -shared -o /tmp/liboracle_asan.so oracle/ffmpc_oracle.cpp
g++ $FL -DFFMPC_EMU -Itests/host_emul -x c++ $CSRC/ffmpc_engine.cu $CSRC/ffmpc_launch_bwd.cu $CSRC/ffmpc_launch_coop_node.cu $CSRC/ffmpc_launch_coop_fwd.cu -shared -o /tmp/libemu_asan.so
# fibers switch stacks under ASan's feet: stack-use-after-return detection and leak checking (python itself) are off
EMU_ASAN=/tmp/libemu_asan.so ORACLE_ASAN=/tmp/liboracle_asan.so LD_PRELOAD=$(g++ -print-file-name=libasan.so) \
  ASAN_OPTIONS=detect_leaks=0:detect_stack_use_after_return=0:halt_on_error=1 UBSAN_OPTIONS=print_stacktrace=1:halt_on_error=1 \
  python tools/asan_case.py
