// ffmpc_launch.h -- host-callable launchers; each family of kernels lives in its own translation
// unit so that it can be compiled with its own ptxas options (see __graft_entry__.build()).
#pragma once
#ifdef FFMPC_EMU
#include "cuda_emu.h"
#else
#include <cuda_runtime.h>
#endif
#include "ffmpc_kernels.cuh"

#ifdef FFMPC_EMU
#define FFMPC_LAUNCH(kernel, grid, block, smem, stream, ...) \
  emu::launch(dim3(grid), dim3(block), (smem), [&]() { kernel(__VA_ARGS__); })
#else
#define FFMPC_LAUNCH(kernel, grid, block, smem, stream, ...) kernel<<<(grid), (block), (smem), (stream)>>>(__VA_ARGS__)
#endif

namespace ffmpc {
#ifndef FFMPC_BWD_WARPS
#define FFMPC_BWD_WARPS 1   // warps (= instances) per block of the Riccati kernel
#endif
constexpr int kBwdWarps = FFMPC_BWD_WARPS;

// ffmpc_engine.cu: sets the text ffmpc_last_error() returns (this thread), returns `code`
int set_error(int code, const char *msg);
// ffmpc_launch_bwd.cu
int prepare_backward();
void launch_backward(const Params &P, cudaStream_t st);
// ffmpc_launch_coop_node.cu / ffmpc_launch_coop_fwd.cu  (lane-cooperative knot evaluation, ffmpc_coop.cuh)
int prepare_coop_node();
int prepare_coop_fwd();
void launch_node_diff_coop(const Params &P, cudaStream_t st);
void launch_node_eval_coop(const Params &P, const NodeEvalArgs &A, double *rec, cudaStream_t st);
int launch_forward_coop(const Params &P, int it, cudaStream_t st);  // returns the number of kernel launches
}  // namespace ffmpc
