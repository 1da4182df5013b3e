// cuda_emu.h -- TEST-ONLY single-process CUDA emulator used to run the real kernel sources of
// franka_force_feedback_mpc_b200/csrc on the CPU (this build container has no GPU).
//
// Every CUDA thread of a block is a ucontext fiber; __syncwarp / __syncthreads / shuffles / votes
// are cooperative barriers between fibers, so warp-cooperative code (and its missing-barrier bugs)
// behaves deterministically.  Blocks run one after another.  The CUDA runtime calls the engine
// uses are mapped to malloc/memcpy.  This header is never part of the product library: the shipped
// libffmpc_b200.so is built by nvcc only and has no host execution path.
#pragma once
#ifndef FFMPC_EMU
#error "cuda_emu.h is only for the FFMPC_EMU test build"
#endif

#include <ucontext.h>

#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <vector>

#define __global__
#define __device__
#define __host__
#define __forceinline__ inline
#define __launch_bounds__(...)
#define __grid_constant__
#define __shared__
#ifndef __restrict__
#define __restrict__
#endif

struct dim3 {
  unsigned x, y, z;
  dim3(unsigned a = 1, unsigned b = 1, unsigned c = 1) : x(a), y(b), z(c) {}
};

namespace emu {

struct State {
  dim3 tid, bid, bdim, gdim;
  int nthreads = 0, cur = 0;
  std::vector<ucontext_t> ctx;
  ucontext_t sched;
  std::vector<char> finished;
  std::vector<char *> stacks;
  std::vector<int> warp_count, warp_gen;
  int blk_count = 0, blk_gen = 0;
  std::vector<uint64_t> xch;  // [warps*32]
  std::vector<uint64_t> xch2; // [warps*32] second payload (all-gather of two values: tensor-core MMA emulation)
  std::vector<int> pred;      // [warps*32]
  char *smem = nullptr;
  size_t smem_cap = 0;
  const std::function<void()> *fn = nullptr;
  long switches = 0;
};
inline State &S() {
  static State s;
  return s;
}
constexpr size_t kStack = 1 << 20;

inline void yield() {
  State &s = S();
  ++s.switches;
  swapcontext(&s.ctx[s.cur], &s.sched);
}
inline int warp_width(int w) {
  State &s = S();
  int n = s.nthreads - w * 32;
  return n > 32 ? 32 : n;
}
inline void warp_sync() {
  State &s = S();
  const int w = s.cur / 32;
  const int gen = s.warp_gen[w];
  if (++s.warp_count[w] == warp_width(w)) {
    s.warp_count[w] = 0;
    ++s.warp_gen[w];
  } else {
    while (s.warp_gen[w] == gen) yield();
  }
}
inline void block_sync() {
  State &s = S();
  const int gen = s.blk_gen;
  if (++s.blk_count == s.nthreads) {
    s.blk_count = 0;
    ++s.blk_gen;
  } else {
    while (s.blk_gen == gen) yield();
  }
}
inline void trampoline() {
  State &s = S();
  (*s.fn)();
  s.finished[s.cur] = 1;
  swapcontext(&s.ctx[s.cur], &s.sched);
}

inline void launch(dim3 grid, dim3 block, size_t smem_bytes, const std::function<void()> &fn) {
  State &s = S();
  const int nt = (int)(block.x * block.y * block.z);
  s.nthreads = nt;
  s.bdim = block;
  s.gdim = grid;
  s.fn = &fn;
  if ((int)s.stacks.size() < nt) {
    const size_t old = s.stacks.size();
    s.stacks.resize(nt);
    for (size_t i = old; i < (size_t)nt; ++i) s.stacks[i] = (char *)malloc(kStack);
  }
  if (s.smem_cap < smem_bytes + 256) {
    free(s.smem);
    s.smem_cap = smem_bytes + 256;
    s.smem = (char *)aligned_alloc(256, (s.smem_cap + 255) / 256 * 256);
  }
  const int nw = (nt + 31) / 32;
  s.ctx.resize(nt);
  s.finished.assign(nt, 0);
  s.warp_count.assign(nw, 0);
  s.warp_gen.assign(nw, 0);
  s.xch.assign((size_t)nw * 32, 0);
  s.xch2.assign((size_t)nw * 32, 0);
  s.pred.assign((size_t)nw * 32, 0);
  for (unsigned b = 0; b < grid.x * grid.y * grid.z; ++b) {
    s.bid = dim3(b % grid.x, (b / grid.x) % grid.y, b / (grid.x * grid.y));
    std::fill(s.finished.begin(), s.finished.end(), 0);
    std::fill(s.warp_count.begin(), s.warp_count.end(), 0);
    s.blk_count = 0;
    memset(s.smem, 0xCD, s.smem_cap);  // poison: uninitialised shared memory reads show up as garbage
    for (int i = 0; i < nt; ++i) {
      getcontext(&s.ctx[i]);
      s.ctx[i].uc_stack.ss_sp = s.stacks[i];
      s.ctx[i].uc_stack.ss_size = kStack;
      s.ctx[i].uc_link = &s.sched;
      makecontext(&s.ctx[i], (void (*)())trampoline, 0);
    }
    int remaining = nt;
    while (remaining > 0) {
      remaining = 0;
      for (int i = 0; i < nt; ++i) {
        if (s.finished[i]) continue;
        s.cur = i;
        s.tid = dim3(i % block.x, (i / block.x) % block.y, i / (block.x * block.y));
        swapcontext(&s.sched, &s.ctx[i]);
        if (!s.finished[i]) ++remaining;
      }
    }
  }
  s.fn = nullptr;
}

template <class T> inline uint64_t to_bits(T v) {
  static_assert(sizeof(T) <= 8, "shuffle payload too large");
  uint64_t b = 0;
  memcpy(&b, &v, sizeof(T));
  return b;
}
template <class T> inline T from_bits(uint64_t b) {
  T v;
  memcpy(&v, &b, sizeof(T));
  return v;
}
template <class T> inline T shfl_idx(T v, int src_lane) {
  State &s = S();
  const int w = s.cur / 32, lane = s.cur % 32;
  s.xch[w * 32 + lane] = to_bits(v);
  warp_sync();
  const int nwd = warp_width(w);
  T r = (src_lane >= 0 && src_lane < nwd) ? from_bits<T>(s.xch[w * 32 + src_lane]) : v;
  warp_sync();
  return r;
}
// every lane receives the (a, b) pair of all 32 lanes of its warp: two barriers instead of one shuffle per element
inline void allgather2(double a, double b, double *all_a, double *all_b) {
  State &s = S();
  const int w = s.cur / 32, lane = s.cur % 32;
  s.xch[w * 32 + lane] = to_bits(a);
  s.xch2[w * 32 + lane] = to_bits(b);
  warp_sync();
  for (int i = 0; i < 32; ++i) {
    all_a[i] = from_bits<double>(s.xch[w * 32 + i]);
    all_b[i] = from_bits<double>(s.xch2[w * 32 + i]);
  }
  warp_sync();
}
inline unsigned ballot(int p) {
  State &s = S();
  const int w = s.cur / 32, lane = s.cur % 32;
  s.pred[w * 32 + lane] = p ? 1 : 0;
  warp_sync();
  unsigned m = 0;
  for (int i = 0; i < warp_width(w); ++i)
    if (s.pred[w * 32 + i]) m |= 1u << i;
  warp_sync();
  return m;
}

}  // namespace emu

#define threadIdx (emu::S().tid)
#define blockIdx (emu::S().bid)
#define blockDim (emu::S().bdim)
#define gridDim (emu::S().gdim)

using std::isfinite;

inline void __syncwarp(unsigned = 0xffffffffu) { emu::warp_sync(); }
inline void __syncthreads() { emu::block_sync(); }
template <class T> inline T __shfl_sync(unsigned, T v, int src, int width = 32) {
  const int lane = emu::S().cur % 32;
  return emu::shfl_idx(v, (lane / width) * width + (src % width));
}
template <class T> inline T __shfl_xor_sync(unsigned, T v, int m, int width = 32) {
  const int lane = emu::S().cur % 32;
  return emu::shfl_idx(v, lane ^ m);
}
template <class T> inline T __shfl_down_sync(unsigned, T v, unsigned d, int width = 32) {
  const int lane = emu::S().cur % 32;
  const int src = lane + (int)d;
  return emu::shfl_idx(v, (src / width == lane / width) ? src : lane);
}
template <class T> inline T __shfl_up_sync(unsigned, T v, unsigned d, int width = 32) {
  const int lane = emu::S().cur % 32;
  const int src = lane - (int)d;
  return emu::shfl_idx(v, (src >= 0 && src / width == lane / width) ? src : lane);
}
inline unsigned __ballot_sync(unsigned, int p) { return emu::ballot(p); }
inline int __any_sync(unsigned, int p) { return emu::ballot(p) != 0; }
inline int __all_sync(unsigned, int p) {
  const int w = emu::S().cur / 32;
  const unsigned full = emu::warp_width(w) == 32 ? 0xffffffffu : ((1u << emu::warp_width(w)) - 1u);
  return emu::ballot(p) == full;
}
inline double rsqrt(double x) { return 1.0 / std::sqrt(x); }
inline int __ffs(int v) { return __builtin_ffs(v); }
inline int __popc(unsigned v) { return __builtin_popcount(v); }
inline int atomicAdd(int *a, int v) {
  int o = *a;
  *a += v;
  return o;
}

// ---------------- CUDA runtime shim ----------------
typedef int cudaError_t;
typedef void *cudaStream_t;
typedef struct EmuEvent { int dummy; } *cudaEvent_t;
enum { cudaSuccess = 0 };
enum cudaMemcpyKind { cudaMemcpyHostToDevice, cudaMemcpyDeviceToHost, cudaMemcpyDeviceToDevice, cudaMemcpyDefault };
inline cudaError_t cudaMalloc(void **p, size_t n) { *p = aligned_alloc(256, (n + 255) / 256 * 256 + 256); return *p ? 0 : 2; }
inline cudaError_t cudaFree(void *p) { free(p); return 0; }
inline cudaError_t cudaMallocHost(void **p, size_t n) { return cudaMalloc(p, n); }
inline cudaError_t cudaFreeHost(void *p) { free(p); return 0; }
inline cudaError_t cudaMemcpy(void *d, const void *s, size_t n, cudaMemcpyKind) { memcpy(d, s, n); return 0; }
inline cudaError_t cudaMemcpyAsync(void *d, const void *s, size_t n, cudaMemcpyKind, cudaStream_t = nullptr) { memcpy(d, s, n); return 0; }
inline cudaError_t cudaMemsetAsync(void *d, int v, size_t n, cudaStream_t = nullptr) { memset(d, v, n); return 0; }
inline cudaError_t cudaMemset(void *d, int v, size_t n) { memset(d, v, n); return 0; }
inline cudaError_t cudaStreamCreate(cudaStream_t *s) { *s = nullptr; return 0; }
constexpr unsigned cudaStreamNonBlocking = 1u, cudaEventDisableTiming = 2u;
inline cudaError_t cudaStreamCreateWithFlags(cudaStream_t *s, unsigned) { *s = (cudaStream_t)(void *)1; return 0; }  // non-null: "created"
inline cudaError_t cudaStreamCreateWithPriority(cudaStream_t *s, unsigned, int) { *s = (cudaStream_t)(void *)1; return 0; }
inline cudaError_t cudaDeviceGetStreamPriorityRange(int *least, int *greatest) { *least = 0; *greatest = -5; return 0; }
inline cudaError_t cudaStreamWaitEvent(cudaStream_t, cudaEvent_t, unsigned = 0) { return 0; }
inline cudaError_t cudaStreamDestroy(cudaStream_t) { return 0; }
inline cudaError_t cudaStreamSynchronize(cudaStream_t) { return 0; }
inline cudaError_t cudaDeviceSynchronize() { return 0; }
inline cudaError_t cudaSetDevice(int) { return 0; }
inline cudaError_t cudaGetDevice(int *d) { *d = 0; return 0; }
inline cudaError_t cudaGetDeviceCount(int *n) { *n = 1; return 0; }
inline cudaError_t cudaGetLastError() { return 0; }
inline cudaError_t cudaPeekAtLastError() { return 0; }
inline const char *cudaGetErrorString(cudaError_t) { return "emu"; }
inline cudaError_t cudaEventCreate(cudaEvent_t *e) { *e = new EmuEvent; return 0; }
inline cudaError_t cudaEventCreateWithFlags(cudaEvent_t *e, unsigned) { *e = new EmuEvent; return 0; }
inline cudaError_t cudaEventDestroy(cudaEvent_t e) { delete e; return 0; }
inline cudaError_t cudaEventRecord(cudaEvent_t, cudaStream_t = nullptr) { return 0; }
inline cudaError_t cudaEventSynchronize(cudaEvent_t) { return 0; }
inline cudaError_t cudaEventElapsedTime(float *ms, cudaEvent_t, cudaEvent_t) { *ms = 0.f; return 0; }
template <class F> inline cudaError_t cudaFuncSetAttribute(F, int, int) { return 0; }
enum { cudaFuncAttributeMaxDynamicSharedMemorySize = 8 };
