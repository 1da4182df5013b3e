// dmma_probe.cu -- micro-benchmark: fp64 tensor-core MMA (mma.sync m8n8k4 / m16n8k8 .f64) against DFMA on sm_100a.
// Question it answers for DESIGN.md: can the 8x8-tiled products of the Riccati sweep run as DMMA at a useful rate,
// and do DMMA and DFMA share one pipe?  Prints instruction rates per SM sub-partition and the TFLOP/s they imply.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/dmma_probe tools/dmma_probe.cu
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ void dmma884(double &d0, double &d1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(d0), "+d"(d1) : "d"(a), "d"(b));
}
__device__ __forceinline__ void dmma1688(double *d, const double *a, const double *b) {
  asm volatile("mma.sync.aligned.m16n8k8.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+d"(d[0]), "+d"(d[1]), "+d"(d[2]), "+d"(d[3])
               : "d"(a[0]), "d"(a[1]), "d"(a[2]), "d"(a[3]), "d"(b[0]), "d"(b[1]));
}

// MODE 0: DFMA only, 1: m8n8k4 only, 2: both interleaved (one DFMA group per DMMA group), 3: m16n8k8
template <int MODE, int ILP>
__global__ void __launch_bounds__(128) k_probe(double *out, int iters, double a, double b) {
  double c[ILP][4], x[ILP];
#pragma unroll
  for (int i = 0; i < ILP; ++i) {
    x[i] = threadIdx.x * 1e-3 + i;
#pragma unroll
    for (int k = 0; k < 4; ++k) c[i][k] = 0.0;
  }
  const double av[4] = {a, a * 0.5, a * 0.25, a * 0.125}, bv[2] = {b, b * 0.5};
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int r = 0; r < 8; ++r) {
#pragma unroll
      for (int i = 0; i < ILP; ++i) {
        if (MODE == 1 || MODE == 2) dmma884(c[i][0], c[i][1], a, b);
        if (MODE == 3) dmma1688(c[i], av, bv);
        if (MODE == 0 || MODE == 2) x[i] = fma(x[i], a, b);
      }
    }
  }
  double s = 0.0;
#pragma unroll
  for (int i = 0; i < ILP; ++i) s += x[i] + c[i][0] + c[i][1] + c[i][2] + c[i][3];
  if (s == 123.456) out[0] = s;
}

template <int MODE, int ILP> void run(const char *name, int sms, int warps_per_sm, double clock_ghz) {
  double *d;
  cudaMalloc(&d, 64);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  const int iters = 2048, threads = 128, blocks = sms * warps_per_sm / 4;
  float best = 1e30f;
  for (int rep = 0; rep < 4; ++rep) {
    cudaEventRecord(e0);
    k_probe<MODE, ILP><<<blocks, threads>>>(d, iters, 0.999, 1e-9);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    if (rep > 0 && ms < best) best = ms;
  }
  const double groups = (double)iters * 8 * ILP;  // per warp
  const double warps_per_smsp = warps_per_sm / 4.0;
  const double cyc = best * 1e-3 * clock_ghz * 1e9;
  const double cyc_per_group = cyc / (groups * warps_per_smsp);  // SMSP cycles per (instruction group) at this occupancy
  double flop_per_group = 0.0;
  if (MODE == 0) flop_per_group = 64;
  if (MODE == 1) flop_per_group = 512;
  if (MODE == 2) flop_per_group = 576;
  if (MODE == 3) flop_per_group = 4096;
  const double tf = flop_per_group * groups * (double)blocks * 4 / (best * 1e-3) * 1e-12;
  printf("%-28s ILP=%d warps/SM=%2d  %8.3f ms  %7.2f SMSP-cycles per group  %7.2f TFLOP/s\n", name, ILP, warps_per_sm, best, cyc_per_group, tf);
  cudaFree(d);
}

int main() {
  cudaDeviceProp p;
  cudaGetDeviceProperties(&p, 0);
  int khz = 0;
  cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
  const double ghz = khz * 1e-6;
  printf("%s, %d SMs, %.3f GHz (attribute)\n", p.name, p.multiProcessorCount, ghz);
  const int sms = p.multiProcessorCount;
  for (int w : {4, 12, 16, 32}) {
    run<0, 1>("DFMA", sms, w, ghz);
    run<0, 4>("DFMA", sms, w, ghz);
    run<1, 1>("DMMA m8n8k4", sms, w, ghz);
    run<1, 4>("DMMA m8n8k4", sms, w, ghz);
    run<1, 8>("DMMA m8n8k4", sms, w, ghz);
    run<3, 1>("DMMA m16n8k8", sms, w, ghz);
    run<3, 4>("DMMA m16n8k8", sms, w, ghz);
    run<2, 4>("DMMA m8n8k4 + DFMA", sms, w, ghz);
  }
  return 0;
}
