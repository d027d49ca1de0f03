// fp64_peak_probe.cu — what the B200's vector FP64 pipe sustains on PLAIN instruction streams, as a hardware denominator
// for the trace kernel's roofline ("nobody has shown what a plain independent DFMA / DADD / DMUL stream
// sustains on this part").  Each thread keeps ILP independent accumulators and runs a long unrolled loop of one
// instruction kind; the result is reported as warp instructions per clock per SM (clock64 inside the kernel, so the
// figure does not depend on what the SM clock was) and as G warp-instructions/s (CUDA events).
// build: nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a fp64_peak_probe.cu -o fp64_peak_probe
#include <stdio.h>
#include <algorithm>
#include <vector>

enum { K_DFMA = 0, K_DADD = 1, K_DMUL = 2, K_MIX = 3 /* DMUL, DADD alternating: what -fmad=false code looks like */, K_FFMA = 4, K_DIV = 5, K_SQRT = 6 };

template <int KIND, int ILP>
__global__ void __launch_bounds__(128) probe(int iters, double seed, double* out, long long* cyc) {
  double a[ILP];
  float f[ILP];
#pragma unroll
  for (int i = 0; i < ILP; ++i) { a[i] = seed + 1e-3 * (threadIdx.x + 7 * i); f[i] = (float)a[i]; }
  const double m = 1.0 + seed * 1e-9, c = seed * 1e-9;
  const float mf = (float)m, cf = (float)c;
  __syncthreads();
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int r = 0; r < 8; ++r) {
#pragma unroll
      for (int i = 0; i < ILP; ++i) {
        if (KIND == K_DFMA) a[i] = __fma_rn(a[i], m, c);
        else if (KIND == K_DADD) a[i] = __dadd_rn(a[i], c);
        else if (KIND == K_DMUL) a[i] = __dmul_rn(a[i], m);
        else if (KIND == K_MIX) a[i] = (r & 1) ? __dadd_rn(a[i], c) : __dmul_rn(a[i], m);
        else if (KIND == K_FFMA) f[i] = __fmaf_rn(f[i], mf, cf);
        else if (KIND == K_DIV) a[i] = __ddiv_rn(m, a[i]);
        else a[i] = __dsqrt_rn(a[i]) + 1.0;
      }
    }
  }
  long long t1 = clock64();
  double s = 0;
#pragma unroll
  for (int i = 0; i < ILP; ++i) s += a[i] + (double)f[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

template <int KIND, int ILP>
void run(const char* name, int sms, int blocks_per_sm, double* d_out, long long* d_cyc) {
  const int iters = KIND >= K_DIV ? 256 : 4096, blocks = sms * blocks_per_sm;
  probe<KIND, ILP><<<blocks, 128>>>(iters, 1.0, d_out, d_cyc);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  cudaEventRecord(e0);
  probe<KIND, ILP><<<blocks, 128>>>(iters, 1.0, d_out, d_cyc);
  cudaEventRecord(e1); cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  std::vector<long long> cyc(blocks);
  cudaMemcpy(cyc.data(), d_cyc, 8 * blocks, cudaMemcpyDeviceToHost);
  std::sort(cyc.begin(), cyc.end());
  const double med = (double)cyc[blocks / 2];
  const double per_warp = (double)iters * 8 * ILP;             /* instructions of the kind, per warp */
  const double per_sm_clk = per_warp * 4 * blocks_per_sm / med; /* 4 warps per block */
  printf("%-5s ILP %d  warps/SM %3d : %6.3f warp-instr/clk/SM  (%5.2f clk per instr per sub-partition)  %8.1f G warp-instr/s  %s\n", name, ILP, 4 * blocks_per_sm,
         per_sm_clk, 4.0 / per_sm_clk, per_warp * 4 * blocks / ms / 1e6, cudaGetErrorString(cudaGetLastError()));
}

int main() {
  cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
  const int sms = p.multiProcessorCount;
  double* d_out; long long* d_cyc;
  cudaMalloc(&d_out, (size_t)sms * 16 * 128 * 8); cudaMalloc(&d_cyc, (size_t)sms * 16 * 8);
  printf("# %s, %d SMs, clock %d MHz (max)\n", p.name, sms, p.clockRate / 1000);
  const int bps[] = {1, 2, 4, 6, 8, 16};
  for (int b : bps) run<K_DFMA, 8>("DFMA", sms, b, d_out, d_cyc);
  for (int b : bps) run<K_DFMA, 2>("DFMA", sms, b, d_out, d_cyc);
  for (int b : bps) run<K_DFMA, 1>("DFMA", sms, b, d_out, d_cyc);
  for (int b : bps) run<K_DADD, 8>("DADD", sms, b, d_out, d_cyc);
  for (int b : bps) run<K_DMUL, 8>("DMUL", sms, b, d_out, d_cyc);
  for (int b : bps) run<K_MIX, 8>("MIX", sms, b, d_out, d_cyc);
  for (int b : bps) run<K_MIX, 1>("MIX", sms, b, d_out, d_cyc);
  for (int b : bps) run<K_FFMA, 8>("FFMA", sms, b, d_out, d_cyc);
  for (int b : bps) run<K_DIV, 4>("DDIV", sms, b, d_out, d_cyc);
  for (int b : bps) run<K_SQRT, 4>("DSQRT", sms, b, d_out, d_cyc);
  return 0;
}
