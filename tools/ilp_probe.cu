// ilp_probe.cu — how much of the geodesic step's time is dependent-issue latency?  Advances R independent rays per
// thread (same arithmetic as the trace kernel's step: next_micro_ray + horizon_hit) at several occupancies.
// build: nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -fmad=false -I ../relativistic-ray-tracer_b200/csrc ilp_probe.cu -o ilp_probe
#include <stdio.h>
#include <vector>
#include "rrt_math.cuh"
#include "rrt_trace.cuh"

template <int R, int MINB>
__global__ void __launch_bounds__(128, MINB) probe(DevHole bh, int n, int nsteps, const double* __restrict__ od, double* __restrict__ out) {
  int t = blockIdx.x * blockDim.x + threadIdx.x;
  d3 o[R], d[R]; double len[R]; int cap[R];
#pragma unroll
  for (int r = 0; r < R; ++r) {
    int i = (t * R + r) % n;
    o[r] = D3(od[6 * i], od[6 * i + 1], od[6 * i + 2]); d[r] = D3(od[6 * i + 3], od[6 * i + 4], od[6 * i + 5]); len[r] = 0.0; cap[r] = 0;
  }
  for (int j = 0; j < nsteps; ++j) {
#pragma unroll
    for (int r = 0; r < R; ++r) { next_micro_ray(bh, o[r], d[r], len[r]); }
#pragma unroll
    for (int r = 0; r < R; ++r) cap[r] += horizon_hit(bh, o[r], d[r], len[r]) ? 1 : 0;
  }
  double acc = 0;
#pragma unroll
  for (int r = 0; r < R; ++r) acc += o[r].x + d[r].y + len[r] + cap[r];
  out[t] = acc;
}

template <int R, int MINB>
void run(DevHole bh, int n, const double* d_od, double* d_out, int sms) {
  int blocks = sms * MINB, nsteps = 63 * 8;
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  probe<R, MINB><<<blocks, 128>>>(bh, n, nsteps, d_od, d_out);
  cudaEventRecord(e0);
  for (int k = 0; k < 5; ++k) probe<R, MINB><<<blocks, 128>>>(bh, n, nsteps, d_od, d_out);
  cudaEventRecord(e1); cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  double chords = 5.0 * blocks * 128.0 * R * nsteps;
  cudaFuncAttributes a; cudaFuncGetAttributes(&a, probe<R, MINB>);
  printf("rays/thread %d  blocks/SM %d  regs %3d  %7.2f G chords/s  (%s)\n", R, MINB, a.numRegs, chords / ms / 1e6, cudaGetErrorString(cudaGetLastError()));
}

int main() {
  cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
  int n = 1 << 16;
  std::vector<double> od(6 * n);
  unsigned s = 12345;
  auto rnd = [&]() { s = s * 1664525u + 1013904223u; return (s >> 8) / 16777216.0; };
  for (int i = 0; i < n; ++i) {
    double x = 2 * rnd() - 1, y = 2 * rnd(), z = 4.8, dx = 0.4 * (rnd() - 0.5), dy = 0.4 * (rnd() - 0.5), dz = -1;
    double l = sqrt(dx * dx + dy * dy + dz * dz);
    od[6 * i] = x; od[6 * i + 1] = y; od[6 * i + 2] = z; od[6 * i + 3] = dx / l; od[6 * i + 4] = dy / l; od[6 * i + 5] = dz / l;
  }
  double *d_od, *d_out;
  cudaMalloc(&d_od, od.size() * 8); cudaMalloc(&d_out, (size_t)p.multiProcessorCount * 16 * 128 * 8);
  cudaMemcpy(d_od, od.data(), od.size() * 8, cudaMemcpyHostToDevice);
  DevHole bh; bh.cx = 0; bh.cy = 1; bh.cz = 0; bh.r = 0.1; bh.r2 = 0.01; bh.h = 0.1; bh.cos_h = cos(0.1); bh.sin_h = sin(0.1); bh.mode = 1; bh.nchords = 63;
  run<1, 6>(bh, n, d_od, d_out, p.multiProcessorCount);
  run<1, 8>(bh, n, d_od, d_out, p.multiProcessorCount);
  run<1, 12>(bh, n, d_od, d_out, p.multiProcessorCount);
  run<1, 16>(bh, n, d_od, d_out, p.multiProcessorCount);
  run<2, 3>(bh, n, d_od, d_out, p.multiProcessorCount);
  run<2, 4>(bh, n, d_od, d_out, p.multiProcessorCount);
  run<2, 6>(bh, n, d_od, d_out, p.multiProcessorCount);
  run<2, 8>(bh, n, d_od, d_out, p.multiProcessorCount);
  run<4, 3>(bh, n, d_od, d_out, p.multiProcessorCount);
  run<4, 4>(bh, n, d_od, d_out, p.multiProcessorCount);
  return 0;
}
