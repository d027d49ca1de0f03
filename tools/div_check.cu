// div_check.cu — brute-force check of the stepper's FMA quotients (rrt_trace.cuh: div_known_rcp, div_by_6) against the device's IEEE
// division, bit for bit.  Operand families: (0) random significands and exponents in the safe range, (1) quotients constructed to sit next to
// a rounding boundary (a = RN(b * (q + ulp(q) / 2))), (2) divisors with significands near 1, 4/3, 2 - 2^-52 (the hard cases of the reciprocal
// theorems), (3) operands at the edges of the safe exponent range and outside it (fallback path), zeros, infinities, NaNs.
// build: nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -fmad=false -I ../relativistic-ray-tracer_b200/csrc div_check.cu -o div_check
// usage: div_check [log2 pairs per family, default 34]
#include <stdio.h>
#include <stdlib.h>
#include "rrt_math.cuh"
#include "rrt_trace.cuh"

__device__ __forceinline__ unsigned long long mix(unsigned long long x) {   /* splitmix64 */
  x += 0x9E3779B97F4A7C15ull; x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull; x = (x ^ (x >> 27)) * 0x94D049BB133111EBull; return x ^ (x >> 31);
}
__device__ __forceinline__ double mk(unsigned long long mant, int exp, bool neg) {   /* (1 + mant / 2^52) * 2^exp */
  unsigned long long bits = ((unsigned long long)(exp + 1023) << 52) | (mant & 0xFFFFFFFFFFFFFull) | (neg ? 0x8000000000000000ull : 0ull);
  return __longlong_as_double((long long)bits);
}
__device__ __forceinline__ bool same_bits(double x, double y) {
  return __double_as_longlong(x) == __double_as_longlong(y) || (x != x && y != y);
}

__global__ void check(int family, unsigned long long per_thread, unsigned long long seed, unsigned long long* bad, double* example) {
  unsigned long long s = mix(seed ^ ((unsigned long long)blockIdx.x * blockDim.x + threadIdx.x) * 0x632BE59BD9B4E019ull);
  unsigned long long nbad = 0;
  for (unsigned long long it = 0; it < per_thread; ++it) {
    const unsigned long long r0 = s = mix(s), r1 = s = mix(s), r2 = s = mix(s);
    double a, b;
    if (family == 0) {
      a = mk(r0, (int)(r2 % 600) - 300, r2 >> 63); b = mk(r1, (int)((r2 >> 16) % 600) - 300, (r2 >> 62) & 1);
    } else if (family == 1) {
      b = mk(r1, (int)((r2 >> 16) % 200) - 100, (r2 >> 62) & 1);
      const double q = mk(r0, (int)(r2 % 200) - 100, r2 >> 63);
      const double half_ulp = mk(0, (int)(r2 % 200) - 100 - 53, false);
      const int k = (int)((r2 >> 40) % 5) - 2;                          /* q + (k + 1/2) ulp, k = -2 .. 2 */
      a = __fma_rn(b, q, b * half_ulp * (double)(2 * k + 1));
    } else if (family == 2) {
      const unsigned long long near[4] = {0ull, 0x5555555555555ull, 0xFFFFFFFFFFFFFull, 0x8000000000000ull};
      const unsigned long long m = near[(r2 >> 20) & 3] + (long long)((r2 >> 24) % 2049) - 1024;
      b = mk(m, (int)((r2 >> 16) % 16) - 8, (r2 >> 62) & 1);
      a = mk(r0, (int)(r2 % 16) - 8, r2 >> 63);
    } else {
      const int ea[8] = {-1022, -470, -466, -464, 464, 466, 470, 1023};
      a = mk(r0, ea[r2 & 7], r2 >> 63); b = mk(r1, ea[(r2 >> 3) & 7], (r2 >> 62) & 1);
      const int sp = (int)((r2 >> 8) % 64);
      if (sp == 0) a = 0.0; if (sp == 1) b = 0.0; if (sp == 2) a = INFINITY; if (sp == 3) b = INFINITY; if (sp == 4) a = NAN; if (sp == 5) b = -0.0;
    }
    const double rb = 1.0 / b;
    const double want = a / b, got = div_known_rcp(a, b, rb);
    const double want6 = a / 6.0, got6 = div_by_6(a);
    if (!same_bits(want, got) || !same_bits(want6, got6)) {
      if (nbad == 0 && atomicAdd(bad + 1, 1ull) == 0) { example[0] = a; example[1] = b; example[2] = want; example[3] = got; example[4] = want6; example[5] = got6; }
      ++nbad;
    }
  }
  if (nbad) atomicAdd(bad, nbad);
}

int main(int argc, char** argv) {
  const int lg = argc > 1 ? atoi(argv[1]) : 34;
  cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
  const int blocks = p.multiProcessorCount * 8, threads = 256;
  const unsigned long long total = 1ull << lg, per_thread = total / ((unsigned long long)blocks * threads) + 1;
  unsigned long long* d_bad; double* d_ex;
  cudaMalloc(&d_bad, 16); cudaMalloc(&d_ex, 48);
  printf("# %s: div_known_rcp / div_by_6 (RRT_FAST_DIV = %d) against IEEE division, %llu pairs per family\n", p.name, RRT_FAST_DIV, per_thread * blocks * threads);
  const char* names[4] = {"random significands, exponents -300..300", "quotients next to a rounding boundary", "divisor significands near 1, 4/3, 3/2, 2 - 2^-52",
                          "edges of the safe exponent range, zeros, infinities, NaNs"};
  unsigned long long all = 0;
  for (int f = 0; f < 4; ++f) {
    cudaMemset(d_bad, 0, 16);
    check<<<blocks, threads>>>(f, per_thread, 1234567ull + f, d_bad, d_ex);
    unsigned long long bad[2]; double ex[6];
    cudaMemcpy(bad, d_bad, 16, cudaMemcpyDeviceToHost); cudaMemcpy(ex, d_ex, 48, cudaMemcpyDeviceToHost);
    printf("family %d (%s): %llu mismatches  %s\n", f, names[f], bad[0], cudaGetErrorString(cudaGetLastError()));
    if (bad[0]) printf("  first: a = %a  b = %a  a/b = %a  got %a   a/6 = %a  got %a\n", ex[0], ex[1], ex[2], ex[3], ex[4], ex[5]);
    all += bad[0];
  }
  printf("total mismatches: %llu\n", all);
  return all ? 1 : 0;
}
