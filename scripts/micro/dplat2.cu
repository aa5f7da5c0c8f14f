// Single-warp FP64 issue behaviour on sm_100a: cycles per DP instruction at ILP 1..8, with
// 1..8 warps per SM sub-partition.  nvcc -O3 -fmad=false.
#include <cstdio>
#include <cuda_runtime.h>
template <int ILP>
__global__ void k(double *out, long long *cyc, double a, double b, int n) {
  double x[ILP];
#pragma unroll
  for (int i = 0; i < ILP; i++) x[i] = a + i;
  __syncthreads();
  long long t0 = clock64();
  for (int it = 0; it < n; it++) {
#pragma unroll
    for (int i = 0; i < ILP; i++) x[i] = x[i] + b;
#pragma unroll
    for (int i = 0; i < ILP; i++) x[i] = x[i] * b;
  }
  long long t1 = clock64();
  double s = 0;
#pragma unroll
  for (int i = 0; i < ILP; i++) s += x[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0 && blockIdx.x == 0) cyc[0] = t1 - t0;
}
template <int ILP>
void run(double *out, long long *cyc, int threads) {
  int n = 4000;
  k<ILP><<<1, threads>>>(out, cyc, 1.0, 1.0000001, n);
  long long h; cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
  printf("threads %4d ILP %d: %.2f cycles per DP warp-instruction per warp (%.2f per SMSP-instr)\n", threads, ILP,
         (double)h / (2.0 * n * ILP), (double)h / (2.0 * n * ILP) / ((threads + 127) / 128));
}
int main() {
  double *out; long long *cyc;
  cudaMalloc(&out, 1 << 20); cudaMalloc(&cyc, 64);
  for (int threads : {32, 128, 256, 512, 1024}) {
    run<1>(out, cyc, threads); run<2>(out, cyc, threads); run<4>(out, cyc, threads); run<8>(out, cyc, threads);
  }
  return 0;
}
