#include <cstdio>
#include <cuda_runtime.h>
__global__ void lat(double *out, long long *cyc, double a, double b, int n) {
  double x = a;
  long long t0 = clock64();
  for (int i = 0; i < n; i++) { x = x + b; x = x + b; x = x + b; x = x + b; }
  long long t1 = clock64();
  double y = a;
  for (int i = 0; i < n; i++) { y = y * b; y = y * b; y = y * b; y = y * b; }
  long long t2 = clock64();
  __shared__ double sh[64];
  sh[threadIdx.x] = x;
  __syncthreads();
  int idx = threadIdx.x;
  long long t3 = clock64();
  double z = 0;
  for (int i = 0; i < n; i++) { z += sh[idx]; idx = (idx + (int)z) & 31; }
  long long t4 = clock64();
  out[threadIdx.x] = x + y + z;
  if (threadIdx.x == 0) { cyc[0] = t1 - t0; cyc[1] = t2 - t1; cyc[2] = t4 - t3; }
}
__global__ void thr(double *out, double a, double b, int n) {
  double x0 = a, x1 = a + 1, x2 = a + 2, x3 = a + 3, x4 = a + 4, x5 = a + 5, x6 = a + 6, x7 = a + 7;
  for (int i = 0; i < n; i++) { x0 = x0 * b + a; x1 = x1 * b + a; x2 = x2 * b + a; x3 = x3 * b + a; x4 = x4 * b + a; x5 = x5 * b + a; x6 = x6 * b + a; x7 = x7 * b + a; }
  out[blockIdx.x * blockDim.x + threadIdx.x] = x0 + x1 + x2 + x3 + x4 + x5 + x6 + x7;
}
int main() {
  double *out; long long *cyc;
  cudaMalloc(&out, 1 << 24); cudaMalloc(&cyc, 64);
  int n = 10000;
  lat<<<1, 32>>>(out, cyc, 1.0, 1.0000001, n);
  long long h[3]; cudaMemcpy(h, cyc, 24, cudaMemcpyDeviceToHost);
  printf("DADD dependent latency %.2f cycles, DMUL %.2f cycles, LDS+DADD chain %.2f\n", h[0] / (4.0 * n), h[1] / (4.0 * n), h[2] / (double)n);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  int blocks = 148 * 8, threads = 256; n = 20000;
  thr<<<blocks, threads>>>(out, 1.0, 0.999, 100);
  cudaEventRecord(e0); thr<<<blocks, threads>>>(out, 1.0, 0.999, n); cudaEventRecord(e1); cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  double ops = (double)blocks * threads * n * 8 * 2;   // mul + add separately (no fma)
  printf("DP throughput (fmad=false): %.2f Tops/s (mul or add)\n", ops / ms / 1e9);
  return 0;
}
