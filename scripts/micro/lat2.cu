// Microbenchmark (GPU box): dependent-issue latencies of the instructions on k_pgsb's critical path.
#include <cstdio>
#include <cuda_runtime.h>
__global__ void lat(double *out, long long *cyc, double a, double b, int n) {
  __shared__ __align__(16) double sh[256];
  for (int i = threadIdx.x; i < 256; i += 32) sh[i] = a;
  __syncwarp();
  long long t[12];
  double x = a;
  t[0] = clock64();
  for (int i = 0; i < n; i++) { x = x + b; x = x + b; x = x + b; x = x + b; }
  t[1] = clock64();
  double y = a;
  for (int i = 0; i < n; i++) { y = y * b; y = y * b; y = y * b; y = y * b; }
  t[2] = clock64();
  // shuffle of a double (2 x SHFL.IDX) dependent chain
  double s = x; int src = (threadIdx.x + 1) & 31;
  for (int i = 0; i < n; i++) { s = __shfl_sync(0xffffffffu, s, src); s = __shfl_sync(0xffffffffu, s, src); s = __shfl_sync(0xffffffffu, s, src); s = __shfl_sync(0xffffffffu, s, src); }
  t[3] = clock64();
  // redux min dependent chain
  int r = threadIdx.x + (int)a;
  for (int i = 0; i < n; i++) { r = __reduce_min_sync(0xffffffffu, r + threadIdx.x); r = __reduce_min_sync(0xffffffffu, r + threadIdx.x); r = __reduce_min_sync(0xffffffffu, r + threadIdx.x); r = __reduce_min_sync(0xffffffffu, r + threadIdx.x); }
  t[4] = clock64();
  // vote chain
  int v = threadIdx.x;
  for (int i = 0; i < n; i++) { v += __any_sync(0xffffffffu, v & 1); v += __any_sync(0xffffffffu, v & 2); v += __any_sync(0xffffffffu, v & 1); v += __any_sync(0xffffffffu, v & 2); }
  t[5] = clock64();
  // LDS.128 pointer chase
  int idx = threadIdx.x; double z = 0;
  for (int i = 0; i < n; i++) { double2 q = *reinterpret_cast<double2 *>(&sh[2 * idx]); z += q.x; idx = (idx + (int)q.y) & 31; }
  t[6] = clock64();
  // select on compare chain: x = (x < b) ? x + b : x   (DSETP + FSEL x2 + DADD)
  double c = a;
  for (int i = 0; i < n; i++) { c = (c < b) ? c : c + b; c = (c < b) ? c : c + b; c = (c < b) ? c : c + b; c = (c < b) ? c : c + b; }
  t[7] = clock64();
  out[threadIdx.x] = x + y + s + r + v + z + c;
  if (threadIdx.x == 0) for (int k = 0; k < 7; k++) cyc[k] = t[k + 1] - t[k];
}
int main() {
  double *out; long long *cyc;
  cudaMalloc(&out, 4096); cudaMalloc(&cyc, 128);
  int n = 10000;
  lat<<<1, 32>>>(out, cyc, 1.0, 1.0000001, n);
  long long h[7]; cudaMemcpy(h, cyc, 56, cudaMemcpyDeviceToHost);
  const char *nm[7] = {"DADD", "DMUL", "SHFL double (2 x SHFL.IDX)", "REDUX.MIN + IADD", "VOTE.ANY + IADD", "LDS.128 + DADD + IADD chase", "DSETP + FSEL + DADD"};
  for (int k = 0; k < 7; k++) printf("%-32s %.2f cycles\n", nm[k], h[k] / ((k == 5 ? 1.0 : 4.0) * n));
  return 0;
}
