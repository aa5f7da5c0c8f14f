// Microbenchmark (GPU box): cycles per hardware cluster barrier (arrive.release + wait.acquire), alone and with a
// DSMEM push before it and a local read after it, for cluster sizes 2..16 and CTA sizes 64..256.
//   nvcc -gencode arch=compute_100a,code=sm_100a -o clbar clbar.cu && ./clbar
#include <cstdio>
#include <cuda_runtime.h>
#include <cooperative_groups.h>
namespace cg = cooperative_groups;

__global__ void k_bar(int n, int push, long long *out, double *sink) {
  cg::cluster_group cluster = cg::this_cluster();
  __shared__ double xs[2][256];
  const int tid = threadIdx.x, CL = (int)cluster.dim_blocks().x;
  xs[0][tid] = 0; xs[1][tid] = 0;
  cluster.sync();
  const unsigned base = (unsigned)__cvta_generic_to_shared(&xs[0][0]);
  double acc = tid;
  const long long t0 = clock64();
  for (int T = 0; T < n; T++) {
    if (push) {
      unsigned remote;
      const unsigned local = base + (unsigned)(((T & 1) * 256 + tid) * 8);
      asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(local), "r"((blockIdx.x + 1) % CL));
      asm volatile("st.shared::cluster.f64 [%0], %1;" ::"r"(remote), "d"(acc) : "memory");
    }
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
    if (push) acc += *(volatile double *)&xs[T & 1][tid];
  }
  const long long t1 = clock64();
  if (tid == 0 && blockIdx.x == 0) out[0] = t1 - t0;
  sink[blockIdx.x * blockDim.x + tid] = acc;
  cluster.sync();
}

int main() {
  long long *out; double *sink;
  cudaMalloc(&out, 8); cudaMalloc(&sink, 16 * 256 * 8);
  cudaFuncSetAttribute(k_bar, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
  const int n = 20000;
  for (int cl : {1, 2, 4, 8, 16})
    for (int th : {64, 128, 256})
      for (int push = 0; push < 2; push++) {
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(cl); cfg.blockDim = dim3(th);
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeClusterDimension;
        at[0].val.clusterDim.x = cl; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
        cfg.attrs = at; cfg.numAttrs = 1;
        cudaError_t e = cudaLaunchKernelEx(&cfg, k_bar, n, push, out, sink);
        cudaDeviceSynchronize();
        long long c = 0;
        cudaMemcpy(&c, out, 8, cudaMemcpyDeviceToHost);
        printf("cluster %2d x %3d threads, push %d: %6.1f cycles per round (%s)\n", cl, th, push, (double)c / n, cudaGetErrorString(e));
      }
  return 0;
}
