// How many clusters of size S (one 512-thread CTA with ~223 KB of shared memory per SM) can be co-resident on this GPU?
//   nvcc -gencode arch=compute_100a,code=sm_100a -o cluster_probe scripts/cluster_probe.cu && ./cluster_probe
#include <cstdio>
#include <cuda_runtime.h>
__global__ void k(float* p) {
  extern __shared__ float s[];
  s[threadIdx.x] = 1.f;
  if (p) p[0] = s[0];
}
int main() {
  const size_t smem = 228608;
  cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  cudaFuncSetAttribute(k, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
  for (int S : {1, 2, 4, 6, 7, 8, 9, 10, 12, 16}) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(S * 64);
    cfg.blockDim = dim3(512);
    cfg.dynamicSmemBytes = smem;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = S;
    at[0].val.clusterDim.y = 1;
    at[0].val.clusterDim.z = 1;
    cfg.attrs = at;
    cfg.numAttrs = 1;
    int n = -1;
    cudaError_t e = cudaOccupancyMaxActiveClusters(&n, k, &cfg);
    printf("cluster size %2d: max active clusters %3d -> %3d SMs (%s)\n", S, n, n * S, cudaGetErrorString(e));
  }
  return 0;
}
