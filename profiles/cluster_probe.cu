// cluster_probe.cu -- how many thread-block clusters of a persistent, shared-memory-filling kernel (one block per SM,
// 384 threads, ~214 KB dynamic shared memory) can be resident on this GPU, per cluster size.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 profiles/cluster_probe.cu -o profiles/cluster_probe
#include <cuda_runtime.h>
#include <cstdio>
__global__ void __launch_bounds__(384, 1) k(int* p) {
  extern __shared__ char s[];
  if (p && threadIdx.x == 0) p[blockIdx.x] = s[0];
}
int main() {
  cudaDeviceProp prop;
  cudaGetDeviceProperties(&prop, 0);
  const int smem = 214 * 1024;
  cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  cudaFuncSetAttribute(k, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
  printf("%s, %d SMs\n", prop.name, prop.multiProcessorCount);
  for (int cs : {1, 2, 4, 8, 16}) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(prop.multiProcessorCount / cs * cs);
    cfg.blockDim = dim3(384);
    cfg.dynamicSmemBytes = smem;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = cs; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    cfg.attrs = at; cfg.numAttrs = 1;
    int n = 0;
    cudaError_t e = cudaOccupancyMaxActiveClusters(&n, k, &cfg);
    printf("cluster size %2d: max active clusters %d (%d blocks)  %s\n", cs, n, n * cs, e == cudaSuccess ? "" : cudaGetErrorString(e));
    cudaGetLastError();
  }
  return 0;
}
