// Probe: how many thread-block clusters of a given size can be co-resident (1 CTA/SM, ~220 KB smem each)?
#include <cstdio>
#include <cuda_runtime.h>
__global__ void k(int * p) { extern __shared__ char s[]; if (p) p[0] = s[0]; }
int main() {
    cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
    cudaFuncSetAttribute(k, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
    for (int cs : {1, 2, 4, 8, 16}) {
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(148 / cs * cs); cfg.blockDim = dim3(320); cfg.dynamicSmemBytes = 220 * 1024;
        cudaLaunchAttribute a[1];
        a[0].id = cudaLaunchAttributeClusterDimension; a[0].val.clusterDim.x = cs; a[0].val.clusterDim.y = 1; a[0].val.clusterDim.z = 1;
        cfg.attrs = a; cfg.numAttrs = 1;
        int n = -1;
        cudaError_t e = cudaOccupancyMaxActiveClusters(&n, k, &cfg);
        printf("cluster %2d: max active clusters %d (%d SMs) %s\n", cs, n, n * cs, cudaGetErrorString(e));
    }
    cudaDeviceProp pr; cudaGetDeviceProperties(&pr, 0);
    printf("SMs %d  L2 %d MB\n", pr.multiProcessorCount, pr.l2CacheSize >> 20);
    return 0;
}
