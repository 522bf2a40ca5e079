// Micro-probe: how many clusters of each size can be co-resident with one 227 KB CTA per SM (GPC structure).
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o cluster_occ cluster_occ.cu && ./cluster_occ
#include <cstdio>
#include <cuda_runtime.h>
__global__ void k(int * p) { extern __shared__ char s[]; if (p) p[0] = s[0]; }
int main() {
    cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    cudaFuncSetAttribute(k, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
    for (int cs = 1; cs <= 16; ++cs) {
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(148 / cs * cs); cfg.blockDim = dim3(640); cfg.dynamicSmemBytes = 227 * 1024;
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeClusterDimension;
        at[0].val.clusterDim.x = cs; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
        cfg.attrs = at; cfg.numAttrs = 1;
        int n = 0;
        cudaError_t e = cudaOccupancyMaxActiveClusters(&n, k, &cfg);
        printf("cluster %2d: %3d clusters = %3d SMs  (%s)\n", cs, n, n * cs, cudaGetErrorString(e));
    }
    return 0;
}
