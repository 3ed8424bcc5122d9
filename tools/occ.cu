// tools/occ.cu -- how many clusters of G CTAs (1024 threads, `smem` bytes each) are co-resident?
#include <cstdio>
#include <cuda_runtime.h>
__global__ void __launch_bounds__(1024, 1) dummy(int *p) { extern __shared__ char s[]; if (p) p[0] = s[0]; }
int main() {
    cudaDeviceProp pr; cudaGetDeviceProperties(&pr, 0);
    printf("%s SMs %d\n", pr.name, pr.multiProcessorCount);
    cudaFuncSetAttribute(dummy, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
    for (int smem : {232448, 200000, 160000}) {
        cudaFuncSetAttribute(dummy, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        for (int G = 1; G <= 16; G++) {
            cudaLaunchConfig_t cfg = {}; cfg.gridDim = dim3(G * 64); cfg.blockDim = dim3(1024); cfg.dynamicSmemBytes = smem;
            cudaLaunchAttribute a[1]; a[0].id = cudaLaunchAttributeClusterDimension; a[0].val.clusterDim.x = G; a[0].val.clusterDim.y = 1; a[0].val.clusterDim.z = 1;
            cfg.attrs = a; cfg.numAttrs = 1;
            int n = 0; cudaError_t e = cudaOccupancyMaxActiveClusters(&n, dummy, &cfg);
            printf("smem %d G %2d: clusters %3d -> SMs %3d %s\n", smem, G, n, n * G, e == cudaSuccess ? "" : cudaGetErrorString(e));
            cudaGetLastError();
        }
    }
    return 0;
}
