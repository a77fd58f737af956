// How many thread-block clusters of each size fit a B200 at one 512-thread CTA per SM, and which SMs a launch lands on.
#include <cstdio>
#include <cuda_runtime.h>
__global__ void __launch_bounds__(512, 1) k(int *smid, int *cid) {
    extern __shared__ float sm[];
    unsigned s, c;
    asm volatile("mov.u32 %0, %%smid;" : "=r"(s));
    asm volatile("mov.u32 %0, %%clusterid.x;" : "=r"(c));
    if (threadIdx.x == 0) { smid[blockIdx.x] = (int)s; cid[blockIdx.x] = (int)c; }
    sm[threadIdx.x] = 0;
}
int main() {
    cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    cudaFuncSetAttribute(k, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
    int *smid, *cid;
    cudaMalloc(&smid, 4096); cudaMalloc(&cid, 4096);
    for (int smem : {0, 110 * 1024, 200 * 1024})
        for (int cs : {1, 2, 4, 8, 16}) {
            cudaLaunchConfig_t cfg = {};
            cfg.gridDim = dim3(cs * 16); cfg.blockDim = dim3(512); cfg.dynamicSmemBytes = smem;
            cudaLaunchAttribute at[1];
            at[0].id = cudaLaunchAttributeClusterDimension; at[0].val.clusterDim.x = cs; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
            cfg.attrs = at; cfg.numAttrs = 1;
            int n = -1;
            cudaError_t e = cudaOccupancyMaxActiveClusters(&n, k, &cfg);
            printf("smem %6d cluster %2d: max active clusters %d (%s) -> %d CTAs\n", smem, cs, n, cudaGetErrorString(e), n * cs);
        }
    for (int cs : {8, 16}) {
        const int grid = cs == 8 ? 128 : 112;
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(grid); cfg.blockDim = dim3(512); cfg.dynamicSmemBytes = 110 * 1024;
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeClusterDimension; at[0].val.clusterDim.x = cs; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
        cfg.attrs = at; cfg.numAttrs = 1;
        cudaError_t e = cudaLaunchKernelEx(&cfg, k, smid, cid);
        cudaDeviceSynchronize();
        int hs[256], hc[256];
        cudaMemcpy(hs, smid, grid * 4, cudaMemcpyDeviceToHost); cudaMemcpy(hc, cid, grid * 4, cudaMemcpyDeviceToHost);
        printf("cluster %d launch (%s): ", cs, cudaGetErrorString(e));
        for (int i = 0; i < grid; i++) { if (i % cs == 0) printf("| c%d: ", hc[i]); printf("%d ", hs[i]); }
        printf("\n");
    }
    return 0;
}
