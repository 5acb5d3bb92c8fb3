// Where do the two warps of a 64-thread CTA land?  Launches the Hanabi kernels' population (2 048 CTAs of 2 warps, ~10 KB
// of shared memory each, cooperative-style: all resident) and prints, for a few SMs, the %warpid (hardware warp slot) of
// warp 0 / warp 1 of every CTA on that SM, plus a histogram of slot % 4 (= the SM's scheduler the warp issues on).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/probes/_bin/warp_slots_probe tools/probes/warp_slots_probe.cu
#include <cstdio>
#include <cuda_runtime.h>

__global__ void __launch_bounds__(64, 14) probe(unsigned *out, volatile unsigned *go)
{
    extern __shared__ unsigned char smem[];
    unsigned smid, slot;
    asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
    asm volatile("mov.u32 %0, %%warpid;" : "=r"(slot));
    if ((threadIdx.x & 31) == 0) {
        out[(blockIdx.x * 2 + (threadIdx.x >> 5)) * 2 + 0] = smid;
        out[(blockIdx.x * 2 + (threadIdx.x >> 5)) * 2 + 1] = slot;
    }
    smem[threadIdx.x] = 1;
    // stay resident until every CTA has started (so that the placement is the steady-state one)
    if (threadIdx.x == 0) {
        atomicAdd((unsigned *)go, 1u);
        while (*go < gridDim.x) { }
    }
    __syncthreads();
}

int main()
{
    const int grid = 2048;
    unsigned *out, *go;
    cudaMalloc(&out, grid * 4 * sizeof(unsigned));
    cudaMalloc(&go, 4);
    cudaMemset(go, 0, 4);
    cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 12 * 1024);
    probe<<<grid, 64, 12 * 1024>>>(out, go);
    if (cudaDeviceSynchronize() != cudaSuccess) { printf("launch failed: %s\n", cudaGetErrorString(cudaGetLastError())); return 1; }
    static unsigned h[grid * 4];
    cudaMemcpy(h, out, sizeof(h), cudaMemcpyDeviceToHost);
    int hist[2][4] = {}, paired = 0, same_group = 0;
    for (int b = 0; b < grid; b++) {
        const unsigned s0 = h[b * 4 + 1], s1 = h[b * 4 + 3];
        hist[0][s0 % 4]++; hist[1][s1 % 4]++;
        paired += (s1 == s0 + 1 && s0 % 2 == 0);
        same_group += (s0 / 4 == s1 / 4);
    }
    printf("slot %% 4 of warp 0: %d %d %d %d   of warp 1: %d %d %d %d\n", hist[0][0], hist[0][1], hist[0][2], hist[0][3], hist[1][0],
           hist[1][1], hist[1][2], hist[1][3]);
    printf("CTAs whose warps sit in slots (2j, 2j+1): %d of %d; in the same group of four: %d\n", paired, grid, same_group);
    for (unsigned sm = 0; sm < 2; sm++) {
        printf("SM %u: (block: slot0 slot1)", sm);
        for (int b = 0; b < grid; b++)
            if (h[b * 4] == sm) printf("  %d: %u %u", b, h[b * 4 + 1], h[b * 4 + 3]);
        printf("\n");
    }
    return 0;
}
