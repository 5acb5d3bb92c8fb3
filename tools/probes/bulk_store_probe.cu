// Microbenchmark behind DESIGN 3.2: how fast can 2048 single-warp CTAs stream 8 KB shared-memory tiles to HBM with
// cp.async.bulk (the fused Overcooked rollout's store pattern), with no game logic at all?
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o bulk_store_probe bulk_store_probe.cu && ./bulk_store_probe
// Variants: pending = bulk groups allowed in flight at the wait (0 = the kernel's wait-after-every-store),
// work = dependent ALU iterations between stores (stands in for the game logic), chunk = bytes per store.
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

template <int PENDING>
__global__ void __launch_bounds__(32) probe(uint8_t *obs, size_t slice_bytes, int slices, uint32_t chunk, int T, int P,
                                            size_t observer_bytes, int work, uint32_t *sink)
{
    extern __shared__ __align__(128) uint8_t smem[];
    const int lane = threadIdx.x;
    for (uint32_t i = lane * 4; i < chunk; i += 128) *reinterpret_cast<uint32_t *>(smem + i) = i;
    __syncwarp();
    uint32_t x = lane;
    int slice = 0;
    for (int t = 0; t < T; t++) {
        for (int k = 0; k < work; k++) x = x * 1664525u + 1013904223u;   // dependent chain
        for (int p = 0; p < P; p++) {
            if (lane == 0) {
                asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(PENDING) : "memory");
            }
            __syncwarp();
            smem[lane * 4] = (uint8_t)x;   // the tile changes between stores
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            __syncwarp();
            if (lane == 0) {
                uint8_t *dst = obs + (size_t)slice * slice_bytes + (size_t)p * observer_bytes + (size_t)blockIdx.x * chunk;
                asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(smem_u32(smem)),
                             "r"(chunk) : "memory");
                asm volatile("cp.async.bulk.commit_group;" ::: "memory");
            }
        }
        if (++slice == slices) slice = 0;
    }
    if (lane == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
    if (x == 0xdeadbeef) *sink = x;
}

template <int PENDING>
static void run(const char *name, uint8_t *obs, int ctas, uint32_t chunk, int work, int T, uint32_t *sink)
{
    const int P = 2, slices = 8;
    const size_t observer = (size_t)ctas * chunk, slice = observer * P;
    cudaFuncSetAttribute(probe<PENDING>, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    float best = 1e30f;
    for (int rep = 0; rep < 4; rep++) {
        cudaEventRecord(e0);
        probe<PENDING><<<ctas, 32, chunk + 128>>>(obs, slice, slices, chunk, T, P, observer, work, sink);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms;
        cudaEventElapsedTime(&ms, e0, e1);
        if (rep && ms < best) best = ms;
    }
    const double bytes = (double)slice * T;
    printf("%-28s ctas %5d chunk %6u work %5d : %7.2f us/step  %7.1f GB/s  (%s)\n", name, ctas, chunk, work,
           1e3 * best / T, bytes / (best * 1e-3) / 1e9, cudaGetErrorString(cudaGetLastError()));
}

int main()
{
    uint8_t *obs;
    uint32_t *sink;
    const size_t cap = (size_t)8 * 2 * 4096 * 33280;
    cudaMalloc(&obs, cap);
    cudaMalloc(&sink, 4);
    const int T = 1000;
    for (int work : {0, 200, 500, 1000, 2000}) {
        run<0>("wait every store", obs, 2048, 8320, work, T, sink);
        run<1>("one store in flight", obs, 2048, 8320, work, T, sink);
    }
    run<0>("wait every store", obs, 4096, 4160, 500, T, sink);
    run<1>("one store in flight", obs, 4096, 4160, 500, T, sink);
    run<0>("wait every store", obs, 1024, 16640, 500, T, sink);
    run<1>("one store in flight", obs, 1024, 16640, 500, T, sink);
    run<0>("wait every store", obs, 4096, 8320, 500, T, sink);
    run<0>("wait every store", obs, 2048, 33280, 500, T / 2, sink);
    run<0>("wait every store", obs, 2072, 8320, 500, T, sink);   // 14 x 148 CTAs
    return 0;
}
