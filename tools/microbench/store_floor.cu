// Microbenchmark: how fast can 34 MB (32768 worlds x 1040 B) be written per launch, back to back, into a ring of
// buffers larger than L2?  Variants: (a) bulk smem->global copies (one warp per 16.6 KB tile, as step_tile does),
// (b) plain 128-bit register stores, (c) bulk with bigger CTAs.  Build: nvcc -arch=sm_100a -O3 -o store_floor store_floor.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); return 1; } } while (0)

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

// (a) one warp per tile, `chunks` bulk copies of `bytes` each
__global__ void __launch_bounds__(32) k_bulk(uint8_t *dst, uint32_t bytes, int chunks, int pdl)
{
    extern __shared__ __align__(128) uint8_t smem[];
    if (pdl) { asm volatile("griddepcontrol.wait;" ::: "memory"); asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
    for (uint32_t i = threadIdx.x; i < bytes / 16; i += 32) reinterpret_cast<uint4 *>(smem)[i] = make_uint4(i, 1, 2, 3);
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncwarp();
    if (threadIdx.x == 0) {
        for (int c = 0; c < chunks; c++) {
            uint8_t *d = dst + ((size_t)c * gridDim.x + blockIdx.x) * bytes;
            asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(d), "r"(smem_u32(smem)), "r"(bytes) : "memory");
        }
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
    }
}

// (a2) as (a) with an L2 evict_first cache hint on the bulk copies
__global__ void __launch_bounds__(32) k_bulk_hint(uint8_t *dst, uint32_t bytes, int chunks, int pdl)
{
    extern __shared__ __align__(128) uint8_t smem[];
    if (pdl) { asm volatile("griddepcontrol.wait;" ::: "memory"); asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
    for (uint32_t i = threadIdx.x; i < bytes / 16; i += 32) reinterpret_cast<uint4 *>(smem)[i] = make_uint4(i, 1, 2, 3);
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncwarp();
    if (threadIdx.x == 0) {
        uint64_t pol;
        asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
        for (int c = 0; c < chunks; c++) {
            uint8_t *d = dst + ((size_t)c * gridDim.x + blockIdx.x) * bytes;
            asm volatile("cp.async.bulk.global.shared::cta.bulk_group.L2::cache_hint [%0], [%1], %2, %3;" ::"l"(d), "r"(smem_u32(smem)), "r"(bytes), "l"(pol) : "memory");
        }
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
    }
}

// (b2) 128-bit streaming stores (st.global.cs)
__global__ void k_stg_cs(uint4 *dst, size_t n16, int pdl)
{
    if (pdl) { asm volatile("griddepcontrol.wait;" ::: "memory"); asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n16; i += (size_t)gridDim.x * blockDim.x)
        __stcs(dst + i, make_uint4((uint32_t)i, 1, 2, 3));
}

// (b) grid-stride 128-bit stores
__global__ void k_stg(uint4 *dst, size_t n16, int pdl)
{
    if (pdl) { asm volatile("griddepcontrol.wait;" ::: "memory"); asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n16; i += (size_t)gridDim.x * blockDim.x)
        dst[i] = make_uint4((uint32_t)i, 1, 2, 3);
}

__global__ void k_empty(int pdl)
{
    if (pdl) { asm volatile("griddepcontrol.wait;" ::: "memory"); asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
}

template <class F> float time_loop(F f, int iters, cudaStream_t st)
{
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    for (int i = 0; i < 20; i++) f(i);
    cudaEventRecord(a, st);
    for (int i = 0; i < iters; i++) f(i);
    cudaEventRecord(b, st);
    cudaEventSynchronize(b);
    float ms; cudaEventElapsedTime(&ms, a, b);
    return ms * 1e3f / iters;
}

template <class... KA, class... A>
void launch(void (*k)(KA...), unsigned grid, unsigned block, size_t smem, cudaStream_t st, int pdl, A... args)
{
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid); cfg.blockDim = dim3(block); cfg.dynamicSmemBytes = smem; cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr; cfg.numAttrs = pdl ? 1 : 0;
    cudaLaunchKernelEx(&cfg, k, KA(args)...);
}

int main()
{
    const size_t N = 32768, SC = 520, P = 2, per = N * SC * P;   // 34 MB
    const int RING = 8;
    uint8_t *buf; CK(cudaMalloc(&buf, per * RING));
    cudaStream_t st; CK(cudaStreamCreate(&st));
    CK(cudaFuncSetAttribute(k_bulk, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024));
    CK(cudaFuncSetAttribute(k_bulk_hint, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024));
    const int iters = 2000;
    for (int pdl = 1; pdl < 2; pdl++) {
        printf("---- pdl=%d\n", pdl);
        float t = time_loop([&](int i) { launch(k_empty, 1024, 32, 0, st, pdl, pdl); }, iters, st);
        printf("empty 1024x32                       %7.2f us\n", t);
        t = time_loop([&](int i) { launch(k_bulk, 1024, 32, 32 * SC, st, pdl, buf + (i % RING) * per, (uint32_t)(32 * SC), 2, pdl); }, iters, st);
        printf("bulk  1024 warps x 2 x 16.6KB       %7.2f us  %7.1f GB/s\n", t, per / t / 1e3);
        t = time_loop([&](int i) { launch(k_bulk_hint, 1024, 32, 32 * SC, st, pdl, buf + (i % RING) * per, (uint32_t)(32 * SC), 2, pdl); }, iters, st);
        printf("bulk+evict_first 1024 x 2 x 16.6KB   %7.2f us  %7.1f GB/s\n", t, per / t / 1e3);
        t = time_loop([&](int i) { launch(k_bulk_hint, 2048, 32, 16 * SC, st, pdl, buf + (i % RING) * per, (uint32_t)(16 * SC), 2, pdl); }, iters, st);
        printf("bulk+evict_first 2048 x 2 x 8.3KB    %7.2f us  %7.1f GB/s\n", t, per / t / 1e3);
        t = time_loop([&](int i) { launch(k_bulk, 2048, 32, 16 * SC, st, pdl, buf + (i % RING) * per, (uint32_t)(16 * SC), 2, pdl); }, iters, st);
        printf("bulk  2048 warps x 2 x 8.3KB        %7.2f us  %7.1f GB/s\n", t, per / t / 1e3);
        t = time_loop([&](int i) { launch(k_stg_cs, 148 * 4, 256, 0, st, pdl, (uint4 *)(buf + (i % RING) * per), per / 16, pdl); }, iters, st);
        printf("stg.cs 592 x 256                    %7.2f us  %7.1f GB/s\n", t, per / t / 1e3);
        t = time_loop([&](int i) { launch(k_bulk, 2048, 32, 32 * SC, st, pdl, buf + (i % RING) * per, (uint32_t)(32 * SC), 1, pdl); }, iters, st);
        printf("bulk  2048 warps x 1 x 16.6KB       %7.2f us  %7.1f GB/s\n", t, per / t / 1e3);
        t = time_loop([&](int i) { launch(k_bulk, 512, 32, 64 * SC, st, pdl, buf + (i % RING) * per, (uint32_t)(64 * SC), 2, pdl); }, iters, st);
        printf("bulk   512 warps x 2 x 33KB         %7.2f us  %7.1f GB/s\n", t, per / t / 1e3);
        t = time_loop([&](int i) { launch(k_bulk, 4096, 32, 8 * SC, st, pdl, buf + (i % RING) * per, (uint32_t)(8 * SC), 2, pdl); }, iters, st);
        printf("bulk  4096 warps x 2 x 4.2KB        %7.2f us  %7.1f GB/s\n", t, per / t / 1e3);
        t = time_loop([&](int i) { launch(k_bulk, 1024, 32, 32 * SC, st, pdl, buf, (uint32_t)(32 * SC), 2, pdl); }, iters, st);
        printf("bulk  1024 warps, SAME buffer (L2)  %7.2f us  %7.1f GB/s\n", t, per / t / 1e3);
        for (int block : {128, 256, 512}) for (int mult : {2, 4, 8}) {
            t = time_loop([&](int i) { launch(k_stg, 148 * mult, block, 0, st, pdl, (uint4 *)(buf + (i % RING) * per), per / 16, pdl); }, iters, st);
            printf("stg   %4d x %3d                     %7.2f us  %7.1f GB/s\n", 148 * mult, block, t, per / t / 1e3);
        }
        t = time_loop([&](int i) { launch(k_stg, 148 * 8, 256, 0, st, pdl, (uint4 *)buf, per / 16, pdl); }, iters, st);
        printf("stg   1184 x 256 SAME buffer (L2)   %7.2f us  %7.1f GB/s\n", t, per / t / 1e3);
    }
    // large single launch for reference: 8 x 34 MB in one kernel
    float t = time_loop([&](int i) { launch(k_stg, 148 * 8, 256, 0, st, 0, (uint4 *)buf, per * RING / 16, 0); }, 200, st);
    printf("stg   one launch of 273 MB          %7.2f us  %7.1f GB/s\n", t, per * RING / t / 1e3);
    CK(cudaDeviceSynchronize());
    printf("done\n");
    return 0;
}
