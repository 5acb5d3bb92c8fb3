// Per-SM store-path probe: 14 single-warp CTAs per SM (the Hanabi writer population) stream stores into an L2-resident
// region (no HBM bound) with 16-byte or 32-byte stores per lane.  Prints bytes per clock per SM.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/probes/_bin/store_width_probe tools/probes/store_width_probe.cu
#include <cstdio>
#include <cuda_runtime.h>

template <int WIDE>
__global__ void __launch_bounds__(32) probe(uint4 *out, int iters, size_t region_chunks)
{
    const size_t warp_base = (size_t)blockIdx.x * 4096;   // each warp cycles through its own 64 KB window
    const int lane = threadIdx.x;
    uint4 v = make_uint4(lane, 1, 2, 3);
    for (int i = 0; i < iters; i++) {
        if (WIDE) {
            const size_t c = (warp_base + (size_t)((i * 64) % 4096) + 2 * lane) % region_chunks;
            asm volatile("st.global.v8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"l"(out + c), "r"(v.x), "r"(v.y), "r"(v.z),
                         "r"(v.w), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
        } else {
            const size_t c = (warp_base + (size_t)((i * 32) % 4096) + lane) % region_chunks;
            out[c] = v;
        }
        v.x += i;
    }
}

int main()
{
    int sms = 0, clk = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
    const size_t region_chunks = (size_t)48 << 16;   // 48 MB of 16-byte chunks: inside the 126 MB L2
    uint4 *buf;
    cudaMalloc(&buf, region_chunks * 16);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    for (int per_sm : {4, 8, 14, 28}) {
        for (int wide = 0; wide < 2; wide++) {
            const int iters = wide ? 20000 : 40000;   // same bytes
            for (int rep = 0; rep < 2; rep++) {
                cudaEventRecord(e0);
                if (wide) probe<1><<<sms * per_sm, 32>>>(buf, iters, region_chunks);
                else probe<0><<<sms * per_sm, 32>>>(buf, iters, region_chunks);
                cudaEventRecord(e1);
                cudaEventSynchronize(e1);
            }
            float ms = 0;
            cudaEventElapsedTime(&ms, e0, e1);
            const double bytes = (double)sms * per_sm * iters * (wide ? 1024.0 : 512.0);
            printf("%2d warps/SM, %2d-byte stores per lane: %7.1f GB/s total, %5.1f B/clk/SM (at %d MHz nominal)  [%s]\n", per_sm,
                   wide ? 32 : 16, bytes / ms / 1e6, bytes / (ms * 1e-3) / sms / (clk * 1e3), clk / 1000, cudaGetErrorString(cudaGetLastError()));
        }
    }
    return 0;
}
