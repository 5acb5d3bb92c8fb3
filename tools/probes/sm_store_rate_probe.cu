// Is every SM's path to memory equally fast?  One CTA per SM (14 warps, the Hanabi writer population) streams 16-byte
// stores (st.global.cs) into its own region of a buffer larger than L2, all SMs at once; prints the time every SM needed
// for the same number of bytes, sorted, with its %smid — and the same again with only every other TPC active.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/probes/_bin/sm_store_rate_probe tools/probes/sm_store_rate_probe.cu
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cuda_runtime.h>

__global__ void __launch_bounds__(448) probe(uint4 *out, size_t chunks_per_cta, int iters, unsigned long long *times, unsigned *smids,
                                             int active_mod)
{
    unsigned smid;
    asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
    if (active_mod > 1 && (smid / 2) % active_mod != 0) return;
    uint4 *mine = out + (size_t)blockIdx.x * chunks_per_cta;
    unsigned long long t0, t1;
    __syncthreads();
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
    const uint4 v = make_uint4(threadIdx.x, 1, 2, 3);
    for (int it = 0; it < iters; it++)
        for (size_t c = threadIdx.x; c < chunks_per_cta; c += blockDim.x) __stcs(mine + c, v);
    __syncthreads();
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
    if (threadIdx.x == 0) { times[blockIdx.x] = t1 - t0; smids[blockIdx.x] = smid; }
}

int main(int argc, char **argv)
{
    int sms = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    const size_t kb = argc > 1 ? atoi(argv[1]) : 4096;   // per CTA and pass: 4 MB = 592 MB in all, > L2; 512 KB = L2-resident
    const size_t chunks = (kb << 10) / 16;
    const int passes = (int)((16u << 20) / (kb << 10));
    printf("%zu KB per SM and pass, %d passes\n", kb, passes);
    uint4 *out;
    unsigned long long *times;
    unsigned *smids;
    cudaMalloc(&out, (size_t)sms * chunks * 16);
    cudaMalloc(&times, sms * 8);
    cudaMalloc(&smids, sms * 4);
    for (int mod = 1; mod <= 2; mod++) {
        cudaMemset(times, 0, sms * 8);
        for (int rep = 0; rep < 2; rep++) probe<<<sms, 448>>>(out, chunks, passes, times, smids, mod);
        if (cudaDeviceSynchronize() != cudaSuccess) { printf("failed: %s\n", cudaGetErrorString(cudaGetLastError())); return 1; }
        std::vector<unsigned long long> t(sms);
        std::vector<unsigned> id(sms);
        cudaMemcpy(t.data(), times, sms * 8, cudaMemcpyDeviceToHost);
        cudaMemcpy(id.data(), smids, sms * 4, cudaMemcpyDeviceToHost);
        std::vector<std::pair<double, unsigned>> v;
        for (int i = 0; i < sms; i++)
            if (t[i]) v.push_back({16.0 * chunks * passes / (double)t[i], id[i]});   // GB/s of this SM
        std::sort(v.begin(), v.end());
        double sum = 0;
        for (auto &p : v) sum += p.first;
        printf("every %d%s TPC active: %zu SMs, per-SM store rate GB/s: min %.1f  p10 %.1f  median %.1f  p90 %.1f  max %.1f  (sum %.0f)\n", mod,
               mod == 1 ? "" : "nd", v.size(), v.front().first, v[v.size() / 10].first, v[v.size() / 2].first, v[v.size() * 9 / 10].first,
               v.back().first, sum);
        printf("  slowest 24 (smid:GB/s):");
        for (int i = 0; i < 24 && i < (int)v.size(); i++) printf(" %u:%.0f", v[i].second, v[i].first);
        printf("\n  fastest 12:");
        for (int i = 0; i < 12 && i < (int)v.size(); i++) printf(" %u:%.0f", v[v.size() - 1 - i].second, v[v.size() - 1 - i].first);
        printf("\n");
    }
    return 0;
}
