// Per-SM store rates, measured once per device: see sm_rates_host.h.
#include <algorithm>
#include <cmath>

#include "common.cuh"
#include "sm_rates_host.h"

namespace b200 {

// One CTA per SM (14 warps, the population of the Hanabi writing warps) streams 16-byte stores into its own slice of
// `scratch`, all SMs at once; ns[%smid] = the time it took.
__global__ void __launch_bounds__(448) sm_rate_kernel(uint4 *scratch, size_t chunks_per_cta, int passes, unsigned long long *ns)
{
    unsigned smid;
    asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
    uint4 *mine = scratch + (size_t)blockIdx.x * chunks_per_cta;
    unsigned long long t0, t1;
    __syncthreads();
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
    const uint4 v = make_uint4(0u, 0u, 0u, 0u);
    for (int it = 0; it < passes; it++)
        for (size_t c = threadIdx.x; c < chunks_per_cta; c += blockDim.x) __stcs(mine + c, v);
    __syncthreads();
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
    if (threadIdx.x == 0 && smid < 256u) ns[smid] = t1 - t0;
}

struct SmRates {
    bool tried = false, ok = false;
    float rate[256] = {};
};

const float *sm_store_rates(void *scratch, size_t scratch_bytes, cudaError_t *err)
{
    static SmRates table[64];
    *err = cudaSuccess;
    int dev = 0;
    cudaGetDevice(&dev);
    const int sms = device_info().sms;
    SmRates &sr = table[dev >= 0 && dev < 64 ? dev : 0];
    std::lock_guard<std::mutex> lock(host_cache_mutex());
    if (sr.tried) return sr.ok ? sr.rate : nullptr;
    const size_t chunks = scratch_bytes / 16 / (size_t)std::max(sms, 1);
    if (sms > 256 || chunks < (128u << 10) / 16 || getenv("B200ENV_NO_SM_BALANCE")) return nullptr;   // (may be tried again)
    sr.tried = true;
    unsigned long long *ns = nullptr, host_ns[256];
    if ((*err = cudaMalloc(&ns, sizeof(host_ns))) != cudaSuccess) return nullptr;
    cudaMemset(ns, 0, sizeof(host_ns));
    const int passes = (int)std::max<size_t>(2, (24u << 20) / (chunks * 16));   // ~24 MB per SM
    for (int rep = 0; rep < 2; rep++)   // (the second run is the measurement: clocks and caches warm)
        sm_rate_kernel<<<sms, 448>>>(reinterpret_cast<uint4 *>(scratch), chunks, passes, ns);
    *err = cudaMemcpy(host_ns, ns, sizeof(host_ns), cudaMemcpyDeviceToHost);
    cudaFree(ns);
    cudaMemset(scratch, 0, scratch_bytes);
    if (*err != cudaSuccess) return nullptr;
    int seen = 0;
    double sum = 0;
    for (int i = 0; i < 256; i++)
        if (host_ns[i]) { sr.rate[i] = 1e6f / (float)host_ns[i]; sum += sr.rate[i]; seen++; }
    sr.ok = seen == sms;   // one CTA per SM, ids below 256
    if (sr.ok) {
        const float mean = (float)(sum / seen);
        for (int i = 0; i < 256; i++) {
            sr.rate[i] /= mean;
            if (sr.rate[i] != 0.f && (sr.rate[i] < 0.5f || sr.rate[i] > 2.f)) sr.ok = false;   // not a rate
        }
    }
    return sr.ok ? sr.rate : nullptr;
}

bool sm_quotas(const float *rates, uint32_t tiles, uint32_t cap, double exponent, uint32_t quota[256])
{
    for (int i = 0; i < 256; i++) quota[i] = 0;
    if (!rates) return false;
    double w[256], W = 0;
    for (int i = 0; i < 256; i++) { w[i] = rates[i] > 0.f ? pow((double)rates[i], exponent) : 0.0; W += w[i]; }
    if (W <= 0) return false;
    double acc = 0;
    uint32_t given = 0;
    for (int i = 0; i < 256; i++) {
        if (w[i] == 0.0) continue;
        acc += w[i] / W * tiles;
        uint32_t upto = (uint32_t)(acc + 0.5);
        if (upto > tiles) upto = tiles;
        const uint32_t q = upto - given;
        given = upto;
        quota[i] = q > cap ? cap : q;
    }
    return true;
}

}   // namespace b200
