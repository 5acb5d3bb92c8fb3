// Host-visible interface of random_policy.cu.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace b200 {
namespace policy {

// counter-based hash: any (seed, step, index) can be regenerated independently (tests replay it on the host)
__host__ __device__ inline uint32_t mix(uint32_t seed, uint32_t step, uint32_t idx)
{
    uint32_t h = seed ^ (step * 0x9E3779B9u) ^ (idx * 0x85EBCA6Bu + 0xC2B2AE35u);
    h ^= h >> 16; h *= 0x7FEB352Du;
    h ^= h >> 15; h *= 0x846CA68Bu;
    h ^= h >> 16;
    return h;
}

// actions i32 [total]; mask i32 [total][M] or nullptr; action i = k-th legal move, k = floor(hash(seed, step, i) * legal / 2^32)
cudaError_t launch_sample(int32_t *actions, const int32_t *mask, uint32_t total, int M, uint32_t seed, uint32_t step,
                          cudaStream_t st);
uint32_t host_mix(uint32_t seed, uint32_t step, uint32_t idx);

}   // namespace policy
}   // namespace b200
