// Host-visible interface of random_policy.cu.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace b200 {
namespace policy {

// actions i32 [total]; mask i32 [total][M] or nullptr; action i = k-th legal move, k = floor(hash(seed, step, i) * legal / 2^32)
cudaError_t launch_sample(int32_t *actions, const int32_t *mask, uint32_t total, int M, uint32_t seed, uint32_t step,
                          cudaStream_t st);
uint32_t host_mix(uint32_t seed, uint32_t step, uint32_t idx);

}   // namespace policy
}   // namespace b200
