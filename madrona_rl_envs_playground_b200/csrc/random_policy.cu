// Uniform random LEGAL action sampler (sm_100a) — the "random agent" of the reference's example scripts
// (scripts/overcooked_example.py:99-105 torch.randint; scripts/hanabi_example.py:64-67 masked argmax of random logits),
// as one small kernel that writes straight into the simulator's action tensor, so that a random-action rollout is a
// device-only loop {sample -> step} with no host round trip.
//
// Counter-based: action = f(seed, step, agent row, world) through a 2-round multiply-xorshift hash, so any (step, world)
// can be regenerated independently — tests replay the same stream on the host.
#include "random_policy_host.h"

namespace b200 {
namespace policy {

// one thread per (agent row, world): rows = P (or 1), mask i32 [rows][N][M] or nullptr (all M actions legal)
__global__ void __launch_bounds__(256) sample_kernel(int32_t *__restrict__ actions, const int32_t *__restrict__ mask,
                                                     const uint32_t total, const int M, const uint32_t seed,
                                                     const uint32_t step)
{
    asm volatile("griddepcontrol.wait;" ::: "memory");
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    const uint32_t i = blockIdx.x * 256u + threadIdx.x;
    if (i >= total) return;
    const uint32_t r = mix(seed, step, i);
    int choice;
    if (!mask) {
        choice = (int)(((uint64_t)r * (uint32_t)M) >> 32);
    } else {
        const int32_t *m = mask + (size_t)i * M;
        int legal = 0;
        for (int k = 0; k < M; k++) legal += (m[k] != 0);
        choice = 0;   // rows with no legal move (the inactive Hanabi player) get action 0; the step ignores them
        if (legal > 0) {
            int pick = (int)(((uint64_t)r * (uint32_t)legal) >> 32);
            for (int k = 0; k < M; k++) {
                if (m[k] != 0 && pick-- == 0) { choice = k; break; }
            }
        }
    }
    actions[i] = choice;
}

uint32_t host_mix(uint32_t seed, uint32_t step, uint32_t idx) { return mix(seed, step, idx); }

cudaError_t launch_sample(int32_t *actions, const int32_t *mask, uint32_t total, int M, uint32_t seed, uint32_t step,
                          cudaStream_t st)
{
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((total + 255) / 256);
    cfg.blockDim = dim3(256);
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    return cudaLaunchKernelEx(&cfg, sample_kernel, actions, mask, total, M, seed, step);
}

}   // namespace policy
}   // namespace b200
