// Host-visible interface of hanabi.cu.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace b200 {
namespace hanabi {

constexpr int GAME_WORDS = 25;   // packed per-world game record in 32-bit words: 13 of deck, 12 of register record (hanabi_game.cuh)

struct LaunchArgs {
    uint32_t *game;            // [GAME_WORDS][N]
    const int32_t *actions;    // [2][N]
    uint8_t *obs, *state;      // [2][N][658], [2][N][783] elements of obs_elem bytes
    int obs_elem;              // 1 = int8 (the reference's export), 2 = bfloat16, 4 = float (policy-ready rows)
    int32_t *mask, *active, *done;   // [2][N][20], [2][N], [N]
    float *reward;             // [2][N]
    unsigned long long *scan_status;   // [ceil(N/32)]
    uint32_t *scan_ticket, *scan_epoch, *cur_episode;
    uint32_t *roll_ring, *roll_step;   // fused random rollout: per-step reset counts (rollout_ring_words), step counter
    uint32_t *roll_quota, *roll_ctl;   // ... tiles per SM id [256], claim counters + grid barrier [8]
    int roll_balanced;                 // roll_quota holds measured shares (prepare_rollout_balance); 0 = equal shares by CTA index
    unsigned long long *trace;   // debug timeline buffer of the next launch (nullptr = off), 8 stamps per tile
    int trace_steps;             // fused rollout: steps to trace, 16 stamps per (step, tile)
    // b200env_random_rollout: the step kernel also draws the next step's random legal actions (0 = off)
    int sample_next;
    uint32_t sample_seed, sample_step;
    uint32_t N;
    int colors, ranks, max_info, max_life;
};

inline uint32_t num_tiles(uint32_t N) { return (N + 31) / 32; }
cudaError_t launch_step(const LaunchArgs &a, int init_only, cudaStream_t stream);
size_t rollout_ring_words(uint32_t N);
// fills a.roll_quota (tiles per SM in proportion to the SM's measured store rate; measured once per device, with
// `scratch` — device memory the caller overwrites afterwards — as the target); synchronises the device
cudaError_t prepare_rollout_balance(LaunchArgs &a, void *scratch, size_t scratch_bytes);
// num_steps x {draw a uniformly random legal move -> step} in one cooperative launch; the action tensor must hold the
// moves of the first step.  cudaErrorCooperativeLaunchTooLarge = the tiles cannot all be resident: launch step by step.
cudaError_t launch_rollout(const LaunchArgs &a, int num_steps, uint32_t seed, uint32_t first_step, cudaStream_t stream);

}   // namespace hanabi
}   // namespace b200
