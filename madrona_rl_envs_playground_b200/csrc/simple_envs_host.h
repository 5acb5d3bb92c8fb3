// Host-visible interface of simple_envs.cu (Cartpole, Acrobat, Balance Beam).
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace b200 {

inline uint32_t num_tiles(uint32_t N) { return (N + 255) / 256; }
// sets the calling thread's b200env_last_error() message and returns `code` (defined in b200envs.cu)
int report_error(int code, const char *msg);
size_t simple_roll_ring_words(uint32_t N);   // words of the fused rollouts' reset-count ring for N worlds

namespace cartpole {
struct LaunchArgs {
    float *state;             // [N][4]
    const int32_t *actions;   // [N]
    float *reward;            // [N]
    int32_t *done;            // [N]
    unsigned long long *scan_status;
    uint32_t *scan_ticket, *scan_epoch, *cur_episode;
    uint32_t *roll_ring, *roll_step;   // fused random rollout: per-step reset counts (roll_ring_words), step counter
    uint32_t N;
    // b200env_random_rollout: the step kernel draws this step's uniformly random action itself (policy::sample_kernel's
    // draw: move = floor(mix(seed, step, flat action index) * moves / 2^32)) and stores it in the action tensor
    int sample;
    uint32_t sample_seed, sample_step;
};
cudaError_t launch_step(const LaunchArgs &a, int init_only, cudaStream_t stream);
// num_steps x {draw a uniformly random action -> step} in one cooperative launch (the world in registers);
// cudaErrorCooperativeLaunchTooLarge = the tiles cannot all be resident: launch step by step
cudaError_t launch_rollout(const LaunchArgs &a, int num_steps, uint32_t seed, uint32_t first_step, cudaStream_t stream);
}   // namespace cartpole

namespace acrobat {
struct LaunchArgs {
    float *state;             // [N][4] theta1, theta2, omega1, omega2
    const int32_t *actions;   // [N]
    float *reward;            // [N]
    int32_t *done;            // [N]
    unsigned long long *scan_status;
    uint32_t *scan_ticket, *scan_epoch, *cur_episode;
    uint32_t *term_status;    // [tiles] last termination of each tile in this launch (epoch-tagged)
    uint32_t *ep_len;         // the Manager-wide episode-length counter (EpisodeManager::episodeLength)
    uint32_t N;
    // b200env_random_rollout: the step kernel draws this step's uniformly random action itself (policy::sample_kernel's
    // draw: move = floor(mix(seed, step, flat action index) * moves / 2^32)) and stores it in the action tensor
    int sample;
    uint32_t sample_seed, sample_step;
};
cudaError_t launch_step(const LaunchArgs &a, int init_only, cudaStream_t stream);
}   // namespace acrobat

namespace balance {
struct LaunchArgs {
    int32_t *obs;             // [2][N][7]
    const int32_t *actions;   // [2][N]
    float *reward;            // [2][N]
    int32_t *done;            // [N]
    int32_t *loc;             // [2][N]
    int32_t *time;            // [N]
    unsigned long long *scan_status;
    uint32_t *scan_ticket, *scan_epoch, *cur_episode;
    uint32_t *roll_ring, *roll_step;   // fused random rollout: per-step reset counts (roll_ring_words), step counter
    uint32_t N;
    // b200env_random_rollout: the step kernel draws this step's uniformly random action itself (policy::sample_kernel's
    // draw: move = floor(mix(seed, step, flat action index) * moves / 2^32)) and stores it in the action tensor
    int sample;
    uint32_t sample_seed, sample_step;
};
cudaError_t launch_step(const LaunchArgs &a, int init_only, cudaStream_t stream);
cudaError_t launch_rollout(const LaunchArgs &a, int num_steps, uint32_t seed, uint32_t first_step, cudaStream_t stream);
}   // namespace balance

}   // namespace b200
