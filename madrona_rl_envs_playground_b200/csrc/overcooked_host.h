// Host-visible interface of overcooked.cu (used by the C-ABI implementation in b200envs.cu).
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace b200 {
namespace oc {

struct Layout;

struct LaunchArgs {
    void *players;          // uint2 [P][N]
    uint32_t *objs;         // [K][N]
    int32_t *timestep;      // [N]
    const int32_t *actions; // [P][N]
    int32_t *rewards;       // [P][N]
    int32_t *done;          // [N]
    uint8_t *obs;           // [P][N][W][H][C]
    int32_t *ep_return;     // [N] or nullptr
    unsigned long long *stats;   // [2] or nullptr
    uint32_t N;
    int force_kernel;       // -1 = auto, 0 = CTA-per-world, 2/4/8 = warp-tile kernel with that PMAX (tests)
};

int pick_kernel(const Layout &L);
cudaError_t launch_step(const Layout &hL, const Layout *dL, const LaunchArgs &a, int init_only, cudaStream_t st);

}   // namespace oc
}   // namespace b200
