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
    uint8_t *obs;           // [P][N][W][H][C] elements of obs_elem bytes
    int obs_elem;           // 1 = int8 (the reference's export), 2 = bfloat16, 4 = float (warp-tile kernels only)
    int32_t *ep_return;     // [N] or nullptr
    unsigned long long *stats;   // [2] or nullptr
    const uint8_t *tile_const[4];   // [Layout][template tile] blocks of the warp-tile kernel for 1 / 2 / 4 / 8 lanes per world (device; null = not usable)
    int lanes;                   // lanes per world of the warp-tile kernel (1, 2, 4 or 8)
    unsigned long long *trace;   // debug timeline buffer of the next launch (nullptr = off)
    uint32_t N;
    int force_kernel;       // -1 = auto, 0 = CTA-per-world, 1 = warp-tile (tests)
};

inline int lanes_index(int lanes) { return lanes == 1 ? 0 : lanes == 2 ? 1 : lanes == 4 ? 2 : 3; }
// obs_elem = bytes per observation element (1, 2 or 4)
int pick_kernel(const Layout &L, uint32_t N, int obs_elem);   // 1 = warp-tile, 0 = CTA-per-world
int pick_lanes(const Layout &L, uint32_t N, int obs_elem);
bool tile_fits(const Layout &L, int lanes, int obs_elem);   // the warp-tile kernel's shared memory fits on an SM
bool world_cta_fits(const Layout &L, int obs_elem);         // ... the CTA-per-world kernel's (one slab of S x C elements at least)
size_t tile_const_bytes(const Layout &L, int lanes, int obs_elem);
// The constant block is fetched by every CTA of every step: TILE_CONST_COPIES copies, tile_const_stride() bytes apart,
// spread those reads over the L2 instead of hammering the few lines of one copy (CTA b reads copy b % TILE_CONST_COPIES)
constexpr int TILE_CONST_COPIES = 64;
size_t tile_const_stride(const Layout &L, int lanes, int obs_elem);
void build_tile_const(const Layout &L, int lanes, int obs_elem, uint8_t *dst);   // host buffer of tile_const_bytes()
// fused T-step rollout (one launch); only for layouts the warp-tile kernel runs
bool rollout_available(const Layout &hL, const LaunchArgs &a);
// actions == nullptr: uniformly random moves drawn in the kernel (the random policy's hash of seed, first_step + t, flat
// action index); the last step's moves are left in a.actions
cudaError_t launch_rollout(const Layout &hL, const LaunchArgs &a, int T, const int32_t *actions, int action_slices,
                           uint8_t *obs, int32_t *rewards, int32_t *done, int out_slices, cudaStream_t st,
                           uint32_t seed = 0, uint32_t first_step = 0);
cudaError_t launch_step(const Layout &hL, const Layout *dL, const LaunchArgs &a, int init_only, cudaStream_t st);
// compat: native obs [P][N][W][H][C] (int8) -> the reference's padded rows [N][S][P][336] (world-major memory)
constexpr int RAW_ROW_BYTES = 5 * 64 + 16;   // sizeof(LocationXObservation), src/overcooked_env/sim.hpp:125-127
cudaError_t launch_export_raw(const Layout &hL, const LaunchArgs &a, uint8_t *dst, cudaStream_t st);

}   // namespace oc
}   // namespace b200
