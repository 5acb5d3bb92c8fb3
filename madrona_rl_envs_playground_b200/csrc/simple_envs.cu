// Cartpole and Balance Beam step kernels (sm_100a): one thread per world, 256-world CTA tiles, fully coalesced SoA-free
// access (the exported tensors ARE the state), in-step reset with reference-ordered episode numbering (common.cuh).
// Both kernels are launch-latency bound at the BASELINE sizes (49 KB and 2.7 MB of traffic per step).
#include "common.cuh"
#include "simple_envs_host.h"

namespace b200 {

constexpr int TILE = 256;

// block-wide exclusive rank of `flag` among the tile's threads + tile total
__device__ __forceinline__ void tile_rank(int flag, uint32_t *s_warp, uint32_t &rank, uint32_t &total)
{
    const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const unsigned m = __ballot_sync(0xffffffffu, flag);
    if (lane == 0) s_warp[warp] = __popc(m);
    __syncthreads();
    uint32_t before = 0, all = 0;
#pragma unroll
    for (int w = 0; w < TILE / 32; w++) {
        const uint32_t c = s_warp[w];
        before += (w < (int)warp) ? c : 0;
        all += c;
    }
    rank = before + __popc(m & ((1u << lane) - 1u));
    total = all;
    __syncthreads();
}

// ------------------------------------------------------------------------------------------------------------
// Cartpole — src/cartpole_env/sim.cpp:9-145.  The reference's constants are double literals: every expression is
// evaluated in double and rounded to float on assignment (sim.cpp:68-84); reproduced literally, compiled with
// --fmad=false.  sinf/cosf differ from glibc by a few ulp: parity tolerance 1e-5 relative (tests/).
// ------------------------------------------------------------------------------------------------------------
namespace cartpole {

#define CP_GRAVITY 9.8
#define CP_MASSCART 1.0
#define CP_MASSPOLE 0.1
#define CP_TOTAL_MASS (CP_MASSPOLE + CP_MASSCART)
#define CP_LENGTH 0.5
#define CP_POLEMASS_LENGTH (CP_MASSPOLE * CP_LENGTH)
#define CP_FORCE_MAG 10
#define CP_TAU 0.02
#define CP_X_THRESHOLD 2.4
#define CP_PI 3.141592653589793238463
#define CP_THETA_THRESHOLD_RADIANS (12 * 2 * CP_PI / 360)

__global__ void __launch_bounds__(TILE) step_kernel(const LaunchArgs a, const EpisodeScan scan, const uint32_t num_tiles,
                                                    const int init_only)
{
    __shared__ uint32_t s_scan[4];
    __shared__ uint32_t s_warp[TILE / 32];
    EpisodeScanTile st;
    st.begin(scan, s_scan);
    const uint32_t n = st.tile * TILE + threadIdx.x;
    const bool live = n < a.N;

    float4 s = make_float4(0.f, 0.f, 0.f, 0.f);   // x, x_dot, theta, theta_dot
    int done = 0;
    if (live) {
        if (init_only) {
            done = 1;
        } else {
            s = reinterpret_cast<const float4 *>(a.state)[n];
            const int action = a.actions[n];
            // actionSystem, sim.cpp:68-84
            const float force = (action == 1 ? CP_FORCE_MAG : -CP_FORCE_MAG);
            const float costheta = cosf(s.z);
            const float sintheta = sinf(s.z);
            const float temp = (force + CP_POLEMASS_LENGTH * s.w * s.w * sintheta) / CP_TOTAL_MASS;
            const float thetaacc = (CP_GRAVITY * sintheta - costheta * temp) /
                                   (CP_LENGTH * (4.0 / 3.0 - CP_MASSPOLE * costheta * costheta / CP_TOTAL_MASS));
            const float xacc = temp - CP_POLEMASS_LENGTH * thetaacc * costheta / CP_TOTAL_MASS;
            s.x = s.x + CP_TAU * s.y;
            s.y = s.y + CP_TAU * xacc;
            s.z = s.z + CP_TAU * s.w;
            s.w = s.w + CP_TAU * thetaacc;
            // checkDone, sim.cpp:86-96
            done = s.x < -CP_X_THRESHOLD || s.x > CP_X_THRESHOLD || s.z < -CP_THETA_THRESHOLD_RADIANS ||
                   s.z > CP_THETA_THRESHOLD_RADIANS;
        }
    }
    uint32_t rank, total;
    tile_rank(done, s_warp, rank, total);
    st.publish(scan, total, num_tiles);
    const uint32_t episode0 = st.exclusive_base(scan, total, num_tiles, s_scan + 3);
    if (live) {
        if (done) {   // resetWorld, sim.cpp:48-66: four draws in the order x, x_dot, theta, theta_dot
            uint32_t rng = rng_make(episode0 + rank);
            const float lo = -0.05f, diff = 0.05f - (-0.05f);
            s.x = lo + rng_rand(rng) * diff;
            s.y = lo + rng_rand(rng) * diff;
            s.z = lo + rng_rand(rng) * diff;
            s.w = lo + rng_rand(rng) * diff;
        }
        reinterpret_cast<float4 *>(a.state)[n] = s;
        a.reward[n] = init_only ? 0.f : 1.f;
        a.done[n] = init_only ? 0 : done;
    }
}

cudaError_t launch_step(const LaunchArgs &a, int init_only, cudaStream_t stream)
{
    EpisodeScan scan{a.scan_status, a.scan_ticket, a.scan_epoch, a.cur_episode};
    const uint32_t tiles = num_tiles(a.N);
    step_kernel<<<tiles, TILE, 0, stream>>>(a, scan, tiles, init_only);
    return cudaGetLastError();
}

}   // namespace cartpole

// ------------------------------------------------------------------------------------------------------------
// Balance Beam — src/balance_beam_env/sim.cpp:46-205.  Tensors: obs i32 [2][N][7], reward f32 [2][N], done i32 [N],
// actions i32 [2][N]; hidden state: loc i32 [2][N], time i32 [N].
// ------------------------------------------------------------------------------------------------------------
namespace balance {

constexpr int BUFFER = 2, NUM_SPACES = 5, TIME = 3;
#define BB_SCALE 0.2

__global__ void __launch_bounds__(TILE) step_kernel(const LaunchArgs a, const EpisodeScan scan, const uint32_t num_tiles,
                                                    const int init_only)
{
    __shared__ uint32_t s_scan[4];
    __shared__ uint32_t s_warp[TILE / 32];
    EpisodeScanTile st;
    st.begin(scan, s_scan);
    const uint32_t n = st.tile * TILE + threadIdx.x;
    const bool live = n < a.N;
    const size_t N = a.N;

    int loc[2] = {0, 0}, time = 0, done = 0;
    int obs[2][2 * TIME + 1];
    float reward = 0.f;
    if (live) {
        if (init_only) {
            done = 1;
        } else {
#pragma unroll
            for (int i = 0; i < 2; i++) {
                loc[i] = a.loc[i * N + n];
#pragma unroll
                for (int k = 0; k < 7; k++) obs[i][k] = a.obs[(i * N + n) * 7 + k];
                const int act = a.actions[i * N + n];   // actionSystem, sim.cpp:83-99
                loc[i] += act == 0 ? -2 : act == 1 ? -1 : act == 2 ? 1 : act == 3 ? 2 : 0;
            }
            time = a.time[n] - 1;                        // timeSystem, sim.cpp:101-104
#pragma unroll
            for (int i = 0; i < 2; i++) {                // observationSystem, sim.cpp:106-118
#pragma unroll
                for (int k = TIME * 2; k > 0; k--) obs[i][k] = obs[i][k - 1];
                obs[i][TIME] = loc[1 - i] + BUFFER;
                obs[i][0] = loc[i] + BUFFER;
                obs[i][2 * TIME] = time;
            }
            // checkDone, sim.cpp:120-155 (double arithmetic rounded to float, as the reference)
            const int d = loc[0] - loc[1];
            reward = (float)(loc[0] == loc[1] ? 1.0 : -abs(d) * BB_SCALE);
#pragma unroll
            for (int i = 0; i < 2; i++) {
                if (loc[i] < 0 || loc[i] >= NUM_SPACES) {
                    done = 1;
                    reward = (float)(-NUM_SPACES * (time + 1) * BB_SCALE);
                }
            }
            if (time == 0) done = 1;
        }
    }
    uint32_t rank, total;
    tile_rank(done, s_warp, rank, total);
    st.publish(scan, total, num_tiles);
    const uint32_t episode0 = st.exclusive_base(scan, total, num_tiles, s_scan + 3);
    if (live) {
        if (done) {   // resetWorld, sim.cpp:46-81: two draws, agent 0 then agent 1
            uint32_t rng = rng_make(episode0 + rank);
            time = TIME - 1;
#pragma unroll
            for (int i = 0; i < 2; i++) loc[i] = (int)((float)NUM_SPACES * rng_rand(rng));
#pragma unroll
            for (int i = 0; i < 2; i++) {
#pragma unroll
                for (int k = 0; k < 2 * TIME; k++) obs[i][k] = 0;
                obs[i][2 * TIME] = time;
                obs[i][0] = loc[i] + BUFFER;
                obs[i][TIME] = loc[1 - i] + BUFFER;
            }
        }
#pragma unroll
        for (int i = 0; i < 2; i++) {
            a.loc[i * N + n] = loc[i];
#pragma unroll
            for (int k = 0; k < 7; k++) a.obs[(i * N + n) * 7 + k] = obs[i][k];
            a.reward[i * N + n] = reward;
        }
        a.time[n] = time;
        a.done[n] = init_only ? 0 : done;
    }
}

cudaError_t launch_step(const LaunchArgs &a, int init_only, cudaStream_t stream)
{
    EpisodeScan scan{a.scan_status, a.scan_ticket, a.scan_epoch, a.cur_episode};
    const uint32_t tiles = num_tiles(a.N);
    step_kernel<<<tiles, TILE, 0, stream>>>(a, scan, tiles, init_only);
    return cudaGetLastError();
}

}   // namespace balance
}   // namespace b200
