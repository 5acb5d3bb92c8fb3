// Cartpole, Acrobat and Balance Beam step kernels (sm_100a): one thread per world, 256-world CTA tiles, fully coalesced SoA-free
// access (the exported tensors ARE the state), in-step reset with reference-ordered episode numbering (common.cuh).
// Both kernels are launch-latency bound at the BASELINE sizes (49 KB and 2.7 MB of traffic per step).
#include "common.cuh"
#include "random_policy_host.h"
#include "simple_envs_host.h"

namespace b200 {

constexpr int TILE = 256;

// The action of flat element `idx` of the action tensor: the caller's, or — inside b200env_random_rollout — the draw
// policy::sample_kernel would make for (seed, step, idx), written back so that the action tensor reads the same.
template <class Args>
__device__ __forceinline__ int take_action(const Args &a, size_t idx, int moves)
{
    if (!a.sample) return a.actions[idx];
    const int move = (int)(((uint64_t)policy::mix(a.sample_seed, a.sample_step, (uint32_t)idx) * (uint32_t)moves) >> 32);
    const_cast<int32_t *>(a.actions)[idx] = move;
    return move;
}

// every step kernel here is launched with the programmatic-stream-serialization attribute: at ~3 us of work per launch
// the launch gap is what a random rollout pays for.  The kernels wait for the previous grid first thing (the episode
// scan's ticket counter is reset by the previous launch's last tile), so only launch and scheduling latency overlap.
template <class... KArgs, class... Args>
static cudaError_t launch_pdl(void (*kernel)(KArgs...), unsigned grid, unsigned block, cudaStream_t st, Args... args)
{
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(block);
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    return cudaLaunchKernelEx(&cfg, kernel, KArgs(args)...);
}
__device__ __forceinline__ void pdl_prologue()
{
    asm volatile("griddepcontrol.wait;" ::: "memory");
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
}

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
    pdl_prologue();
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
            const int action = take_action(a, n, 2);
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
    return launch_pdl(step_kernel, tiles, TILE, stream, a, scan, tiles, init_only);
}

}   // namespace cartpole

// ------------------------------------------------------------------------------------------------------------
// Acrobat — src/acrobat_env/sim.cpp:12-240: a two-link pendulum, one RK4 step of 0.2 s per env
// step.  Everything is float except the three sub-expressions the reference writes with double operands (`cos` of
// theta2 inside d2, `pi / 2.0` inside phi2, `dt / 2.0` and `dt / 6.0`); reproduced literally, --fmad=false.  Parity
// tolerance 1e-5 relative (sinf / cosf differ from glibc by a few ulp).
//
// The episode-length counter is ONE word shared by all worlds of a Manager (EpisodeManager::episodeLength,
// init.hpp:9): every world-step increments it (sim.cpp:175), any reset zeroes it (:57), a world that finds it above
// 500 is truncated (:210).  In the reference's deterministic order (one worker, worlds ascending, whole task graph per
// world) the counter after world w is therefore
//     0                          if w terminated,
//     (w - t) mod 501            if t < w is the last world that terminated earlier in this step,
//     (c_in + w + 1) mod 501     if none did (c_in = the counter when the step began),
// and w is truncated exactly when that value is 0.  So the only cross-world input is "the last termination before me":
// an exclusive max-scan over worlds — in the tile by ballots, across tiles by a second status array read in ticket
// order (one word per tile, tagged with the launch epoch like EpisodeScan's) — followed by the usual ordered episode
// numbering of the worlds that reset.
// ------------------------------------------------------------------------------------------------------------
namespace acrobat {

struct Vec5 { float v[5]; };
#define AC_PI 3.14159265358979323846264338327950288f   // madrona::math::pi (float), math.hpp:26

// ds_dt, sim.cpp:71-97 (powf(x, 2) == x * x rounded once)
__device__ __forceinline__ Vec5 ds_dt(const Vec5 &y)
{
    const float m1 = 1.0f, m2 = 1.0f, l1 = 1.0f, lc1 = 0.5f, lc2 = 0.5f, I1 = 1.0f, I2 = 1.0f, g = 9.8f;
    const float theta1 = y.v[0], theta2 = y.v[1], dtheta1 = y.v[2], dtheta2 = y.v[3], a = y.v[4];
    const float s2 = sinf(theta2);
    const float d1 = (m1 * (lc1 * lc1) + m2 * ((l1 * l1) + (lc2 * lc2) + 2 * l1 * lc2 * cosf(theta2)) + I1 + I2);
    const float d2 = m2 * ((lc2 * lc2) + l1 * lc2 * cos((double)theta2)) + I2;        // ::cos(double) in the reference
    const float phi2 = m2 * lc2 * g * cosf(theta1 + theta2 - AC_PI / 2.0);            // double subtraction, then cosf
    const float phi1 = (-m2 * l1 * lc2 * (dtheta2 * dtheta2) * s2 - 2 * m2 * l1 * lc2 * dtheta2 * dtheta1 * s2 +
                        (m1 * lc1 + m2 * l1) * g * cosf(theta1 - AC_PI / 2) + phi2);
    const float ddtheta2 = (a + d2 / d1 * phi1 - m2 * l1 * lc2 * (dtheta1 * dtheta1) * s2 - phi2) /
                           (m2 * (lc2 * lc2) + I2 - (d2 * d2) / d1);
    const float ddtheta1 = -(d2 * ddtheta2 + phi1) / d1;
    return Vec5{{dtheta1, dtheta2, ddtheta1, ddtheta2, 0.f}};
}
__device__ __forceinline__ Vec5 mul_const(Vec5 x, float c)
{
#pragma unroll
    for (int i = 0; i < 5; i++) x.v[i] *= c;
    return x;
}
__device__ __forceinline__ Vec5 add(Vec5 a, const Vec5 &b)
{
#pragma unroll
    for (int i = 0; i < 5; i++) a.v[i] += b.v[i];
    return a;
}
__device__ __forceinline__ float wrap(float x, float m, float M)
{
    const float diff = M - m;
    while (x > M) x -= diff;
    while (x < m) x += diff;
    return x;
}

constexpr uint32_t MAX_EPISODE_LENGTH = 500;   // sim.cpp:210

__global__ void __launch_bounds__(TILE) step_kernel(const LaunchArgs a, const EpisodeScan scan, const uint32_t num_tiles,
                                                    const int init_only)
{
    __shared__ uint32_t s_scan[4];
    __shared__ uint32_t s_warp[TILE / 32];
    __shared__ int s_wterm[TILE / 32];
    __shared__ uint32_t s_cin;
    __shared__ long long s_lower;
    pdl_prologue();
    EpisodeScanTile st;
    st.begin(scan, s_scan);
    if (threadIdx.x == 0) s_cin = *(volatile uint32_t *)a.ep_len;   // read before this tile publishes anything
    const uint32_t n = st.tile * TILE + threadIdx.x;
    const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const bool live = n < a.N;

    float4 s = make_float4(0.f, 0.f, 0.f, 0.f);   // theta1, theta2, omega1, omega2
    int term = 0;
    if (live && !init_only) {
        s = reinterpret_cast<const float4 *>(a.state)[n];
        const int action = take_action(a, n, 3);
        // actionSystem, sim.cpp:172-196: rk4 over {0, dt} (:117-147)
        const float torque = action == 0 ? -1.0f : action == 1 ? 0.0f : 1.0f;
        const Vec5 y0{{s.x, s.y, s.z, s.w, torque}};
        const float dt = 0.2f - 0.0f;
        const float dt2 = dt / 2.0;
        const float dt6 = dt / 6.0;
        const Vec5 k1 = ds_dt(y0);
        const Vec5 k2 = ds_dt(add(y0, mul_const(k1, dt2)));
        const Vec5 k3 = ds_dt(add(y0, mul_const(k2, dt2)));
        const Vec5 k4 = ds_dt(add(y0, mul_const(k3, dt)));
        const Vec5 ns = add(y0, mul_const(add(k1, add(mul_const(k2, 2), add(mul_const(k3, 2), k4))), dt6));
        s.x = wrap(ns.v[0], -AC_PI, AC_PI);
        s.y = wrap(ns.v[1], -AC_PI, AC_PI);
        s.z = fminf(fmaxf(ns.v[2], -(4 * AC_PI)), 4 * AC_PI);
        s.w = fminf(fmaxf(ns.v[3], -(9 * AC_PI)), 9 * AC_PI);
        // checkDone, sim.cpp:198-214: the free end above the target height
        term = -cosf(s.x) - cosf(s.y + s.x) > 1.0;
    }

    // ---- the last termination before each world: in the tile ... ----
    const unsigned tm = __ballot_sync(0xffffffffu, term);
    if (lane == 0) s_wterm[warp] = tm ? (int)(warp * 32 + 31 - __clz(tm)) : -1;
    __syncthreads();
    int pred = -1, tile_last = -1;
#pragma unroll
    for (int w = 0; w < TILE / 32; w++) {
        const int t = s_wterm[w];
        if (t >= 0) {
            tile_last = t;
            if (w < (int)warp) pred = t;
        }
    }
    const unsigned below = tm & ((1u << lane) - 1u);
    if (below) pred = (int)(warp * 32 + 31 - __clz(below));
    // ---- ... and in the lower tiles (ticket order = world order) ----
    const uint32_t tag = st.epoch & 0xFFFFu;
    if (threadIdx.x == 0)
        *(volatile uint32_t *)(a.term_status + st.tile) = (tag << 16) | (tile_last >= 0 ? 0x100u | (uint32_t)tile_last : 0u);
    if (warp == 0) {
        long long best = -1;
        for (uint32_t i0 = 4u * lane; i0 < st.tile; i0 += 128u) {
            uint32_t v[4];
            bool ok;
            do {
                asm volatile("ld.volatile.global.v4.u32 {%0, %1, %2, %3}, [%4];"
                             : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3])
                             : "l"(a.term_status + i0)
                             : "memory");
                ok = true;
#pragma unroll
                for (int c = 0; c < 4; c++) ok = ok && (i0 + c >= st.tile || (v[c] >> 16) == tag);
            } while (!ok);
#pragma unroll
            for (int c = 0; c < 4; c++)
                if (i0 + c < st.tile && (v[c] & 0x100u)) best = (long long)(i0 + c) * TILE + (long long)(v[c] & 0xFFu);
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) best = max(best, __shfl_xor_sync(0xffffffffu, best, o));
        if (lane == 0) s_lower = best;
    }
    __syncthreads();

    // ---- the shared counter after this world, truncation, reset ----
    uint32_t c_after = 0;
    if (!init_only && !term) {
        const uint32_t period = MAX_EPISODE_LENGTH + 1;
        if (pred >= 0) c_after = (uint32_t)((int)threadIdx.x - pred) % period;
        else if (s_lower >= 0) c_after = (uint32_t)(((long long)n - s_lower) % period);
        else c_after = (uint32_t)(((unsigned long long)s_cin + n + 1ull) % period);
    }
    const int done = live && (init_only || term || c_after == 0);
    if (n == a.N - 1) *(volatile uint32_t *)a.ep_len = c_after;   // every tile read the old value before it published

    uint32_t rank, total;
    tile_rank(done, s_warp, rank, total);
    st.publish(scan, total, num_tiles);
    const uint32_t episode0 = st.exclusive_base(scan, total, num_tiles, s_scan + 3);
    if (live) {
        if (done) {   // resetWorld, sim.cpp:46-69: four draws in the order theta1, theta2, omega1, omega2
            uint32_t rng = rng_make(episode0 + rank);
            const float lo = -0.1f, diff = 0.1f - (-0.1f);
            s.x = lo + rng_rand(rng) * diff;
            s.y = lo + rng_rand(rng) * diff;
            s.z = lo + rng_rand(rng) * diff;
            s.w = lo + rng_rand(rng) * diff;
        }
        reinterpret_cast<float4 *>(a.state)[n] = s;
        a.reward[n] = init_only ? 0.f : -1.f;
        a.done[n] = init_only ? 0 : done;
    }
}

cudaError_t launch_step(const LaunchArgs &a, int init_only, cudaStream_t stream)
{
    EpisodeScan scan{a.scan_status, a.scan_ticket, a.scan_epoch, a.cur_episode};
    const uint32_t tiles = num_tiles(a.N);
    return launch_pdl(step_kernel, tiles, TILE, stream, a, scan, tiles, init_only);
}

}   // namespace acrobat

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
    pdl_prologue();
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
                const int act = take_action(a, i * N + n, 4);   // actionSystem, sim.cpp:83-99
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
    return launch_pdl(step_kernel, tiles, TILE, stream, a, scan, tiles, init_only);
}

}   // namespace balance
}   // namespace b200
