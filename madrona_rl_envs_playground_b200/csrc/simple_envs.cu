// Cartpole, Acrobat and Balance Beam step kernels (sm_100a): one thread per world, 256-world CTA tiles, fully coalesced SoA-free
// access (the exported tensors ARE the state), in-step reset with reference-ordered episode numbering (common.cuh).
// Both kernels are launch-latency bound at the BASELINE sizes (49 KB and 2.7 MB of traffic per step).
#include <cstdio>
#include <cstdlib>
#include "common.cuh"
#include "random_policy_host.h"
#include "simple_envs_host.h"

namespace b200 {

constexpr int TILE = 256;

size_t simple_roll_ring_words(uint32_t N) { return roll_ring_words(num_tiles(N)); }

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

// Episode numbers inside a fused rollout of thread-per-world tiles (the ring protocol of common.cuh, block-wide):
// number(t, tile_count) publishes the tile's reset count of step t, learns the resets of all tiles in step t-1 and of
// the lower tiles in step t, and returns the episode index of the tile's first resetting world (all threads call).
struct RolloutEpisodes {
    uint32_t base, step_base;
    bool small;   // few tiles: one_group() below; many: the accumulator protocol of common.cuh (fewer words polled)
    __device__ __forceinline__ void begin(const RolloutScan &S, uint32_t num_tiles, uint32_t one_group_max)
    {
        small = num_tiles <= one_group_max;
        base = *(volatile uint32_t *)S.cur_episode;
        step_base = *(volatile uint32_t *)S.step_base;
    }
    // A few tiles (at most SCAN_GROUP): no accumulators at all.  The tile writes its status word and the
    // warp reads, in ONE polling loop of two independent 128-bit loads per lane, every tile's word of the previous
    // step (their sum = that step's resets; all present = nobody is more than a step behind) and the lower tiles'
    // words of this step — one store-to-load hop between tiles per step.
    __device__ __forceinline__ void one_group(const RolloutScan &S, uint32_t num_tiles, uint32_t tile, uint32_t gstep,
                                              bool want_prev, uint32_t now_end, uint32_t &prev, uint32_t &lower)
    {
        const unsigned lane = threadIdx.x;
        const uint32_t *now = roll_slot(S, num_tiles, gstep).status, *old = roll_slot(S, num_tiles, gstep - 1u).status;
        const uint32_t tag_now = (gstep + 1u) & 0xFFFFu, tag_old = gstep & 0xFFFFu, i0 = 4u * lane;
        const bool need_old = want_prev && i0 < num_tiles, need_now = i0 < now_end;
        uint32_t a[4] = {0, 0, 0, 0}, b[4] = {0, 0, 0, 0};
        bool ok;
        do {
            if (need_old)
                asm volatile("ld.volatile.global.v4.u32 {%0, %1, %2, %3}, [%4];"
                             : "=r"(a[0]), "=r"(a[1]), "=r"(a[2]), "=r"(a[3]) : "l"(old + i0) : "memory");
            if (need_now)
                asm volatile("ld.volatile.global.v4.u32 {%0, %1, %2, %3}, [%4];"
                             : "=r"(b[0]), "=r"(b[1]), "=r"(b[2]), "=r"(b[3]) : "l"(now + i0) : "memory");
            ok = true;
#pragma unroll
            for (int c = 0; c < 4; c++) {
                ok = ok && (!need_old || i0 + c >= num_tiles || (a[c] >> 16) == tag_old);
                ok = ok && (!need_now || i0 + c >= now_end || (b[c] >> 16) == tag_now);
            }
        } while (!ok);
        prev = lower = 0;
#pragma unroll
        for (int c = 0; c < 4; c++) {
            prev += (need_old && i0 + c < num_tiles) ? (a[c] & 0xFFFFu) : 0u;
            lower += (need_now && i0 + c < now_end) ? (b[c] & 0xFFFFu) : 0u;
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            prev += __shfl_xor_sync(0xffffffffu, prev, o);
            lower += __shfl_xor_sync(0xffffffffu, lower, o);
        }
    }
    __device__ __forceinline__ uint32_t number(const RolloutScan &S, uint32_t num_tiles, uint32_t tile, int t,
                                               uint32_t tile_count, uint32_t *smem2)
    {
        const uint32_t gstep = step_base + (uint32_t)t;
        if (threadIdx.x < 32) {
            uint32_t prev, lower;
            if (small) {
                if (threadIdx.x == 0)
                    *(volatile uint32_t *)(roll_slot(S, num_tiles, gstep).status + tile) = (((gstep + 1u) & 0xFFFFu) << 16) | tile_count;
                one_group(S, num_tiles, tile, gstep, t > 0, tile_count ? tile : 0u, prev, lower);
            } else {
                if (threadIdx.x == 0) roll_publish(S, num_tiles, tile, gstep, tile_count);
                __syncwarp();
                prev = t > 0 ? roll_step_total(S, num_tiles, gstep - 1u, threadIdx.x) : 0u;
                lower = tile_count ? roll_lower_tiles(S, num_tiles, tile, gstep, threadIdx.x) : 0u;
            }
            if (threadIdx.x == 0) { smem2[0] = prev; smem2[1] = lower; }
        }
        __syncthreads();
        base += smem2[0];
        const uint32_t first = base + smem2[1];
        __syncthreads();   // smem2 is rewritten in the next step
        return first;
    }
    // after the last step: tile 0 leaves the counters for the next launch (every tile read them in begin())
    __device__ __forceinline__ void finish(const RolloutScan &S, uint32_t num_tiles, uint32_t tile, int T)
    {
        if (tile == 0 && threadIdx.x < 32) {
            const uint32_t last = step_base + (uint32_t)T - 1u;
            uint32_t total, unused;
            if (small) one_group(S, num_tiles, tile, last + 1u, true, 0u, total, unused);
            else total = roll_step_total(S, num_tiles, last, threadIdx.x);
            if (threadIdx.x == 0) {
                *(volatile uint32_t *)S.cur_episode = base + total;
                *(volatile uint32_t *)S.step_base = step_base + (uint32_t)T;
            }
        }
    }
};

// Largest grid numbered with one_group(): measured on B200 (profiles/r01v_simple_rollouts.txt) Cartpole steps in 1.6 us
// on 4 tiles against 2.9 us with accumulators, 2.2 against 2.8 on 32 tiles, 2.6 against 2.5 on 64, 3.3 against 2.6 on
// 128 (every tile polls every tile's word).  B200ENV_ROLL_ONE_GROUP overrides (tuning).
static uint32_t one_group_max_tiles()
{
    static const uint32_t v = [] {
        const char *e = getenv("B200ENV_ROLL_ONE_GROUP");
        const long x = e ? strtol(e, nullptr, 10) : 32;
        return (uint32_t)(x < 0 ? 0 : x > (long)SCAN_GROUP ? (long)SCAN_GROUP : x);
    }();
    return v;
}

// cooperative launch: fails with cudaErrorCooperativeLaunchTooLarge instead of deadlocking when the tiles cannot all be
// resident (the caller then launches step by step)
template <class... KArgs, class... Args>
static cudaError_t launch_cooperative(void (*kernel)(KArgs...), unsigned grid, unsigned block, cudaStream_t st, Args... args)
{
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(block);
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeCooperative;
    attr[0].val.cooperative = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    return cudaLaunchKernelEx(&cfg, kernel, KArgs(args)...);
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

// actionSystem + checkDone, sim.cpp:68-96: one Euler step of (x, x_dot, theta, theta_dot); returns done
__device__ __forceinline__ int dynamics(float4 &s, int action)
{
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
    return s.x < -CP_X_THRESHOLD || s.x > CP_X_THRESHOLD || s.z < -CP_THETA_THRESHOLD_RADIANS ||
           s.z > CP_THETA_THRESHOLD_RADIANS;
}
// resetWorld, sim.cpp:48-66: four draws in the order x, x_dot, theta, theta_dot
__device__ __forceinline__ void fresh(float4 &s, uint32_t episode)
{
    uint32_t rng = rng_make(episode);
    const float lo = -0.05f, diff = 0.05f - (-0.05f);
    s.x = lo + rng_rand(rng) * diff;
    s.y = lo + rng_rand(rng) * diff;
    s.z = lo + rng_rand(rng) * diff;
    s.w = lo + rng_rand(rng) * diff;
}

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
            done = dynamics(s, take_action(a, n, 2));
        }
    }
    uint32_t rank, total;
    tile_rank(done, s_warp, rank, total);
    st.publish(scan, total, num_tiles);
    const uint32_t episode0 = st.exclusive_base(scan, total, num_tiles, s_scan + 3);
    if (live) {
        if (done) fresh(s, episode0 + rank);
        reinterpret_cast<float4 *>(a.state)[n] = s;
        a.reward[n] = init_only ? 0.f : 1.f;
        a.done[n] = init_only ? 0 : done;
    }
}

// Fused random rollout: T x {draw a random action, step} in ONE cooperative launch, the world in registers, the
// exported tensors rewritten every step.  See rollout_episodes below.
__global__ void __launch_bounds__(TILE) rollout_kernel(const LaunchArgs a, const RolloutScan S, const uint32_t seed,
                                                       const uint32_t first_step, const uint32_t num_tiles, const int T,
                                                       const uint32_t one_group_max)
{
    __shared__ uint32_t s_warp[TILE / 32];
    __shared__ uint32_t s_eps[2];
    const uint32_t tile = blockIdx.x, n = tile * TILE + threadIdx.x;
    const bool live = n < a.N;
    RolloutEpisodes ep;
    ep.begin(S, num_tiles, one_group_max);
    float4 s = live ? reinterpret_cast<const float4 *>(a.state)[n] : make_float4(0.f, 0.f, 0.f, 0.f);
    for (int t = 0; t < T; t++) {
        int done = 0, action = 0;
        if (live) {
            action = (int)(((uint64_t)policy::mix(seed, first_step + (uint32_t)t, n) * 2u) >> 32);
            done = dynamics(s, action);
        }
        uint32_t rank, total;
        tile_rank(done, s_warp, rank, total);
        const uint32_t episode0 = ep.number(S, num_tiles, tile, t, total, s_eps);
        if (live) {
            if (done) fresh(s, episode0 + rank);
            reinterpret_cast<float4 *>(a.state)[n] = s;
            const_cast<int32_t *>(a.actions)[n] = action;
            a.reward[n] = 1.f;
            a.done[n] = done;
        }
    }
    ep.finish(S, num_tiles, tile, T);
}

cudaError_t launch_rollout(const LaunchArgs &a, int num_steps, uint32_t seed, uint32_t first_step, cudaStream_t stream)
{
    if (num_steps <= 0) return cudaSuccess;
    const uint32_t tiles = num_tiles(a.N);
    return launch_cooperative(rollout_kernel, tiles, TILE, stream, a, RolloutScan{a.roll_ring, a.roll_step, a.cur_episode},
                              seed, first_step, tiles, num_steps, one_group_max_tiles());
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

struct World {   // one world in registers: positions, time left, both players' observation rows
    int loc[2], time;
    int obs[2][2 * TIME + 1];
};

// actionSystem, timeSystem, observationSystem, checkDone (sim.cpp:83-155); returns done
__device__ __forceinline__ int dynamics(World &w, int act0, int act1, float &reward)
{
    const int act[2] = {act0, act1};
#pragma unroll
    for (int i = 0; i < 2; i++) w.loc[i] += act[i] == 0 ? -2 : act[i] == 1 ? -1 : act[i] == 2 ? 1 : act[i] == 3 ? 2 : 0;
    w.time -= 1;                                     // timeSystem, sim.cpp:101-104
#pragma unroll
    for (int i = 0; i < 2; i++) {                    // observationSystem, sim.cpp:106-118
#pragma unroll
        for (int k = TIME * 2; k > 0; k--) w.obs[i][k] = w.obs[i][k - 1];
        w.obs[i][TIME] = w.loc[1 - i] + BUFFER;
        w.obs[i][0] = w.loc[i] + BUFFER;
        w.obs[i][2 * TIME] = w.time;
    }
    // checkDone, sim.cpp:120-155 (double arithmetic rounded to float, as the reference)
    int done = 0;
    const int d = w.loc[0] - w.loc[1];
    reward = (float)(w.loc[0] == w.loc[1] ? 1.0 : -abs(d) * BB_SCALE);
#pragma unroll
    for (int i = 0; i < 2; i++) {
        if (w.loc[i] < 0 || w.loc[i] >= NUM_SPACES) {
            done = 1;
            reward = (float)(-NUM_SPACES * (w.time + 1) * BB_SCALE);
        }
    }
    if (w.time == 0) done = 1;
    return done;
}
// resetWorld, sim.cpp:46-81: two draws, agent 0 then agent 1
__device__ __forceinline__ void fresh(World &w, uint32_t episode)
{
    uint32_t rng = rng_make(episode);
    w.time = TIME - 1;
#pragma unroll
    for (int i = 0; i < 2; i++) w.loc[i] = (int)((float)NUM_SPACES * rng_rand(rng));
#pragma unroll
    for (int i = 0; i < 2; i++) {
#pragma unroll
        for (int k = 0; k < 2 * TIME; k++) w.obs[i][k] = 0;
        w.obs[i][2 * TIME] = w.time;
        w.obs[i][0] = w.loc[i] + BUFFER;
        w.obs[i][TIME] = w.loc[1 - i] + BUFFER;
    }
}
// The observation rows of a tile are 7 words per world: a thread reading or writing its own row strides 28 B (a warp
// touches 28 sectors per instruction).  They go through shared memory instead (s_obs: 2 x TILE x 7 words, a stride of 7
// words is bank-conflict free) so that every global access is a fully coalesced 128 B per warp.  ALL threads call.
constexpr int OBS_WORDS = 2 * TIME + 1;
__device__ __forceinline__ void load_world(const LaunchArgs &a, size_t N, uint32_t tile0, bool live, World &w, int32_t *s_obs)
{
    const uint32_t n = tile0 + threadIdx.x;
    const uint32_t words = min((uint32_t)TILE, a.N - tile0) * OBS_WORDS;
#pragma unroll
    for (int i = 0; i < 2; i++) {
        const int32_t *src = a.obs + (i * N + tile0) * OBS_WORDS;
#pragma unroll
        for (int j = 0; j < OBS_WORDS; j++)
            if (j * TILE + threadIdx.x < words) s_obs[i * TILE * OBS_WORDS + j * TILE + threadIdx.x] = src[j * TILE + threadIdx.x];
    }
    __syncthreads();
    if (live) {
#pragma unroll
        for (int i = 0; i < 2; i++) {
            w.loc[i] = a.loc[i * N + n];
#pragma unroll
            for (int k = 0; k < OBS_WORDS; k++) w.obs[i][k] = s_obs[(i * TILE + threadIdx.x) * OBS_WORDS + k];
        }
        w.time = a.time[n];
    }
    // (s_obs is rewritten by store_exported, behind the block barriers of tile_rank)
}
__device__ __forceinline__ void store_exported(const LaunchArgs &a, size_t N, uint32_t tile0, bool live, const World &w,
                                               float reward, int done, int32_t *s_obs)
{
    const uint32_t n = tile0 + threadIdx.x;
    const uint32_t words = min((uint32_t)TILE, a.N - tile0) * OBS_WORDS;
    if (live) {
#pragma unroll
        for (int i = 0; i < 2; i++) {
#pragma unroll
            for (int k = 0; k < OBS_WORDS; k++) s_obs[(i * TILE + threadIdx.x) * OBS_WORDS + k] = w.obs[i][k];
            a.reward[i * N + n] = reward;
        }
        a.done[n] = done;
    }
    __syncthreads();
#pragma unroll
    for (int i = 0; i < 2; i++) {
        int32_t *dst = a.obs + (i * N + tile0) * OBS_WORDS;
#pragma unroll
        for (int j = 0; j < OBS_WORDS; j++)
            if (j * TILE + threadIdx.x < words) dst[j * TILE + threadIdx.x] = s_obs[i * TILE * OBS_WORDS + j * TILE + threadIdx.x];
    }
    // (a fused rollout stages the next step's rows in s_obs behind the block barriers of that step's tile_rank)
}
__device__ __forceinline__ void store_hidden(const LaunchArgs &a, size_t N, uint32_t n, const World &w)
{
    a.loc[n] = w.loc[0];
    a.loc[N + n] = w.loc[1];
    a.time[n] = w.time;
}

__global__ void __launch_bounds__(TILE) step_kernel(const LaunchArgs a, const EpisodeScan scan, const uint32_t num_tiles,
                                                    const int init_only)
{
    __shared__ uint32_t s_scan[4];
    __shared__ uint32_t s_warp[TILE / 32];
    __shared__ int32_t s_obs[2 * TILE * OBS_WORDS];
    pdl_prologue();
    EpisodeScanTile st;
    st.begin(scan, s_scan);
    const uint32_t n = st.tile * TILE + threadIdx.x;
    const bool live = n < a.N;
    const size_t N = a.N;

    const uint32_t tile0 = st.tile * TILE;
    World w = {};
    int done = 0;
    float reward = 0.f;
    if (!init_only) load_world(a, N, tile0, live, w, s_obs);
    if (live) {
        if (init_only) {
            done = 1;
        } else {
            const int act0 = take_action(a, n, 4), act1 = take_action(a, N + n, 4);
            done = dynamics(w, act0, act1, reward);
        }
    }
    uint32_t rank, total;
    tile_rank(done, s_warp, rank, total);
    st.publish(scan, total, num_tiles);
    const uint32_t episode0 = st.exclusive_base(scan, total, num_tiles, s_scan + 3);
    if (live) {
        if (done) fresh(w, episode0 + rank);
        store_hidden(a, N, n, w);
    }
    store_exported(a, N, tile0, live, w, reward, init_only ? 0 : done, s_obs);
}

// Fused random rollout: T x {draw both players' random actions, step} in ONE cooperative launch, the world in
// registers, the exported tensors rewritten every step.
__global__ void __launch_bounds__(TILE) rollout_kernel(const LaunchArgs a, const RolloutScan S, const uint32_t seed,
                                                       const uint32_t first_step, const uint32_t num_tiles, const int T,
                                                       const uint32_t one_group_max)
{
    __shared__ uint32_t s_warp[TILE / 32];
    __shared__ uint32_t s_eps[2];
    __shared__ int32_t s_obs[2 * TILE * OBS_WORDS];
    const uint32_t tile = blockIdx.x, tile0 = tile * TILE, n = tile0 + threadIdx.x;
    const bool live = n < a.N;
    const size_t N = a.N;
    RolloutEpisodes ep;
    ep.begin(S, num_tiles, one_group_max);
    World w = {};
    load_world(a, N, tile0, live, w, s_obs);
    for (int t = 0; t < T; t++) {
        int done = 0, act0 = 0, act1 = 0;
        float reward = 0.f;
        if (live) {
            act0 = (int)(((uint64_t)policy::mix(seed, first_step + (uint32_t)t, n) * 4u) >> 32);
            act1 = (int)(((uint64_t)policy::mix(seed, first_step + (uint32_t)t, (uint32_t)N + n) * 4u) >> 32);
            done = dynamics(w, act0, act1, reward);
        }
        uint32_t rank, total;
        tile_rank(done, s_warp, rank, total);
        const uint32_t episode0 = ep.number(S, num_tiles, tile, t, total, s_eps);
        if (live) {
            if (done) fresh(w, episode0 + rank);
            const_cast<int32_t *>(a.actions)[n] = act0;
            const_cast<int32_t *>(a.actions)[N + n] = act1;
        }
        store_exported(a, N, tile0, live, w, reward, done, s_obs);
    }
    if (live) store_hidden(a, N, n, w);
    ep.finish(S, num_tiles, tile, T);
}

cudaError_t launch_step(const LaunchArgs &a, int init_only, cudaStream_t stream)
{
    EpisodeScan scan{a.scan_status, a.scan_ticket, a.scan_epoch, a.cur_episode};
    const uint32_t tiles = num_tiles(a.N);
    return launch_pdl(step_kernel, tiles, TILE, stream, a, scan, tiles, init_only);
}

cudaError_t launch_rollout(const LaunchArgs &a, int num_steps, uint32_t seed, uint32_t first_step, cudaStream_t stream)
{
    if (num_steps <= 0) return cudaSuccess;
    const uint32_t tiles = num_tiles(a.N);
    return launch_cooperative(rollout_kernel, tiles, TILE, stream, a, RolloutScan{a.roll_ring, a.roll_step, a.cur_episode},
                              seed, first_step, tiles, num_steps, one_group_max_tiles());
}

}   // namespace balance
}   // namespace b200

// ------------------------------------------------------------------------------------------------------------
// b200env_cartpole_step_noalloc — a Cartpole step for callers that cannot allocate (CUDA graph capture, tight serving
// loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
// The caller passes every input and output; the only memory the step needs beyond them — the ordered
// episode numbering's scratch (tile status words, group sums, accumulators, ticket, epoch: EpisodeScan, common.cuh) —
// lives in the file-scope device array below, which every GPU that loads the library holds until exit (32 KB).
//   * NOALLOC_MAX_WORLDS bounds the batch (the array is sized for its 4 096 tiles); larger batches are refused.
//   * every call shares the array and finds the previous call's contents in it, so each call first clears the words it
//     will use (a tiny kernel on the caller's stream: status / sums / accumulators / ticket = 0, epoch = 1 — a zeroed
//     status word never carries tag 1);
//   * calls that overlap on the GPU (two streams, two host threads, two simulators) RACE on the array: serialise them.
// In host code `g_noalloc_scratch` is only a host-side stand-in for the symbol: its device address comes from
// cudaGetSymbolAddress, once per device with that device current, and is what the kernels are given.
// ------------------------------------------------------------------------------------------------------------
namespace b200 {
namespace cartpole {

constexpr uint32_t NOALLOC_MAX_WORLDS = 1u << 20;
constexpr uint32_t NOALLOC_SCRATCH_WORDS = 8192;                 // 32 KB
constexpr uint32_t NOALLOC_TICKET = NOALLOC_SCRATCH_WORDS - 64;  // word index of the ticket counter; epoch follows it
static __device__ __align__(256) uint32_t g_noalloc_scratch[NOALLOC_SCRATCH_WORDS];

// words [0, used) hold what EpisodeScanTile::publish_count / exclusive_base_warp index: u32 status [tiles], group sums
// at (tiles + 3) & ~3, two accumulators per group behind them, + the slack of the 128-bit polling loads
static uint32_t noalloc_words_used(uint32_t tiles)
{
    const uint32_t groups = (tiles + SCAN_GROUP - 1) / SCAN_GROUP;
    return ((tiles + 3u) & ~3u) + ((groups + 3u) & ~3u) + 2u * groups + 16u;
}
static_assert(((NOALLOC_MAX_WORLDS / TILE + 3u) & ~3u) + 64u + 2u * 64u + 16u <= NOALLOC_TICKET,
              "scratch array too small for NOALLOC_MAX_WORLDS");

__global__ void __launch_bounds__(TILE) noalloc_clear_kernel(uint32_t *scratch, const uint32_t used)
{
    for (uint32_t i = threadIdx.x; i < used; i += TILE) scratch[i] = 0u;
    if (threadIdx.x == 0) {
        scratch[NOALLOC_TICKET] = 0u;       // next tile index
        scratch[NOALLOC_TICKET + 1] = 1u;   // launch epoch: tag 1, never found in a zeroed status word
    }
}

// the step kernel's arithmetic (dynamics / fresh / ordered episode numbering), on caller-supplied tensors
__global__ void __launch_bounds__(TILE) step_noalloc_kernel(float *__restrict__ state, const int32_t *__restrict__ actions,
                                                            float *__restrict__ reward, int32_t *__restrict__ done_out,
                                                            const EpisodeScan scan, const uint32_t N,
                                                            const uint32_t num_tiles)
{
    __shared__ uint32_t s_scan[4];
    __shared__ uint32_t s_warp[TILE / 32];
    EpisodeScanTile st;
    st.begin(scan, s_scan);
    const uint32_t n = st.tile * TILE + threadIdx.x;
    const bool live = n < N;
    float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
    int done = 0;
    if (live) {
        s = reinterpret_cast<const float4 *>(state)[n];
        done = dynamics(s, actions[n]);
    }
    uint32_t rank, total;
    tile_rank(done, s_warp, rank, total);
    st.publish(scan, total, num_tiles);
    const uint32_t episode0 = st.exclusive_base(scan, total, num_tiles, s_scan + 3);
    if (live) {
        if (done) fresh(s, episode0 + rank);
        reinterpret_cast<float4 *>(state)[n] = s;
        reward[n] = 1.f;
        done_out[n] = done;
    }
}

// device address of the scratch array on the CURRENT device (cudaGetSymbolAddress once per device)
static cudaError_t noalloc_scratch_address(uint32_t **out)
{
    static uint32_t *addr[64] = {};
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if (dev < 0 || dev >= 64) return cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> lock(host_cache_mutex());
    if (!addr[dev]) {
        void *p = nullptr;
        e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
        if (e != cudaSuccess) return e;
        addr[dev] = static_cast<uint32_t *>(p);
    }
    *out = addr[dev];
    return cudaSuccess;
}

}   // namespace cartpole
}   // namespace b200

extern "C" uint32_t b200env_cartpole_step_noalloc_max_worlds(void) { return b200::cartpole::NOALLOC_MAX_WORLDS; }

extern "C" int b200env_cartpole_step_noalloc(int gpu_id, void *cuda_stream, uint32_t num_worlds, float *state,
                                             const int32_t *actions, float *reward, int32_t *done,
                                             uint32_t *episode_counter)
{
    using namespace b200::cartpole;
    if (!state || !actions || !reward || !done || !episode_counter) return b200::report_error(-1, "null argument");
    if (num_worlds == 0) return b200::report_error(-1, "num_worlds must be > 0");
    if (num_worlds > NOALLOC_MAX_WORLDS) {
        char msg[160];
        snprintf(msg, sizeof(msg), "b200env_cartpole_step_noalloc handles at most %u worlds per call (static scratch), not %u",
                 NOALLOC_MAX_WORLDS, num_worlds);
        return b200::report_error(-1, msg);
    }
    int prev = -1;
    if (cudaGetDevice(&prev) != cudaSuccess) prev = -1;
    if (prev != gpu_id && cudaSetDevice(gpu_id) != cudaSuccess) return b200::report_error(-2, "cudaSetDevice failed");
    uint32_t *scratch = nullptr;
    cudaError_t e = noalloc_scratch_address(&scratch);
    if (e == cudaSuccess) {
        const uint32_t tiles = b200::num_tiles(num_worlds);
        cudaStream_t st = (cudaStream_t)cuda_stream;
        noalloc_clear_kernel<<<1, b200::TILE, 0, st>>>(scratch, noalloc_words_used(tiles));
        b200::EpisodeScan scan{reinterpret_cast<unsigned long long *>(scratch), scratch + NOALLOC_TICKET,
                               scratch + NOALLOC_TICKET + 1, episode_counter};
        step_noalloc_kernel<<<tiles, b200::TILE, 0, st>>>(state, actions, reward, done, scan, num_worlds, tiles);
        e = cudaGetLastError();
    }
    if (prev >= 0 && prev != gpu_id) cudaSetDevice(prev);
    if (e != cudaSuccess) {
        char msg[200];
        snprintf(msg, sizeof(msg), "b200env_cartpole_step_noalloc: %s", cudaGetErrorString(e));
        return b200::report_error(-2, msg);
    }
    return 0;
}
