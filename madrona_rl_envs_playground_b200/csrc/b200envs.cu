// C-ABI implementation (include/b200envs.h): buffer ownership, tensor table, step dispatch, host-buffer step,
// checkpointing.  The kernels live in overcooked.cu / hanabi.cu / simple_envs.cu.
#include "../../include/b200envs.h"

#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "hanabi_host.h"
#include "overcooked.cuh"
#include "overcooked_host.h"
#include "random_policy_host.h"
#include "simple_envs_host.h"

namespace {

thread_local std::string g_error;

int fail(int code, const char *fmt, ...)
{
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    g_error = buf;
    return code;
}

#define CU(call)                                                                                          \
    do {                                                                                                  \
        cudaError_t e_ = (call);                                                                          \
        if (e_ != cudaSuccess)                                                                            \
            return fail(B200ENV_ERR_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, \
                        __LINE__);                                                                        \
    } while (0)

struct DeviceGuard {
    int prev = -1;
    bool ok = true;
    explicit DeviceGuard(int dev)
    {
        if (cudaGetDevice(&prev) != cudaSuccess) prev = -1;
        if (prev != dev) ok = cudaSetDevice(dev) == cudaSuccess;
    }
    ~DeviceGuard()
    {
        int cur = -1;
        if (prev >= 0 && cudaGetDevice(&cur) == cudaSuccess && cur != prev) cudaSetDevice(prev);
    }
};

struct Region {   // one device allocation that is part of the checkpointable state
    void *ptr;
    size_t bytes;
};

}   // namespace

namespace b200 {
int report_error(int code, const char *msg) { return fail(code, "%s", msg); }
}   // namespace b200

struct b200env {
    int kind = -1, gpu = 0;
    uint32_t N = 0;
    int P = 1;
    std::vector<void *> allocs;
    std::vector<Region> state;          // checkpoint regions, in a fixed order
    b200env_tensor_desc tensors[B200ENV_T_NUM_SLOTS];
    bool has[B200ENV_T_NUM_SLOTS] = {};
    size_t action_count = 0;

    // overcooked
    b200::oc::Layout layout;
    b200::oc::Layout *d_layout = nullptr;
    b200::oc::LaunchArgs oc{};
    bool fused_rollout = true;                 // b200env_overcooked_rollout: one fused launch (else one launch per step)
    unsigned long long *trace_buf = nullptr;   // debug timeline (b200env_debug_trace)
    size_t trace_stride = 0;
    int trace_steps = 0;
    // hanabi / balance / cartpole
    b200::hanabi::LaunchArgs hb{};
    b200::balance::LaunchArgs bb;
    b200::cartpole::LaunchArgs cp;
    b200::acrobat::LaunchArgs ac;

    // pipelined host-buffer steps (b200env_step_host_async / _wait): own streams, two snapshot buffers per tensor slot
    struct HostPipe {
        cudaStream_t step_stream = nullptr, copy_stream = nullptr;
        cudaEvent_t snapped[2] = {}, copied[2] = {};
        void *stage[2][B200ENV_T_NUM_SLOTS] = {};
        unsigned issued = 0, waited = 0;   // steps enqueued / waited for
    } *pipe = nullptr;

    template <class T>
    int alloc(T **out, size_t count, bool checkpoint)
    {
        void *p = nullptr;
        const size_t bytes = count * sizeof(T);
        cudaError_t e = cudaMalloc(&p, bytes ? bytes : 16);
        if (e != cudaSuccess) return fail(B200ENV_ERR_CUDA, "cudaMalloc(%zu) failed: %s", bytes, cudaGetErrorString(e));
        e = cudaMemset(p, 0, bytes ? bytes : 16);
        if (e != cudaSuccess) return fail(B200ENV_ERR_CUDA, "cudaMemset failed: %s", cudaGetErrorString(e));
        allocs.push_back(p);
        if (checkpoint) state.push_back({p, bytes});
        *out = (T *)p;
        return B200ENV_OK;
    }

    void set_tensor(int slot, void *ptr, int dtype, std::initializer_list<int64_t> dims)
    {
        b200env_tensor_desc &t = tensors[slot];
        t.ptr = ptr;
        t.dtype = dtype;
        t.ndim = (int32_t)dims.size();
        int i = 0;
        for (int64_t d : dims) t.dims[i++] = d;
        for (; i < 6; i++) t.dims[i] = 1;
        t.device = gpu;
        has[slot] = true;
    }

    int launch(cudaStream_t st, int init_only)
    {
        cudaError_t e = cudaSuccess;
        switch (kind) {
        case B200ENV_OVERCOOKED: e = b200::oc::launch_step(layout, d_layout, oc, init_only, st); break;
        case B200ENV_HANABI: e = b200::hanabi::launch_step(hb, init_only, st); break;
        case B200ENV_BALANCE: e = b200::balance::launch_step(bb, init_only, st); break;
        case B200ENV_CARTPOLE: e = b200::cartpole::launch_step(cp, init_only, st); break;
        case B200ENV_ACROBAT: e = b200::acrobat::launch_step(ac, init_only, st); break;
        default: return fail(B200ENV_ERR_INVALID, "bad simulator handle");
        }
        if (e != cudaSuccess) return fail(B200ENV_ERR_CUDA, "kernel launch failed: %s", cudaGetErrorString(e));
        return B200ENV_OK;
    }
};

namespace {

#define TRY(expr)                      \
    do {                               \
        int rc_ = (expr);              \
        if (rc_ != B200ENV_OK) return rc_; \
    } while (0)

// id tensors: agent_id[p][n] = p, world_id[p][n] = n
int fill_ids(b200env *env, int P)
{
    int32_t *wid = nullptr, *aid = nullptr;
    TRY(env->alloc(&wid, (size_t)P * env->N, false));
    TRY(env->alloc(&aid, (size_t)P * env->N, false));
    std::vector<int32_t> h((size_t)P * env->N);
    for (int p = 0; p < P; p++)
        for (uint32_t n = 0; n < env->N; n++) h[(size_t)p * env->N + n] = (int32_t)n;
    CU(cudaMemcpy(wid, h.data(), h.size() * 4, cudaMemcpyHostToDevice));
    for (int p = 0; p < P; p++)
        for (uint32_t n = 0; n < env->N; n++) h[(size_t)p * env->N + n] = p;
    CU(cudaMemcpy(aid, h.data(), h.size() * 4, cudaMemcpyHostToDevice));
    env->set_tensor(B200ENV_T_WORLD_ID, wid, B200ENV_I32, {P, env->N});
    env->set_tensor(B200ENV_T_AGENT_ID, aid, B200ENV_I32, {P, env->N});
    return B200ENV_OK;
}

int fill_ones(int32_t *dev, size_t count)
{
    std::vector<int32_t> h(count, 1);
    CU(cudaMemcpy(dev, h.data(), count * 4, cudaMemcpyHostToDevice));
    return B200ENV_OK;
}

// scan bookkeeping for the RNG envs: status array (zero => epoch 0 never matches), ticket = 0, epoch = 1, episode = 0
template <class Args>
int alloc_scan(b200env *env, Args &a, uint32_t tiles)
{
    TRY(env->alloc(&a.scan_status, (size_t)tiles + 16, false));   // tile words + group words, see EpisodeScan
    TRY(env->alloc(&a.scan_ticket, 1, false));
    TRY(env->alloc(&a.scan_epoch, 1, false));
    TRY(env->alloc(&a.cur_episode, 1, true));
    const uint32_t one = 1;
    CU(cudaMemcpy(a.scan_epoch, &one, 4, cudaMemcpyHostToDevice));
    return B200ENV_OK;
}

int finish_create(b200env *env, b200env **out)
{
    // initial reset + initial observation (the reference does this in Sim::Sim)
    TRY(env->launch(0, 1));
    CU(cudaStreamSynchronize(0));
    *out = env;
    return B200ENV_OK;
}

}   // namespace

extern "C" {

const char *b200env_last_error(void) { return g_error.c_str(); }
const char *b200env_version(void) { return "b200envs 0.1 (sm_100a)"; }

int b200env_kind(const b200env *env) { return env ? env->kind : B200ENV_ERR_INVALID; }
uint32_t b200env_num_worlds(const b200env *env) { return env ? env->N : 0; }

void b200env_destroy(b200env *env)
{
    if (!env) return;
    DeviceGuard g(env->gpu);
    cudaDeviceSynchronize();
    if (env->pipe) {
        for (int b = 0; b < 2; b++) {
            if (env->pipe->snapped[b]) cudaEventDestroy(env->pipe->snapped[b]);
            if (env->pipe->copied[b]) cudaEventDestroy(env->pipe->copied[b]);
        }
        if (env->pipe->step_stream) cudaStreamDestroy(env->pipe->step_stream);
        if (env->pipe->copy_stream) cudaStreamDestroy(env->pipe->copy_stream);
        delete env->pipe;
    }
    for (void *p : env->allocs) cudaFree(p);
    delete env;
}

// rules 0 = src/overcooked_env, 1 = src/overcooked2_env ("Simplecooked")
static int overcooked_create(const b200env_overcooked_config *c, int gpu_id, uint32_t num_worlds, b200env **out, int rules,
                             int obs_dtype = B200ENV_I8)
{
    using namespace b200::oc;
    if (obs_dtype != B200ENV_I8 && obs_dtype != B200ENV_BF16 && obs_dtype != B200ENV_F32)
        return fail(B200ENV_ERR_INVALID, "observation dtype must be B200ENV_I8, B200ENV_BF16 or B200ENV_F32");
    const int es = obs_dtype == B200ENV_I8 ? 1 : obs_dtype == B200ENV_BF16 ? 2 : 4;
    if (!c || !out || !c->terrain || !c->start_player_x || !c->start_player_y || !c->recipe_values || !c->recipe_times)
        return fail(B200ENV_ERR_INVALID, "null argument");
    const int64_t S = c->height * c->width, P = c->num_players;
    // the reference stores size / positions in uint8_t (sim.hpp:86,150) and caps players at 64 (sim.hpp:14)
    if (c->height <= 0 || c->width <= 0 || S > 255) return fail(B200ENV_ERR_INVALID, "grid must have 1..255 cells");
    if (P < 1 || P > 64) return fail(B200ENV_ERR_INVALID, "num_players must be 1..64");
    // Simplecooked: MAX_NUM_PLAYERS 2, MAX_SIZE 100 (src/overcooked2_env/sim.hpp:13-14)
    if (rules && (P > 2 || S > 100)) return fail(B200ENV_ERR_INVALID, "Simplecooked allows at most 2 players and 100 cells");
    if (num_worlds == 0) return fail(B200ENV_ERR_INVALID, "num_worlds must be > 0");
    DeviceGuard g(gpu_id);
    if (!g.ok) return fail(B200ENV_ERR_CUDA, "cudaSetDevice(%d) failed", gpu_id);

    b200env *env = new b200env;
    env->kind = B200ENV_OVERCOOKED; env->gpu = gpu_id; env->N = num_worlds; env->P = (int)P;
    Layout &L = env->layout;
    memset(&L, 0, sizeof(L));
    L.W = (int)c->width; L.H = (int)c->height; L.S = (int)S; L.P = (int)P; L.C = 5 * (int)P + (rules ? 10 : 16);
    L.rules = rules;
    L.dish_rew = (uint8_t)c->dish_pickup_rew;
    L.horizon = (int)c->horizon;
    L.placement_rew = (uint8_t)c->placement_in_pot_rew;   // WorldState narrows to uint8_t, sim.cpp:603-605
    L.soup_rew = (uint8_t)c->soup_pickup_rew;             // dish_pickup_rew is accepted and never used, sim.cpp:240-243
    for (int cell = 0; cell < L.S; cell++) {
        int64_t t = c->terrain[cell];
        if (t < 0 || t > 6) { delete env; return fail(B200ENV_ERR_INVALID, "terrain[%d] = %lld out of range", cell, (long long)t); }
        if (rules) {   // AIR, POT, COUNTER, ONION_SOURCE, DISH_SOURCE, SERVING, TOMATO_SOURCE -> this library's order
            static const int64_t from_simple[7] = {0, 1, 2, 3, 5, 6, 4};
            t = from_simple[t];
        }
        const int x = cell % L.W, y = cell / L.W;
        uint32_t slot = 0xFFu;
        if (t == POT || t == COUNTER) {
            slot = (uint32_t)L.K;
            L.cell_of_slot[L.K] = (uint8_t)cell;
            if (t == POT) L.pot_slot[L.npots++] = (uint8_t)L.K;
            L.K++;
        }
        L.cell[cell] = (uint32_t)t | (slot << 8) | ((uint32_t)(x * L.H + y) << 16);
    }
    for (int p = 0; p < L.P; p++) {
        const int64_t x = c->start_player_x[p], y = c->start_player_y[p];
        if (x < 0 || x >= L.W || y < 0 || y >= L.H) { delete env; return fail(B200ENV_ERR_INVALID, "start position of player %d outside the grid", p); }
        L.start_pos[p] = (uint8_t)(y * L.W + x);
    }
    for (int r = 0; r < 16; r++) {
        L.recipe_values[r] = (uint8_t)c->recipe_values[r];
        L.recipe_times[r] = (uint8_t)c->recipe_times[r];
    }
    // movement never bounds-checks (layout borders are walls in every reference layout); reject layouts where a
    // player could step off the grid instead of reading out of bounds
    for (int cell = 0; cell < L.S; cell++) {
        const int x = cell % L.W, y = cell / L.W;
        if ((L.cell[cell] & 0xFFu) == AIR && (x == 0 || y == 0 || x == L.W - 1 || y == L.H - 1)) {
            delete env;
            return fail(B200ENV_ERR_INVALID, "walkable cell (%d,%d) on the grid border", x, y);
        }
    }

    const size_t N = num_worlds, SC = (size_t)L.S * L.C;
    int rc = B200ENV_OK;
    uint2 *players = nullptr; int32_t *actions = nullptr, *active = nullptr, *mask = nullptr;
    auto chk = [&](int r) { if (rc == B200ENV_OK) rc = r; };
    chk(env->alloc(&env->d_layout, 1, false));
    chk(env->alloc(&players, (size_t)P * N, true));
    chk(env->alloc(&env->oc.objs, (size_t)(L.K ? L.K : 1) * N, true));
    chk(env->alloc(&env->oc.timestep, N, true));
    chk(env->alloc(&actions, (size_t)P * N, true));
    chk(env->alloc(&env->oc.rewards, (size_t)P * N, true));
    chk(env->alloc(&env->oc.done, N, true));
    chk(env->alloc(&env->oc.obs, (size_t)P * N * SC * es, true));
    env->oc.obs_elem = es;
    chk(env->alloc(&env->oc.ep_return, N, true));
    chk(env->alloc(&env->oc.stats, 2, true));
    chk(env->alloc(&active, (size_t)P * N, false));
    chk(env->alloc(&mask, (size_t)P * N * 6, false));
    if (rc == B200ENV_OK) rc = fill_ids(env, (int)P);
    if (rc == B200ENV_OK) rc = fill_ones(active, (size_t)P * N);       // ActiveAgent is always 1, sim.cpp:619
    if (rc == B200ENV_OK) rc = fill_ones(mask, (size_t)P * N * 6);     // every move always valid, sim.cpp:616-618
    if (rc == B200ENV_OK && cudaMemcpy(env->d_layout, &L, sizeof(L), cudaMemcpyHostToDevice) != cudaSuccess)
        rc = fail(B200ENV_ERR_CUDA, "layout upload failed");
    // constant blocks of the warp-tile kernel, one per lanes-per-world variant that fits in shared memory
    for (int gi = 0; gi < 4 && rc == B200ENV_OK && P <= 8; gi++) {
        const int G = 1 << gi;
        if (!tile_fits(L, G, es)) continue;
        const size_t bytes = tile_const_bytes(L, G, es), stride = tile_const_stride(L, G, es);
        std::vector<uint8_t> blk(stride * TILE_CONST_COPIES);
        build_tile_const(L, G, es, blk.data());
        for (int c = 1; c < TILE_CONST_COPIES; c++) memcpy(blk.data() + c * stride, blk.data(), bytes);
        uint8_t *d_blk = nullptr;
        rc = env->alloc(&d_blk, blk.size(), false);
        if (rc == B200ENV_OK && cudaMemcpy(d_blk, blk.data(), blk.size(), cudaMemcpyHostToDevice) != cudaSuccess)
            rc = fail(B200ENV_ERR_CUDA, "tile constant upload failed");
        env->oc.tile_const[gi] = d_blk;
    }
    env->oc.lanes = pick_lanes(L, num_worlds, es);
    if (rc == B200ENV_OK && rules && pick_kernel(L, num_worlds, es) != 1)
        rc = fail(B200ENV_ERR_NOSLOT, "the Simplecooked rules run on the warp-tile kernel only and this layout does not fit it");
    if (rc == B200ENV_OK && pick_kernel(L, num_worlds, es) != 1 && !world_cta_fits(L, es))
        rc = fail(B200ENV_ERR_NOSLOT, "one observer's slab of this layout (cells x channels x element size) does not fit in "
                                      "shared memory with this observation type; use a narrower one");
    if (rc != B200ENV_OK) { b200env_destroy(env); return rc; }
    env->oc.players = players; env->oc.actions = actions; env->oc.N = num_worlds; env->oc.force_kernel = -1;
    env->action_count = (size_t)P * N;

    env->set_tensor(B200ENV_T_DONE, env->oc.done, B200ENV_I32, {(int64_t)N});
    env->set_tensor(B200ENV_T_ACTIVE_AGENT, active, B200ENV_I32, {P, (int64_t)N});
    env->set_tensor(B200ENV_T_ACTION, actions, B200ENV_I32, {P, (int64_t)N, 1});
    env->set_tensor(B200ENV_T_OBSERVATION, env->oc.obs, obs_dtype, {P, (int64_t)N, L.W, L.H, L.C});
    env->set_tensor(B200ENV_T_ACTION_MASK, mask, B200ENV_I32, {P, (int64_t)N, 6});
    env->set_tensor(B200ENV_T_REWARD, env->oc.rewards, B200ENV_I32, {P, (int64_t)N});
    rc = finish_create(env, out);
    if (rc != B200ENV_OK) b200env_destroy(env);
    return rc;
}

int b200env_overcooked_create(const b200env_overcooked_config *c, int gpu_id, uint32_t num_worlds, b200env **out)
{
    return overcooked_create(c, gpu_id, num_worlds, out, 0);
}

int b200env_overcooked2_create(const b200env_overcooked_config *c, int gpu_id, uint32_t num_worlds, b200env **out)
{
    return overcooked_create(c, gpu_id, num_worlds, out, 1);
}

int b200env_overcooked_create_typed(const b200env_overcooked_config *c, int rules, int obs_dtype, int gpu_id,
                                    uint32_t num_worlds, b200env **out)
{
    if (rules != 0 && rules != 1) return fail(B200ENV_ERR_INVALID, "rules must be 0 (Overcooked) or 1 (Simplecooked)");
    return overcooked_create(c, gpu_id, num_worlds, out, rules, obs_dtype);
}

static int hanabi_create(const b200env_hanabi_config *c, int obs_dtype, int gpu_id, uint32_t num_worlds, b200env **out)
{
    using namespace b200::hanabi;
    if (!c || !out) return fail(B200ENV_ERR_INVALID, "null argument");
    if (obs_dtype != B200ENV_I8 && obs_dtype != B200ENV_BF16 && obs_dtype != B200ENV_F32)
        return fail(B200ENV_ERR_INVALID, "observation dtype must be B200ENV_I8, B200ENV_BF16 or B200ENV_F32");
    const size_t es = obs_dtype == B200ENV_I8 ? 1 : obs_dtype == B200ENV_BF16 ? 2 : 4;
    if (c->players != 2) return fail(B200ENV_ERR_INVALID, "only 2-player Hanabi is supported (as in the reference's tensor shapes)");
    if (c->colors < 1 || c->colors > 5 || c->ranks < 2 || c->ranks > 5)
        return fail(B200ENV_ERR_INVALID, "colors must be 1..5 and ranks 2..5");
    if (c->max_information_tokens < 1 || c->max_information_tokens > 8 || c->max_life_tokens < 1 || c->max_life_tokens > 3)
        return fail(B200ENV_ERR_INVALID, "information tokens must be 1..8 and life tokens 1..3");
    if (2 * c->ranks * c->colors < 10)   // (the reference deals from an empty deck there: undefined, sim.cpp:45-52 / 508-519)
        return fail(B200ENV_ERR_INVALID, "the deck (2 * ranks * colors cards) must hold the 10 cards of the deal");
    if (num_worlds == 0) return fail(B200ENV_ERR_INVALID, "num_worlds must be > 0");
    DeviceGuard g(gpu_id);
    if (!g.ok) return fail(B200ENV_ERR_CUDA, "cudaSetDevice(%d) failed", gpu_id);

    b200env *env = new b200env;
    env->kind = B200ENV_HANABI; env->gpu = gpu_id; env->N = num_worlds; env->P = 2;
    LaunchArgs &a = env->hb;
    memset(&a, 0, sizeof(a));
    a.N = num_worlds; a.colors = (int)c->colors; a.ranks = (int)c->ranks;
    a.max_info = (int)c->max_information_tokens; a.max_life = (int)c->max_life_tokens;
    const size_t N = num_worlds;
    int rc = B200ENV_OK;
    int32_t *actions = nullptr;
    auto chk = [&](int r) { if (rc == B200ENV_OK) rc = r; };
    chk(env->alloc(&a.game, (size_t)GAME_WORDS * N, true));
    chk(env->alloc(&actions, 2 * N, true));
    chk(env->alloc(&a.obs, 2 * N * 658 * es, true));
    chk(env->alloc(&a.state, 2 * N * 783 * es, true));
    a.obs_elem = (int)es;
    chk(env->alloc(&a.mask, 2 * N * 20, true));
    chk(env->alloc(&a.active, 2 * N, true));
    chk(env->alloc(&a.done, N, true));
    chk(env->alloc(&a.reward, 2 * N, true));
    if (rc == B200ENV_OK) rc = alloc_scan(env, a, num_tiles(num_worlds));
    chk(env->alloc(&a.roll_ring, rollout_ring_words(num_worlds), false));
    chk(env->alloc(&a.roll_step, 1, false));
    chk(env->alloc(&a.roll_quota, 256, false));
    chk(env->alloc(&a.roll_ctl, 8, false));
    if (rc == B200ENV_OK) rc = fill_ids(env, 2);
    if (rc == B200ENV_OK) {   // fused rollouts deal tiles to SMs by their store rate (state tensor = scratch, re-zeroed)
        cudaError_t e = prepare_rollout_balance(a, a.state, 2 * N * 783 * es);
        if (e != cudaSuccess) rc = fail(B200ENV_ERR_CUDA, "SM rate measurement failed: %s", cudaGetErrorString(e));
    }
    if (rc != B200ENV_OK) { b200env_destroy(env); return rc; }
    a.actions = actions;
    env->action_count = 2 * N;
    env->set_tensor(B200ENV_T_DONE, a.done, B200ENV_I32, {(int64_t)N});
    env->set_tensor(B200ENV_T_ACTIVE_AGENT, a.active, B200ENV_I32, {2, (int64_t)N});
    env->set_tensor(B200ENV_T_ACTION, actions, B200ENV_I32, {2, (int64_t)N, 1});
    env->set_tensor(B200ENV_T_OBSERVATION, a.obs, obs_dtype, {2, (int64_t)N, 658});
    env->set_tensor(B200ENV_T_AGENT_STATE, a.state, obs_dtype, {2, (int64_t)N, 783});
    env->set_tensor(B200ENV_T_ACTION_MASK, a.mask, B200ENV_I32, {2, (int64_t)N, 20});
    env->set_tensor(B200ENV_T_REWARD, a.reward, B200ENV_F32, {2, (int64_t)N});
    rc = finish_create(env, out);
    if (rc != B200ENV_OK) b200env_destroy(env);
    return rc;
}

int b200env_hanabi_create(const b200env_hanabi_config *c, int gpu_id, uint32_t num_worlds, b200env **out)
{
    return hanabi_create(c, B200ENV_I8, gpu_id, num_worlds, out);
}

int b200env_hanabi_create_typed(const b200env_hanabi_config *c, int obs_dtype, int gpu_id, uint32_t num_worlds, b200env **out)
{
    return hanabi_create(c, obs_dtype, gpu_id, num_worlds, out);
}

int b200env_balance_create(int gpu_id, uint32_t num_worlds, b200env **out)
{
    using namespace b200::balance;
    if (!out) return fail(B200ENV_ERR_INVALID, "null argument");
    if (num_worlds == 0) return fail(B200ENV_ERR_INVALID, "num_worlds must be > 0");
    DeviceGuard g(gpu_id);
    if (!g.ok) return fail(B200ENV_ERR_CUDA, "cudaSetDevice(%d) failed", gpu_id);
    b200env *env = new b200env;
    env->kind = B200ENV_BALANCE; env->gpu = gpu_id; env->N = num_worlds; env->P = 2;
    LaunchArgs &a = env->bb;
    memset(&a, 0, sizeof(a));
    a.N = num_worlds;
    const size_t N = num_worlds;
    int rc = B200ENV_OK;
    int32_t *actions = nullptr, *active = nullptr, *mask = nullptr;
    auto chk = [&](int r) { if (rc == B200ENV_OK) rc = r; };
    chk(env->alloc(&a.obs, 2 * N * 7, true));
    chk(env->alloc(&actions, 2 * N, true));
    chk(env->alloc(&a.reward, 2 * N, true));
    chk(env->alloc(&a.done, N, true));
    chk(env->alloc(&a.loc, 2 * N, true));
    chk(env->alloc(&a.time, N, true));
    chk(env->alloc(&active, 2 * N, false));
    chk(env->alloc(&mask, 2 * N * 4, false));
    if (rc == B200ENV_OK) rc = alloc_scan(env, a, b200::num_tiles(num_worlds));
    chk(env->alloc(&a.roll_ring, b200::simple_roll_ring_words(num_worlds), false));
    chk(env->alloc(&a.roll_step, 1, false));
    if (rc == B200ENV_OK) rc = fill_ids(env, 2);
    if (rc == B200ENV_OK) rc = fill_ones(active, 2 * N);        // sim.cpp:200
    if (rc == B200ENV_OK) rc = fill_ones(mask, 2 * N * 4);      // sim.cpp:196-198
    if (rc != B200ENV_OK) { b200env_destroy(env); return rc; }
    a.actions = actions;
    env->action_count = 2 * N;
    env->set_tensor(B200ENV_T_DONE, a.done, B200ENV_I32, {(int64_t)N});
    env->set_tensor(B200ENV_T_ACTIVE_AGENT, active, B200ENV_I32, {2, (int64_t)N});
    env->set_tensor(B200ENV_T_ACTION, actions, B200ENV_I32, {2, (int64_t)N, 1});
    env->set_tensor(B200ENV_T_OBSERVATION, a.obs, B200ENV_I32, {2, (int64_t)N, 7});
    env->set_tensor(B200ENV_T_ACTION_MASK, mask, B200ENV_I32, {2, (int64_t)N, 4});
    env->set_tensor(B200ENV_T_REWARD, a.reward, B200ENV_F32, {2, (int64_t)N});
    rc = finish_create(env, out);
    if (rc != B200ENV_OK) b200env_destroy(env);
    return rc;
}

int b200env_cartpole_create(int gpu_id, uint32_t num_worlds, b200env **out)
{
    using namespace b200::cartpole;
    if (!out) return fail(B200ENV_ERR_INVALID, "null argument");
    if (num_worlds == 0) return fail(B200ENV_ERR_INVALID, "num_worlds must be > 0");
    DeviceGuard g(gpu_id);
    if (!g.ok) return fail(B200ENV_ERR_CUDA, "cudaSetDevice(%d) failed", gpu_id);
    b200env *env = new b200env;
    env->kind = B200ENV_CARTPOLE; env->gpu = gpu_id; env->N = num_worlds; env->P = 1;
    LaunchArgs &a = env->cp;
    memset(&a, 0, sizeof(a));
    a.N = num_worlds;
    const size_t N = num_worlds;
    int rc = B200ENV_OK;
    int32_t *actions = nullptr;
    auto chk = [&](int r) { if (rc == B200ENV_OK) rc = r; };
    chk(env->alloc(&a.state, 4 * N, true));
    chk(env->alloc(&actions, N, true));
    chk(env->alloc(&a.reward, N, true));
    chk(env->alloc(&a.done, N, true));
    if (rc == B200ENV_OK) rc = alloc_scan(env, a, b200::num_tiles(num_worlds));
    chk(env->alloc(&a.roll_ring, b200::simple_roll_ring_words(num_worlds), false));
    chk(env->alloc(&a.roll_step, 1, false));
    if (rc == B200ENV_OK) rc = fill_ids(env, 1);
    if (rc != B200ENV_OK) { b200env_destroy(env); return rc; }
    a.actions = actions;
    env->action_count = N;
    env->has[B200ENV_T_AGENT_ID] = false;
    env->set_tensor(B200ENV_T_DONE, a.done, B200ENV_I32, {(int64_t)N, 1});
    env->set_tensor(B200ENV_T_ACTION, actions, B200ENV_I32, {(int64_t)N, 1});
    env->set_tensor(B200ENV_T_AGENT_STATE, a.state, B200ENV_F32, {(int64_t)N, 4});
    env->set_tensor(B200ENV_T_REWARD, a.reward, B200ENV_F32, {(int64_t)N, 1});
    env->tensors[B200ENV_T_WORLD_ID].ndim = 2;
    env->tensors[B200ENV_T_WORLD_ID].dims[0] = (int64_t)N;
    env->tensors[B200ENV_T_WORLD_ID].dims[1] = 1;
    rc = finish_create(env, out);
    if (rc != B200ENV_OK) b200env_destroy(env);
    return rc;
}

int b200env_acrobat_create(int gpu_id, uint32_t num_worlds, b200env **out)
{
    using namespace b200::acrobat;
    if (!out) return fail(B200ENV_ERR_INVALID, "null argument");
    if (num_worlds == 0) return fail(B200ENV_ERR_INVALID, "num_worlds must be > 0");
    DeviceGuard g(gpu_id);
    if (!g.ok) return fail(B200ENV_ERR_CUDA, "cudaSetDevice(%d) failed", gpu_id);
    b200env *env = new b200env;
    env->kind = B200ENV_ACROBAT; env->gpu = gpu_id; env->N = num_worlds; env->P = 1;
    LaunchArgs &a = env->ac;
    memset(&a, 0, sizeof(a));
    a.N = num_worlds;
    const size_t N = num_worlds;
    int rc = B200ENV_OK;
    int32_t *actions = nullptr;
    auto chk = [&](int r) { if (rc == B200ENV_OK) rc = r; };
    chk(env->alloc(&a.state, 4 * N, true));
    chk(env->alloc(&actions, N, true));
    chk(env->alloc(&a.reward, N, true));
    chk(env->alloc(&a.done, N, true));
    chk(env->alloc(&a.ep_len, 1, true));
    chk(env->alloc(&a.term_status, (size_t)b200::num_tiles(num_worlds) + 16, false));
    if (rc == B200ENV_OK) rc = alloc_scan(env, a, b200::num_tiles(num_worlds));
    if (rc == B200ENV_OK) rc = fill_ids(env, 1);
    if (rc != B200ENV_OK) { b200env_destroy(env); return rc; }
    a.actions = actions;
    env->action_count = N;
    env->has[B200ENV_T_AGENT_ID] = false;
    // src/acrobat_env/mgr.cpp:172-201: reset i32 [N,1], action i32 [N,1], state f32 [N,4], reward f32 [N,1], world id
    env->set_tensor(B200ENV_T_DONE, a.done, B200ENV_I32, {(int64_t)N, 1});
    env->set_tensor(B200ENV_T_ACTION, actions, B200ENV_I32, {(int64_t)N, 1});
    env->set_tensor(B200ENV_T_AGENT_STATE, a.state, B200ENV_F32, {(int64_t)N, 4});
    env->set_tensor(B200ENV_T_REWARD, a.reward, B200ENV_F32, {(int64_t)N, 1});
    env->tensors[B200ENV_T_WORLD_ID].ndim = 2;
    env->tensors[B200ENV_T_WORLD_ID].dims[0] = (int64_t)N;
    env->tensors[B200ENV_T_WORLD_ID].dims[1] = 1;
    rc = finish_create(env, out);
    if (rc != B200ENV_OK) b200env_destroy(env);
    return rc;
}

#define NO_STEPS_IN_FLIGHT(env)                                                                              \
    do {                                                                                                     \
        if ((env)->pipe && (env)->pipe->issued != (env)->pipe->waited)                                       \
            return fail(B200ENV_ERR_INVALID, "b200env_step_host_async steps are in flight: b200env_step_host_wait first"); \
    } while (0)

int b200env_step_async(b200env *env, void *cuda_stream)
{
    if (!env) return fail(B200ENV_ERR_INVALID, "null simulator");
    NO_STEPS_IN_FLIGHT(env);
    DeviceGuard g(env->gpu);
    return env->launch((cudaStream_t)cuda_stream, 0);
}

int b200env_step(b200env *env)
{
    if (!env) return fail(B200ENV_ERR_INVALID, "null simulator");
    NO_STEPS_IN_FLIGHT(env);
    DeviceGuard g(env->gpu);
    TRY(env->launch(0, 0));
    CU(cudaStreamSynchronize(0));
    return B200ENV_OK;
}

// The reference-raw Overcooked tensors are 13x the native observation and nobody on the native
// path wants them: allocated and filled when first asked for.
static int ensure_compat_tensors(b200env *env, int slot)
{
    if (env->kind != B200ENV_OVERCOOKED || env->has[slot]) return B200ENV_OK;
    if (env->oc.obs_elem != 1) return fail(B200ENV_ERR_NOSLOT, "the reference-raw export exists for int8 observations only");
    DeviceGuard g(env->gpu);
    const int64_t P = env->layout.P, S = env->layout.S, N = env->N;
    const size_t rows = (size_t)P * S * N;
    if (slot == B200ENV_T_OBSERVATION_RAW) {
        uint8_t *raw = nullptr;
        TRY(env->alloc(&raw, rows * b200::oc::RAW_ROW_BYTES, false));
        env->set_tensor(slot, raw, B200ENV_I8, {P * S, N, b200::oc::RAW_ROW_BYTES});
        return B200ENV_OK;
    }
    // flat row k = world * S*P + cell * P + player (creation order inside a world, worlds concatenated:
    // madrona/src/core/state.cpp:511-554; probe output in SURVEY Appendix A) holds player * S + cell and world
    int32_t *lw = nullptr, *li = nullptr;
    TRY(env->alloc(&lw, rows, false));
    TRY(env->alloc(&li, rows, false));
    std::vector<int32_t> h(rows);
    for (size_t k = 0; k < rows; k++) h[k] = (int32_t)(k / (size_t)(S * P));
    CU(cudaMemcpy(lw, h.data(), rows * 4, cudaMemcpyHostToDevice));
    for (size_t k = 0; k < rows; k++) {
        const size_t r = k % (size_t)(S * P);
        h[k] = (int32_t)((r % (size_t)P) * (size_t)S + r / (size_t)P);
    }
    CU(cudaMemcpy(li, h.data(), rows * 4, cudaMemcpyHostToDevice));
    env->set_tensor(B200ENV_T_LOCATION_WORLD_ID, lw, B200ENV_I32, {P * S, N});
    env->set_tensor(B200ENV_T_LOCATION_ID, li, B200ENV_I32, {P * S, N});
    return B200ENV_OK;
}

int b200env_overcooked_export_raw(b200env *env, void *cuda_stream, void *dst)
{
    if (!env || env->kind != B200ENV_OVERCOOKED) return fail(B200ENV_ERR_INVALID, "not an Overcooked simulator");
    if (env->oc.obs_elem != 1) return fail(B200ENV_ERR_NOSLOT, "the reference-raw export exists for int8 observations only");
    NO_STEPS_IN_FLIGHT(env);
    if (!dst) {
        TRY(ensure_compat_tensors(env, B200ENV_T_OBSERVATION_RAW));
        dst = env->tensors[B200ENV_T_OBSERVATION_RAW].ptr;
    }
    DeviceGuard g(env->gpu);
    cudaError_t e = b200::oc::launch_export_raw(env->layout, env->oc, (uint8_t *)dst, (cudaStream_t)cuda_stream);
    if (e != cudaSuccess) return fail(B200ENV_ERR_CUDA, "export launch failed: %s", cudaGetErrorString(e));
    return B200ENV_OK;
}

int b200env_tensor(const b200env *env, int slot, b200env_tensor_desc *out)
{
    if (!env || !out) return fail(B200ENV_ERR_INVALID, "null argument");
    if (slot >= B200ENV_T_LOCATION_WORLD_ID && slot <= B200ENV_T_OBSERVATION_RAW && env->kind == B200ENV_OVERCOOKED)
        TRY(ensure_compat_tensors(const_cast<b200env *>(env), slot));
    if (slot < 0 || slot >= B200ENV_T_NUM_SLOTS || !env->has[slot])
        return fail(B200ENV_ERR_NOSLOT, "tensor slot %d is not exported by this environment", slot);
    *out = env->tensors[slot];
    return B200ENV_OK;
}

static size_t tensor_bytes(const b200env_tensor_desc &t)
{
    size_t n = t.dtype == B200ENV_I8 ? 1 : t.dtype == B200ENV_BF16 ? 2 : 4;
    for (int i = 0; i < t.ndim; i++) n *= (size_t)t.dims[i];
    return n;
}

int b200env_step_host(b200env *env, const int32_t *actions, const int *slots, void *const *host_ptrs, int num_slots)
{
    if (!env || !actions || (num_slots > 0 && (!slots || !host_ptrs))) return fail(B200ENV_ERR_INVALID, "null argument");
    for (int i = 0; i < num_slots; i++)
        if (slots[i] < 0 || slots[i] >= B200ENV_T_NUM_SLOTS || !env->has[slots[i]] || !host_ptrs[i])
            return fail(B200ENV_ERR_NOSLOT, "bad output slot %d", slots[i]);
    NO_STEPS_IN_FLIGHT(env);
    DeviceGuard g(env->gpu);
    CU(cudaMemcpyAsync(env->tensors[B200ENV_T_ACTION].ptr, actions, env->action_count * 4, cudaMemcpyHostToDevice, 0));
    TRY(env->launch(0, 0));
    for (int i = 0; i < num_slots; i++) {
        const b200env_tensor_desc &t = env->tensors[slots[i]];
        CU(cudaMemcpyAsync(host_ptrs[i], t.ptr, tensor_bytes(t), cudaMemcpyDeviceToHost, 0));
    }
    CU(cudaStreamSynchronize(0));
    return B200ENV_OK;
}

int b200env_step_host_async(b200env *env, const int32_t *actions, const int *slots, void *const *host_ptrs, int num_slots)
{
    if (!env || !actions || (num_slots > 0 && (!slots || !host_ptrs))) return fail(B200ENV_ERR_INVALID, "null argument");
    for (int i = 0; i < num_slots; i++)
        if (slots[i] < 0 || slots[i] >= B200ENV_T_NUM_SLOTS || !env->has[slots[i]] || !host_ptrs[i])
            return fail(B200ENV_ERR_NOSLOT, "bad output slot %d", slots[i]);
    DeviceGuard g(env->gpu);
    if (!env->pipe) {
        b200env::HostPipe *hp = new b200env::HostPipe;
        env->pipe = hp;
        CU(cudaStreamCreateWithFlags(&hp->step_stream, cudaStreamNonBlocking));
        CU(cudaStreamCreateWithFlags(&hp->copy_stream, cudaStreamNonBlocking));
        for (int b = 0; b < 2; b++) {
            CU(cudaEventCreateWithFlags(&hp->snapped[b], cudaEventDisableTiming));
            CU(cudaEventCreateWithFlags(&hp->copied[b], cudaEventDisableTiming));
        }
    }
    b200env::HostPipe &hp = *env->pipe;
    if (hp.issued - hp.waited >= 2) return fail(B200ENV_ERR_INVALID, "two steps already in flight: b200env_step_host_wait first");
    if (hp.issued == hp.waited) CU(cudaStreamSynchronize(0));   // first of a run: after whatever the synchronous calls did
    const int b = (int)(hp.issued & 1u);
    for (int i = 0; i < num_slots; i++)
        if (!hp.stage[b][slots[i]]) {
            uint8_t *p = nullptr;
            TRY(env->alloc(&p, tensor_bytes(env->tensors[slots[i]]), false));
            hp.stage[b][slots[i]] = p;
        }
    // step stream: actions in, step, snapshot (device to device: ~20 us for 34 MB) — the snapshot buffer was released
    // by the copy of two steps ago, which the caller has already waited for
    CU(cudaMemcpyAsync(env->tensors[B200ENV_T_ACTION].ptr, actions, env->action_count * 4, cudaMemcpyHostToDevice, hp.step_stream));
    TRY(env->launch(hp.step_stream, 0));
    for (int i = 0; i < num_slots; i++) {
        const b200env_tensor_desc &t = env->tensors[slots[i]];
        CU(cudaMemcpyAsync(hp.stage[b][slots[i]], t.ptr, tensor_bytes(t), cudaMemcpyDeviceToDevice, hp.step_stream));
    }
    CU(cudaEventRecord(hp.snapped[b], hp.step_stream));
    // copy stream: the snapshot crosses PCIe while the step stream runs the next step
    CU(cudaStreamWaitEvent(hp.copy_stream, hp.snapped[b], 0));
    for (int i = 0; i < num_slots; i++)
        CU(cudaMemcpyAsync(host_ptrs[i], hp.stage[b][slots[i]], tensor_bytes(env->tensors[slots[i]]), cudaMemcpyDeviceToHost,
                           hp.copy_stream));
    CU(cudaEventRecord(hp.copied[b], hp.copy_stream));
    hp.issued++;
    return B200ENV_OK;
}

int b200env_step_host_wait(b200env *env)
{
    if (!env) return fail(B200ENV_ERR_INVALID, "null simulator");
    if (!env->pipe || env->pipe->issued == env->pipe->waited) return fail(B200ENV_ERR_INVALID, "no step in flight");
    DeviceGuard g(env->gpu);
    b200env::HostPipe &hp = *env->pipe;
    CU(cudaEventSynchronize(hp.copied[hp.waited & 1u]));
    hp.waited++;
    if (hp.issued == hp.waited) CU(cudaStreamSynchronize(hp.step_stream));   // results visible to every stream again
    return B200ENV_OK;
}

// ---- benchmark / test helper: a stream gate the host opens ----
static __global__ void stream_gate_kernel(const volatile int32_t *flag, unsigned long long timeout_ns)
{
    unsigned long long t0, t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
    while (*flag == 0) {
        __nanosleep(500);
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        if (t - t0 > timeout_ns) break;
    }
}

int b200env_debug_stream_gate(int gpu_id, void *cuda_stream, const int32_t *host_flag, uint32_t timeout_ms)
{
    if (!host_flag) return fail(B200ENV_ERR_INVALID, "null flag");
    DeviceGuard g(gpu_id);
    if (!g.ok) return fail(B200ENV_ERR_CUDA, "cudaSetDevice(%d) failed", gpu_id);
    if (timeout_ms == 0 || timeout_ms > 5000) timeout_ms = 5000;
    void *dev_flag = nullptr;   // fails (instead of faulting on the device) when the flag is not pinned host memory
    CU(cudaHostGetDevicePointer(&dev_flag, const_cast<int32_t *>(host_flag), 0));
    stream_gate_kernel<<<1, 1, 0, (cudaStream_t)cuda_stream>>>((const volatile int32_t *)dev_flag,
                                                               (unsigned long long)timeout_ms * 1000000ull);
    CU(cudaGetLastError());
    return B200ENV_OK;
}

int b200env_overcooked_step_into(b200env *env, void *cuda_stream, const int32_t *actions, void *obs_out,
                                 int32_t *rewards_out, int32_t *done_out)
{
    if (!env || env->kind != B200ENV_OVERCOOKED) return fail(B200ENV_ERR_INVALID, "not an Overcooked simulator");
    NO_STEPS_IN_FLIGHT(env);
    DeviceGuard g(env->gpu);
    b200::oc::LaunchArgs a = env->oc;
    if (actions) a.actions = actions;
    if (obs_out) a.obs = (uint8_t *)obs_out;
    if (rewards_out) a.rewards = rewards_out;
    if (done_out) a.done = done_out;
    cudaError_t e = b200::oc::launch_step(env->layout, env->d_layout, a, 0, (cudaStream_t)cuda_stream);
    if (e != cudaSuccess) return fail(B200ENV_ERR_CUDA, "kernel launch failed: %s", cudaGetErrorString(e));
    return B200ENV_OK;
}

int b200env_overcooked_rollout(b200env *env, void *cuda_stream, int num_steps, const int32_t *actions,
                               int action_slices, void *obs_out, int32_t *rewards_out, int32_t *done_out,
                               int out_slices)
{
    if (!env || env->kind != B200ENV_OVERCOOKED) return fail(B200ENV_ERR_INVALID, "not an Overcooked simulator");
    if (num_steps < 0 || !actions || action_slices < 1 || out_slices < 1)
        return fail(B200ENV_ERR_INVALID, "rollout needs an action ring and at least one output slice");
    NO_STEPS_IN_FLIGHT(env);
    DeviceGuard g(env->gpu);
    const size_t PN = (size_t)env->P * env->N;
    const size_t obs_slice = PN * (size_t)env->layout.S * env->layout.C * (size_t)env->oc.obs_elem;   // bytes
    if (num_steps == 0) return B200ENV_OK;
    if (env->fused_rollout && !env->trace_buf && b200::oc::rollout_available(env->layout, env->oc)) {
        cudaError_t e = b200::oc::launch_rollout(env->layout, env->oc, num_steps, actions, action_slices,
                                                 (uint8_t *)obs_out, rewards_out, done_out, out_slices,
                                                 (cudaStream_t)cuda_stream);
        if (e != cudaSuccess) return fail(B200ENV_ERR_CUDA, "rollout launch failed: %s", cudaGetErrorString(e));
        return B200ENV_OK;
    }
    b200::oc::LaunchArgs a = env->oc;
    for (int t = 0; t < num_steps; t++) {
        const size_t o = (size_t)(t % out_slices);
        if (env->trace_buf) a.trace = (t < env->trace_steps) ? env->trace_buf + (size_t)t * env->trace_stride : nullptr;
        a.actions = actions + (size_t)(t % action_slices) * PN;
        if (obs_out) a.obs = (uint8_t *)obs_out + o * obs_slice;
        if (rewards_out) a.rewards = rewards_out + o * PN;
        if (done_out) a.done = done_out + o * env->N;
        cudaError_t e = b200::oc::launch_step(env->layout, env->d_layout, a, 0, (cudaStream_t)cuda_stream);
        if (e != cudaSuccess) return fail(B200ENV_ERR_CUDA, "kernel launch failed at step %d: %s", t, cudaGetErrorString(e));
    }
    return B200ENV_OK;
}

int b200env_random_rollout(b200env *env, void *cuda_stream, int num_steps, uint32_t seed, uint32_t first_step)
{
    if (!env) return fail(B200ENV_ERR_INVALID, "null simulator");
    if (num_steps < 0) return fail(B200ENV_ERR_INVALID, "negative step count");
    NO_STEPS_IN_FLIGHT(env);
    DeviceGuard g(env->gpu);
    cudaStream_t st = (cudaStream_t)cuda_stream;
    int32_t *actions = (int32_t *)env->tensors[B200ENV_T_ACTION].ptr;
    const int32_t *mask = nullptr;
    int M = 2;   // cartpole: push left / right (src/cartpole_env/sim.cpp:70)
    if (env->kind == B200ENV_OVERCOOKED) M = 6;          // all six moves always legal (sim.cpp:616-618): skip the mask
    else if (env->kind == B200ENV_BALANCE) M = 4;        // always legal (src/balance_beam_env/sim.cpp:196-198)
    else if (env->kind == B200ENV_ACROBAT) M = 3;        // torque -1 / 0 / +1 (src/acrobat_env/sim.cpp:179)
    else if (env->kind == B200ENV_HANABI) { M = 20; mask = (const int32_t *)env->tensors[B200ENV_T_ACTION_MASK].ptr; }
    // Hanabi: the step kernel itself draws the next step's actions from the legal moves it has just computed, so only
    // the first step of the rollout needs the sampler kernel
    const bool in_kernel = env->kind == B200ENV_HANABI;
    // Cartpole / Acrobat / Balance Beam: the step kernel draws THIS step's action itself (no sampler launch at all)
    const bool self_sampling = env->kind == B200ENV_CARTPOLE || env->kind == B200ENV_ACROBAT || env->kind == B200ENV_BALANCE;
    int rc = B200ENV_OK;
    // (a debug trace with 16 stamps per tile is the fused kernel's; the 8-stamp one is the per-step kernel's)
    const bool fused_trace = env->trace_buf && env->trace_stride == 16 * (size_t)b200::hanabi::num_tiles(env->N);
    if (in_kernel && env->fused_rollout && (!env->trace_buf || fused_trace) && num_steps > 0) {
        // one cooperative launch for all steps (game records stay on chip, the game logic runs ahead of the row writes)
        cudaError_t e = b200::policy::launch_sample(actions, mask, (uint32_t)env->action_count, M, seed, first_step, st);
        if (e != cudaSuccess) return fail(B200ENV_ERR_CUDA, "sampler launch failed: %s", cudaGetErrorString(e));
        env->hb.trace = fused_trace ? env->trace_buf : nullptr;
        env->hb.trace_steps = fused_trace ? env->trace_steps : 0;
        e = b200::hanabi::launch_rollout(env->hb, num_steps, seed, first_step, st);
        env->hb.trace = nullptr;
        if (e == cudaSuccess) return B200ENV_OK;
        if (e != cudaErrorCooperativeLaunchTooLarge)
            return fail(B200ENV_ERR_CUDA, "rollout launch failed: %s", cudaGetErrorString(e));
        cudaGetLastError();   // too many tiles to be resident at once: step by step below
    }
    if (env->kind == B200ENV_OVERCOOKED && env->fused_rollout && num_steps > 0 && !env->trace_buf &&
        b200::oc::rollout_available(env->layout, env->oc)) {
        // the fused T-step kernel with the draw made in the kernel: no sampler launches, no action tensor traffic
        cudaError_t e = b200::oc::launch_rollout(env->layout, env->oc, num_steps, nullptr, 1, nullptr, nullptr, nullptr, 1,
                                                 st, seed, first_step);
        if (e != cudaSuccess) return fail(B200ENV_ERR_CUDA, "rollout launch failed: %s", cudaGetErrorString(e));
        return B200ENV_OK;
    }
    if ((env->kind == B200ENV_CARTPOLE || env->kind == B200ENV_BALANCE) && env->fused_rollout && num_steps > 0) {
        // one cooperative launch for all steps: the worlds stay in registers, only the exported tensors move
        cudaError_t e = env->kind == B200ENV_CARTPOLE ? b200::cartpole::launch_rollout(env->cp, num_steps, seed, first_step, st)
                                                      : b200::balance::launch_rollout(env->bb, num_steps, seed, first_step, st);
        if (e == cudaSuccess) return B200ENV_OK;
        if (e != cudaErrorCooperativeLaunchTooLarge)
            return fail(B200ENV_ERR_CUDA, "rollout launch failed: %s", cudaGetErrorString(e));
        cudaGetLastError();   // too many tiles to be resident at once: step by step below
    }
    for (int t = 0; t < num_steps && rc == B200ENV_OK; t++) {
        if (self_sampling) {
            env->cp.sample = env->ac.sample = env->bb.sample = 1;
            env->cp.sample_seed = env->ac.sample_seed = env->bb.sample_seed = seed;
            env->cp.sample_step = env->ac.sample_step = env->bb.sample_step = first_step + (uint32_t)t;
        } else if (t == 0 || !in_kernel) {
            cudaError_t e = b200::policy::launch_sample(actions, mask, (uint32_t)env->action_count, M, seed,
                                                        first_step + (uint32_t)t, st);
            if (e != cudaSuccess) return fail(B200ENV_ERR_CUDA, "sampler launch failed: %s", cudaGetErrorString(e));
        }
        if (in_kernel) {
            env->hb.sample_next = 1;
            env->hb.sample_seed = seed;
            env->hb.sample_step = first_step + (uint32_t)t + 1u;
            env->hb.trace = (env->trace_buf && t < env->trace_steps) ? env->trace_buf + (size_t)t * env->trace_stride : nullptr;
        }
        rc = env->launch(st, 0);
    }
    env->hb.sample_next = 0;
    env->hb.trace = nullptr;
    env->cp.sample = env->ac.sample = env->bb.sample = 0;
    return rc;
}

uint32_t b200env_random_action_hash(uint32_t seed, uint32_t step, uint32_t index)
{
    return b200::policy::host_mix(seed, step, index);
}

int b200env_set_rollout_fused(b200env *env, int fused)
{
    if (!env) return fail(B200ENV_ERR_INVALID, "null simulator");
    env->fused_rollout = fused != 0;
    return B200ENV_OK;
}

int b200env_hanabi_set_rollout_fused(b200env *env, int fused)
{
    if (!env || env->kind != B200ENV_HANABI) return fail(B200ENV_ERR_INVALID, "not a Hanabi simulator");
    env->fused_rollout = fused != 0;
    return B200ENV_OK;
}

int b200env_overcooked_set_rollout_fused(b200env *env, int fused)
{
    if (!env || env->kind != B200ENV_OVERCOOKED) return fail(B200ENV_ERR_INVALID, "not an Overcooked simulator");
    env->fused_rollout = fused != 0;
    return B200ENV_OK;
}

int b200env_debug_trace(b200env *env, void *device_buf, size_t u64_per_step, int steps)
{
    if (!env || (env->kind != B200ENV_OVERCOOKED && env->kind != B200ENV_HANABI))
        return fail(B200ENV_ERR_INVALID, "no timeline for this simulator");
    env->trace_buf = (unsigned long long *)device_buf;
    env->trace_stride = u64_per_step;
    env->trace_steps = device_buf ? steps : 0;
    return B200ENV_OK;
}

int b200env_overcooked_force_kernel(b200env *env, int which)
{
    if (!env || env->kind != B200ENV_OVERCOOKED) return fail(B200ENV_ERR_INVALID, "not an Overcooked simulator");
    if (which != -1 && which != 0 && which != 1) return fail(B200ENV_ERR_INVALID, "kernel selector must be -1, 0 or 1");
    if (which == 0 && env->layout.rules)
        return fail(B200ENV_ERR_NOSLOT, "the Simplecooked rules run on the warp-tile kernel only");
    if (which == 1 && !env->oc.tile_const[b200::oc::lanes_index(env->oc.lanes)])
        return fail(B200ENV_ERR_NOSLOT, "the warp-tile kernel cannot run this layout (more than 8 players or a staging "
                                        "tile that does not fit in shared memory)");
    env->oc.force_kernel = which;
    return B200ENV_OK;
}

int b200env_overcooked_set_lanes(b200env *env, int lanes)
{
    if (!env || env->kind != B200ENV_OVERCOOKED) return fail(B200ENV_ERR_INVALID, "not an Overcooked simulator");
    if (lanes == 0) lanes = b200::oc::pick_lanes(env->layout, env->N, env->oc.obs_elem);
    if (lanes != 1 && lanes != 2 && lanes != 4 && lanes != 8)
        return fail(B200ENV_ERR_INVALID, "lanes per world must be 0 (auto), 1, 2, 4 or 8");
    if (!env->oc.tile_const[b200::oc::lanes_index(lanes)])
        return fail(B200ENV_ERR_NOSLOT, "%d lanes per world: fewer lanes than players, or the staging tile does not fit "
                                        "in shared memory", lanes);
    env->oc.lanes = lanes;
    return B200ENV_OK;
}

size_t b200env_state_bytes(const b200env *env)
{
    size_t n = 0;
    if (env) for (const Region &r : env->state) n += r.bytes;
    return n;
}

int b200env_get_state(b200env *env, void *host_buf, size_t bytes)
{
    if (!env || !host_buf) return fail(B200ENV_ERR_INVALID, "null argument");
    if (bytes != b200env_state_bytes(env)) return fail(B200ENV_ERR_INVALID, "state blob must be %zu bytes", b200env_state_bytes(env));
    DeviceGuard g(env->gpu);
    CU(cudaDeviceSynchronize());
    uint8_t *dst = (uint8_t *)host_buf;
    for (const Region &r : env->state) {
        CU(cudaMemcpy(dst, r.ptr, r.bytes, cudaMemcpyDeviceToHost));
        dst += r.bytes;
    }
    return B200ENV_OK;
}

int b200env_set_state(b200env *env, const void *host_buf, size_t bytes)
{
    if (!env || !host_buf) return fail(B200ENV_ERR_INVALID, "null argument");
    if (bytes != b200env_state_bytes(env)) return fail(B200ENV_ERR_INVALID, "state blob must be %zu bytes", b200env_state_bytes(env));
    DeviceGuard g(env->gpu);
    CU(cudaDeviceSynchronize());
    const uint8_t *src = (const uint8_t *)host_buf;
    for (const Region &r : env->state) {
        CU(cudaMemcpy(r.ptr, src, r.bytes, cudaMemcpyHostToDevice));
        src += r.bytes;
    }
    return B200ENV_OK;
}

int b200env_episode_stats(b200env *env, int64_t out[2])
{
    if (!env || !out) return fail(B200ENV_ERR_INVALID, "null argument");
    out[0] = out[1] = 0;
    DeviceGuard g(env->gpu);
    if (env->kind == B200ENV_OVERCOOKED) {
        CU(cudaMemcpy(out, env->oc.stats, 16, cudaMemcpyDeviceToHost));
        return B200ENV_OK;
    }
    // the environments that seed every episode from one counter (init.hpp EpisodeManager::curEpisode): episodes started
    const uint32_t *counter = env->kind == B200ENV_HANABI ? env->hb.cur_episode : env->kind == B200ENV_BALANCE ? env->bb.cur_episode
                              : env->kind == B200ENV_CARTPOLE ? env->cp.cur_episode : nullptr;
    if (counter) {
        uint32_t v = 0;
        CU(cudaMemcpy(&v, counter, 4, cudaMemcpyDeviceToHost));
        out[0] = (int64_t)v;
    }
    return B200ENV_OK;
}

}   // extern "C"
