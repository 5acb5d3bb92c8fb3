/* b200envs — C ABI of the B200-native batched environment stepper.
 *
 * Drop-in boundary for the env step path of willwng/madrona_rl_envs_playground: these entry points are what the
 * reference's per-environment nanobind modules bind (src/<env>/bindings.cpp -> Manager in src/<env>/mgr.{hpp,cpp}):
 *
 *   reference interface                                              replaced by
 *   ---------------------------------------------------------------  ---------------------------------------------
 *   Manager::Manager(Config)     overcooked mgr.cpp:138-192,          b200env_overcooked_create / _hanabi_create /
 *                                bindings.cpp:15-53; hanabi           _balance_create / _cartpole_create /
 *                                                                     _acrobat_create
 *                                mgr.cpp:124-170; balance / cartpole
 *   Manager::step()              mgr.cpp:196-199 (all envs)           b200env_step (synchronous, like the reference's
 *                                                                     cuGraphLaunch + cudaStreamSynchronize,
 *                                                                     madrona/src/mw/cuda_exec.cpp:1781-1786)
 *   Manager::*Tensor()           overcooked mgr.cpp:201-259,          b200env_tensor(slot)
 *                                hanabi mgr.cpp:172-234, ...
 *   py::Tensor -> DLPack view    madrona/src/python/bindings.cpp:     the (ptr, dtype, dims, device) record returned by
 *                                109-124                              b200env_tensor; the library owns the memory for
 *                                                                     the simulator's lifetime, pointers never move
 *   Manager::~Manager()                                               b200env_destroy
 *
 * Plain pointers and sizes only; no torch / C++ types.  Every function returns B200ENV_OK (0) or a negative error
 * code and never aborts (the reference FATALs); b200env_last_error() returns the message of the calling thread's last
 * failure.  A simulator is bound to one GPU and must be driven from one host thread at a time.
 *
 * Tensor layouts are agent-major and dense.  The Overcooked observation tensor is
 * int8 [P][N][W][H][5P+16] — the per-player (N, W, H, C) arrays that envs/overcooked_env.py:92-102 produces after
 * its scatter / reshape / transpose — so Python needs no indexing pass.  world_id / agent_id tensors are provided so
 * that the reference's generic scatter-by-ID wrapper (pantheonrl_extension/vectorenv.py:262-346) keeps working.
 */
#ifndef B200ENVS_H
#define B200ENVS_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct b200env b200env;

enum {
    B200ENV_OK = 0,
    B200ENV_ERR_INVALID = -1,   /* bad argument / unsupported configuration */
    B200ENV_ERR_CUDA = -2,      /* a CUDA call failed (message in b200env_last_error) */
    B200ENV_ERR_NOSLOT = -3     /* the environment does not export that tensor */
};

enum { B200ENV_OVERCOOKED = 0, B200ENV_HANABI = 1, B200ENV_BALANCE = 2, B200ENV_CARTPOLE = 3, B200ENV_ACROBAT = 4 };

enum { B200ENV_I8 = 0, B200ENV_I32 = 1, B200ENV_F32 = 2, B200ENV_BF16 = 3 };

/* Tensor slots: union of the reference's ExportID enums (src/overcooked_env/sim.hpp:23-38, src/hanabi_env/sim.hpp:34-45,
 * src/balance_beam_env/sim.hpp:24-34, src/cartpole_env/sim.hpp:17-24). */
enum {
    B200ENV_T_DONE = 0,        /* i32 [N]            (cartpole: [N,1], its reset_tensor) */
    B200ENV_T_ACTIVE_AGENT,    /* i32 [P,N] */
    B200ENV_T_ACTION,          /* i32 [P,N,1]        (cartpole: [N,1]) — the caller writes actions here */
    B200ENV_T_OBSERVATION,     /* overcooked i8 (or bf16 / f32, b200env_overcooked_create_typed) [P,N,W,H,C];
                                  hanabi i8 [2,N,658]; balance i32 [2,N,7] */
    B200ENV_T_ACTION_MASK,     /* i32 [P,N,moves] */
    B200ENV_T_REWARD,          /* overcooked i32 [P,N]; hanabi / balance f32 [2,N]; cartpole f32 [N,1] */
    B200ENV_T_WORLD_ID,        /* i32 [P,N] = n      (cartpole: [N,1]) */
    B200ENV_T_AGENT_ID,        /* i32 [P,N] = p */
    B200ENV_T_AGENT_STATE,     /* hanabi i8 [2,N,783]; cartpole f32 [N,4] (its state_tensor) */
    /* Reference-raw Overcooked export (compat; allocated on first request, see b200env_overcooked_export_raw): */
    B200ENV_T_LOCATION_WORLD_ID, /* i32 [P*S,N]: Manager::locationWorldIDTensor, src/overcooked_env/mgr.cpp:249-253 */
    B200ENV_T_LOCATION_ID,       /* i32 [P*S,N]: Manager::locationIDTensor, mgr.cpp:255-259 */
    B200ENV_T_OBSERVATION_RAW,   /* i8  [P*S,N,336]: Manager::observationTensor's padded rows, mgr.cpp:219-223 */
    B200ENV_T_NUM_SLOTS
};

typedef struct {
    void *ptr;          /* device pointer, valid until b200env_destroy */
    int32_t dtype;      /* B200ENV_I8 / I32 / F32 / BF16 */
    int32_t ndim;
    int64_t dims[6];    /* C-contiguous */
    int32_t device;     /* CUDA device ordinal */
} b200env_tensor_desc;

/* Mirrors Overcooked::Manager::Config (src/overcooked_env/mgr.hpp:16-34); vectors are (pointer, implied length):
 * terrain has height*width entries (TerrainT values, row-major), start_player_x/y have num_players entries,
 * recipe_values / recipe_times have 16 entries indexed 4*onions+tomatoes. */
typedef struct {
    const int64_t *terrain;
    int64_t height, width, num_players;
    const int64_t *start_player_x, *start_player_y;
    int64_t placement_in_pot_rew, dish_pickup_rew, soup_pickup_rew;
    const int64_t *recipe_values, *recipe_times;
    int64_t horizon;
} b200env_overcooked_config;

/* Mirrors Hanabi::Manager::Config (src/hanabi_env/mgr.hpp). Only players == 2 is supported, like the reference's
 * exported tensor shapes (src/hanabi_env/mgr.cpp:172-234 hard-code 2). */
typedef struct {
    uint32_t colors, ranks, players, max_information_tokens, max_life_tokens;
} b200env_hanabi_config;

int b200env_overcooked_create(const b200env_overcooked_config *cfg, int gpu_id, uint32_t num_worlds, b200env **out);
/* src/overcooked2_env ("Simplecooked", the JS-compatible Overcooked the reference's trainer imports): replaces
 * Simplecooked::Manager::Manager(Config) (src/overcooked2_env/mgr.cpp:138-192, bindings.cpp:10-59).  Same Config; terrain
 * values in ITS enum order AIR, POT, COUNTER, ONION_SOURCE, DISH_SOURCE, SERVING, TOMATO_SOURCE (sim.hpp:40); at most 2
 * players and 100 cells (sim.hpp:13-14); observation int8 [P][N][W][H][5P+10].  The simulator is of kind
 * B200ENV_OVERCOOKED: every b200env_overcooked_* call works on it (CTA-per-world kernel excepted). */
int b200env_overcooked2_create(const b200env_overcooked_config *cfg, int gpu_id, uint32_t num_worlds, b200env **out);
/* Policy-ready observations (the step after the path: every consumer in the reference starts with `obs.float()`,
 * pantheonrl_extension/vectoragent.py:353, centralized_agent.py:257, and CleanRLNetwork flattens that to
 * prod(obs_shape) floats, vectoragent.py:66-87): the same simulator with the observation tensor — and every obs_out
 * buffer of the calls below — holding bfloat16 (B200ENV_BF16) or float (B200ENV_F32) instead of int8, element for
 * element the value `.float()` of the int8 tensor would give (bfloat16 represents every int8 value exactly), written
 * by the step kernels themselves: the int8 tensor, the cast kernel and its 1 + 4 bytes per element of extra HBM
 * traffic disappear, and a bf16 policy reads 2 bytes per element instead of 4.  rules: 0 = src/overcooked_env,
 * 1 = src/overcooked2_env (rules 1 only on layouts the warp-tile kernels run: at most 8 players, staging tile
 * <= 56 KB).  Many-agent / large-grid layouts go through the CTA-per-world kernel, which needs one observer's slab
 * (cells x channels x element size) in shared memory: 255 cells x 64 players x float does not fit and returns
 * B200ENV_ERR_NOSLOT.  obs_dtype = B200ENV_I8 is b200env_overcooked_create / _overcooked2_create. */
int b200env_overcooked_create_typed(const b200env_overcooked_config *cfg, int rules, int obs_dtype, int gpu_id,
                                    uint32_t num_worlds, b200env **out);
/* 2 players (the reference's exported shapes hard-code them, src/hanabi_env/mgr.cpp:172-234), 1-5 colours, 2-5 ranks with
 * a deck (2 * ranks * colors cards) of at least the 10 cards of the deal — the reference deals from an empty deck
 * otherwise (sim.cpp:45-52, 508-519) —, 1-8 information and 1-3 life tokens; anything else is B200ENV_ERR_INVALID.
 * With 256 worlds per SM or more the first create on a device measures every SM's store rate (~2 ms, device
 * synchronised; the fused random rollout deals its tiles of worlds to the SMs in that proportion).  Environment
 * variable B200ENV_HANABI_NO_BALANCE=1 keeps equal shares. */
int b200env_hanabi_create(const b200env_hanabi_config *cfg, int gpu_id, uint32_t num_worlds, b200env **out);
/* Policy-ready Hanabi rows, as b200env_overcooked_create_typed: the observation (658) and agent-state (783) tensors hold
 * bfloat16 (B200ENV_BF16) or float (B200ENV_F32) elements with the value `.float()` of the int8 tensor would give — what
 * every consumer computes first (pantheonrl_extension/hanabi_agent.py:257,292) — written by the step and fused-rollout
 * kernels themselves from the same bit rows.  Masks, flags, rewards keep their types.  obs_dtype = B200ENV_I8 is
 * b200env_hanabi_create. */
int b200env_hanabi_create_typed(const b200env_hanabi_config *cfg, int obs_dtype, int gpu_id, uint32_t num_worlds,
                                b200env **out);
int b200env_balance_create(int gpu_id, uint32_t num_worlds, b200env **out);
int b200env_cartpole_create(int gpu_id, uint32_t num_worlds, b200env **out);
/* src/acrobat_env (two-link pendulum, RK4): replaces Acrobat::Manager::Manager(Config) (src/acrobat_env/mgr.cpp:139-170,
 * bindings.cpp:10-33).  Tensors as Cartpole's (mgr.cpp:172-201): DONE i32 [N,1] (reset_tensor), ACTION i32 [N,1] in 0..2
 * (torque -1 / 0 / +1), AGENT_STATE f32 [N,4] = theta1, theta2, omega1, omega2 (observation_tensor), REWARD f32 [N,1],
 * WORLD_ID.  The reference keeps ONE episode-length counter for all worlds of a Manager (init.hpp:9); the kernel
 * reproduces its one-worker order exactly (see csrc/simple_envs.cu). */
int b200env_acrobat_create(int gpu_id, uint32_t num_worlds, b200env **out);
void b200env_destroy(b200env *env);

int b200env_kind(const b200env *env);
uint32_t b200env_num_worlds(const b200env *env);

/* One step for every world: reads the action tensor, updates the worlds, rewrites the exported tensors.
 * b200env_step launches on the legacy default stream and returns after completion (Manager::step semantics), so
 * results are visible to any stream.  b200env_step_async only enqueues on `cuda_stream` (a cudaStream_t; NULL = legacy
 * default stream) — the caller orders its own reads. */
int b200env_step(b200env *env);
int b200env_step_async(b200env *env, void *cuda_stream);

int b200env_tensor(const b200env *env, int slot, b200env_tensor_desc *out);

/* Host-buffer step (the end-to-end path): copies `actions` (int32, same element order as the action tensor) host->
 * device, steps, and copies each requested tensor device->host into host_ptrs[i] (dense, tensor order).  Returns when
 * all copies are complete.  Pinned host memory gives full-speed copies; pageable memory works too. */
int b200env_step_host(b200env *env, const int32_t *actions, const int *slots, void *const *host_ptrs, int num_slots);

/* Pipelined host-buffer steps, for callers whose next actions do not depend on this step's outputs being on the host
 * yet (random / scripted / replayed action streams, or a policy that lags one step): _async enqueues {actions H2D,
 * step, device-side snapshot of the requested tensors} on the simulator's own step stream and the D2H copies of the
 * snapshot on its copy stream, and returns at once; the NEXT step then runs while this step's outputs are still
 * crossing PCIe.  At most two steps may be in flight.  _wait blocks until the OLDEST step in flight has landed in its
 * host buffers (returns B200ENV_ERR_INVALID when none is).  `actions` and the host buffers of a step must stay untouched
 * until its _wait; use pinned memory (pageable memory makes every copy synchronous).  While steps are in flight the
 * other step entry points refuse to run.  Outputs are bit-for-bit those of b200env_step_host. */
int b200env_step_host_async(b200env *env, const int32_t *actions, const int *slots, void *const *host_ptrs, int num_slots);
int b200env_step_host_wait(b200env *env);

/* Overcooked only (compat): writes the observation in the REFERENCE's exported layout — padded 336-byte rows
 * (sizeof(LocationXObservation), src/overcooked_env/sim.hpp:125-127) in world-major (world, cell, player) memory order
 * under the declared shape [P*S, N, 336] (mgr.cpp:219-223; madrona/src/core/state.cpp:511-554 concatenates the worlds'
 * tables) — from the native [P,N,W,H,C] tensor, into `dst` (device, NULL = the simulator's own
 * B200ENV_T_OBSERVATION_RAW tensor, allocated on first use).  Bytes [C, 336) of a row are zero (uninitialised in the
 * reference, sim.cpp:638-640).  With B200ENV_T_LOCATION_ID / _LOCATION_WORLD_ID (flat row k = world*S*P + cell*P +
 * player holds player*S + cell and world) the unmodified wrapper's scatter (envs/overcooked_env.py:92-102) rebuilds the
 * native tensor.  Not a hot path: one extra pass over 13x the bytes, which is exactly what the native layout avoids.
 * int8 observations only. */
int b200env_overcooked_export_raw(b200env *env, void *cuda_stream, void *dst);

/* Benchmark / test helper: enqueues on `cuda_stream` of device `gpu_id` a one-thread kernel that spins until
 * *host_flag (int32 in pinned, device-visible host memory) becomes non-zero, or `timeout_ms` have passed.  Work enqueued
 * behind it starts the moment the host sets the flag, so a timed region holds no host enqueue latency.  Enqueue only
 * kernels that have run before in this process behind a closed gate: the first launch of a kernel loads its module
 * (CUDA lazy loading), which waits for running kernels and would block the host until the gate times out. */
int b200env_debug_stream_gate(int gpu_id, void *cuda_stream, const int32_t *host_flag, uint32_t timeout_ms);

/* Overcooked only: one asynchronous step with caller-supplied device buffers (any may be NULL = the simulator's own
 * tensor): actions i32 [P,N], obs_out [P,N,W,H,C] of the observation tensor's dtype, rewards_out i32 [P,N], done_out i32 [N].  Lets a rollout
 * collector write step t straight into slice t of its [T,...] buffers (what CleanPPOAgent allocates,
 * pantheonrl_extension/vectoragent.py:172-180). */
int b200env_overcooked_step_into(b200env *env, void *cuda_stream, const int32_t *actions, void *obs_out,
                                 int32_t *rewards_out, int32_t *done_out);

/* Overcooked only: `num_steps` asynchronous steps enqueued back to back on `cuda_stream` — the rollout collector's
 * loop (scripts/overcooked_example.py:99-116 with the actions known in advance) without a host round trip per step.
 * Step t reads actions[t % action_slices] (i32 [action_slices][P][N]) and writes slice t % out_slices of
 * obs_out ([out_slices][P][N][W][H][C], the observation dtype), rewards_out (i32 [out_slices][P][N]) and done_out
 * (i32 [out_slices][N]); a NULL output = the simulator's own tensor.
 * Layouts the warp-tile kernel runs (<= 8 players, small grids) take ONE fused launch that keeps the world state on
 * chip for all num_steps steps; others (and b200env_overcooked_set_rollout_fused(env, 0)) take one launch per step.
 * Results are identical either way. */
int b200env_overcooked_rollout(b200env *env, void *cuda_stream, int num_steps, const int32_t *actions,
                               int action_slices, void *obs_out, int32_t *rewards_out, int32_t *done_out,
                               int out_slices);

int b200env_overcooked_set_rollout_fused(b200env *env, int fused);

/* Any environment: `num_steps` x {sample a uniformly random LEGAL action for every agent into the action tensor, step},
 * enqueued on `cuda_stream` with no host round trip — the random agents of the reference's example scripts
 * (scripts/overcooked_example.py:99-105, scripts/hanabi_example.py:64-67) as a device-only loop.  Counter-based:
 * flat action element i at step s picks legal move number floor(b200env_random_action_hash(seed, s, i) * legal / 2^32)
 * (the inactive Hanabi player has no legal move and gets 0, which the step ignores).  Step numbers run from first_step.
 * Hanabi: after the first step the draw is made inside the step kernel (only the player to move gets an action), and
 * by default all num_steps steps run in ONE cooperative launch that keeps the game records on chip and lets the game
 * logic run ahead of the observation writes; b200env_hanabi_set_rollout_fused(env, 0) — or a batch too large for all
 * its tiles to be resident at once — takes one launch per step.  Tensors after the call are identical either way. */
int b200env_random_rollout(b200env *env, void *cuda_stream, int num_steps, uint32_t seed, uint32_t first_step);
int b200env_hanabi_set_rollout_fused(b200env *env, int fused);
/* Any environment: fused multi-step launches of b200env_random_rollout / b200env_overcooked_rollout on (default) or off
 * (one launch per step).  Cartpole and Balance Beam also run b200env_random_rollout as ONE cooperative launch with the
 * worlds in registers; Cartpole, Acrobat and Balance Beam step kernels draw the random action themselves either way
 * (no sampler launch).  Tensors after the call are identical in every mode. */
int b200env_set_rollout_fused(b200env *env, int fused);
uint32_t b200env_random_action_hash(uint32_t seed, uint32_t step, uint32_t index);

/* Overcooked only: choose the step kernel (-1 auto, 0 CTA-per-world, 1 warp-tile). Tests. */
int b200env_overcooked_force_kernel(b200env *env, int which);

/* Overcooked only: lanes per world of the warp-tile kernel (0 auto, 1, 2, 4 or 8; at least one per
 * player).  Results are identical for every value; extra lanes only put more warps on each SM (tuning / tests). */
int b200env_overcooked_set_lanes(b200env *env, int lanes);

/* Debug (Overcooked warp-tile kernel): while set, step t < steps of each b200env_overcooked_rollout call makes every
 * CTA write 8 %globaltimer stamps (entry, dependency resolved, inputs landed, state written, first observer queued,
 * first observer read, exit) at device_buf[t * u64_per_step + 8 * cta].  NULL switches it off.
 * Hanabi: the same for step t < steps of each b200env_random_rollout call, indexed by tile (32 worlds): ticket taken,
 * record loaded, reset count published, bit rows composed, rows written, episode numbers known, resets done, exit.
 * With u64_per_step = 16 * tiles the fused Hanabi rollout is traced instead: 16 stamps per (step, tile) for the first
 * `steps` steps of each call (game warp 0-6, writing warp 8-11; see RolloutTrace in csrc/hanabi.cu). */
int b200env_debug_trace(b200env *env, void *device_buf, size_t u64_per_step, int steps);

/* Checkpoint: size of / copy out / copy in the complete simulator state (hidden state + exported tensors + episode
 * counter) as one opaque host blob; blobs are only valid for a simulator created with the same configuration. */
size_t b200env_state_bytes(const b200env *env);
int b200env_get_state(b200env *env, void *host_buf, size_t bytes);
int b200env_set_state(b200env *env, const void *host_buf, size_t bytes);

/* Per-GPU episode statistics accumulated on the device since creation.  Overcooked: out[0] = finished episodes,
 * out[1] = sum of their returns.  Hanabi, Balance Beam, Cartpole: out[0] = episodes STARTED (the Manager-wide counter the
 * reference seeds every episode from, init.hpp EpisodeManager::curEpisode; the initial reset of every world included),
 * out[1] = 0.  Synchronises the device.  Multi-GPU jobs all-reduce these 16 bytes themselves (NCCL), off the step path. */
int b200env_episode_stats(b200env *env, int64_t out[2]);

/* Cartpole step that ALLOCATES NOTHING.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid
 * it).  One step of `num_worlds` Cartpole worlds (the same arithmetic as b200env_step on a Cartpole simulator) on
 * caller-owned DEVICE buffers of GPU `gpu_id`, enqueued on `cuda_stream` (NULL = legacy default stream), no host
 * synchronisation: state f32 [N][4] in/out (x, x_dot, theta, theta_dot; worlds that finish are reset in place), actions
 * i32 [N] in (0 / 1), reward f32 [N] out, done i32 [N] out, episode_counter u32 [1] in/out — the Manager-wide episode
 * counter that seeds resets (src/cartpole_env/sim.cpp:51-53); a batch whose N worlds have each been reset once holds N.
 * The scratch of the ordered episode numbering lives in a file-scope static device array (32 KB, held by every GPU that
 * loads the library until exit): num_worlds is limited to b200env_cartpole_step_noalloc_max_worlds() = 1 048 576 and
 * larger batches return B200ENV_ERR_INVALID.  Every call shares that array and clears the part it uses first, so
 * results do not depend on earlier calls — but calls that OVERLAP on the GPU (different streams, threads or
 * processes' contexts on one device are separate; same context is not) race on it: serialise them.
 * Calls no cudaMalloc / cudaFree / allocating library routine. */
int b200env_cartpole_step_noalloc(int gpu_id, void *cuda_stream, uint32_t num_worlds, float *state,
                                  const int32_t *actions, float *reward, int32_t *done, uint32_t *episode_counter);
uint32_t b200env_cartpole_step_noalloc_max_worlds(void);

const char *b200env_last_error(void);
const char *b200env_version(void);

#ifdef __cplusplus
}
#endif
#endif
