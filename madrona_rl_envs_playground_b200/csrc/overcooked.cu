// Overcooked step kernels (sm_100a).  See overcooked.cuh for the per-world logic and DESIGN.md for the layout.
//
// HBM layout (structure of arrays over worlds, N = worlds on this GPU):
//   players  uint2 [P][N]   .x = held Object (packed), .y = position | orientation << 8
//   objs     u32   [K][N]   Object on each object-capable cell (counters and pots only)
//   timestep i32   [N]
//   actions  i32   [P][N]   (exported, written by the caller)      rewards i32 [P][N]     done i32 [N]
//   obs      u8    [P][N][W][H][C]   — exactly the tensor envs/overcooked_env.py:96 hands to the policy, so the
//                                      Python wrapper needs no gather / scatter / transpose pass.
//
// Three kernels:
//   step_tile<G>     one step; one warp = one tile of 32/G consecutive worlds, lane = (world, agent).  Agent state in
//                    registers, object slots in shared memory; cross-agent rules resolved with shuffles / ballots.
//                    The observation tile is a TMA-loaded template + scattered dynamic bytes in shared memory, handed
//                    to the copy engine (cp.async.bulk) once per observer.  P <= 8, small grids.
//   rollout_tile<G>  T steps in one launch with the worlds kept on chip (registers + shared memory) and the staging
//                    tile updated incrementally.
//   step_world_cta   one CTA = one world, for many-agent / large-grid layouts (P up to 64, S up to 255).  The
//                    observer-independent slab is built once per world in two shared buffers; per observer only the
//                    player bytes are patched before the slab (S*C contiguous bytes) is bulk-stored.
#include <cstdlib>
#include <cstring>

#include "overcooked.cuh"
#include "overcooked_host.h"
#include "random_policy_host.h"

namespace b200 {
namespace oc {

template <int WPW>
struct SmemObjsT {   // object slots of one world inside a [K][WPW] shared tile
    uint32_t *base;
    __device__ __forceinline__ uint32_t &operator[](int k) { return base[k * WPW]; }
};

struct Buffers {
    uint2 *players;
    uint32_t *objs;
    int32_t *timestep;
    const int32_t *actions;
    int32_t *rewards;
    int32_t *done;
    uint8_t *obs;
    // episode statistics (optional, off the parity path): finished episodes, sum of episode returns
    int32_t *ep_return;              // [N] running return of the current episode
    unsigned long long *stats;       // [2]
    uint32_t N;
    int32_t P;
    unsigned long long *trace;       // debug: 8 globaltimer stamps per CTA (nullptr = off)
};

__device__ __forceinline__ void trace_stamp(const Buffers &B, int slot)
{
    if (B.trace && threadIdx.x == 0) {
        unsigned long long t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        B.trace[(size_t)blockIdx.x * 8 + slot] = t;
    }
}

// Constant block of the warp-tile kernel, one per (simulator, G) in global memory: [Layout][template tile].  The
// template is the observation tile of 32/G worlds ([32/G][W][H][C] bytes) with only the static terrain one-hot set
// (sim.cpp:642-645); every CTA pulls the whole block into shared memory with ONE bulk copy (TMA engine) and then
// only scatters the dynamic bytes on top.
struct TileDims {
    uint32_t SC;            // bytes of one (observer, world) slab = S * C * sizeof(element)
    uint32_t stage_bytes;   // (32/G) * SC rounded up to 16
    uint32_t const_bytes;   // sizeof(Layout) + stage_bytes
    uint32_t const_stride;  // distance between the TILE_CONST_COPIES copies of the block in global memory
    int32_t K;              // object slots
};

// Counter / pot interaction of one agent (do_counter_pot_interaction, sim.cpp:286-338): `cell` is the Object on the
// faced cell, `h` the held Object; returns the shaping reward.
__device__ __forceinline__ int counter_pot_interaction(const Layout &L, int terrain, uint32_t &cell, uint32_t &h)
{
    const uint32_t c = cell;
    int rew = 0;
    if (terrain == COUNTER) {
        if (obj_name(h) != NONE && obj_name(c) == NONE) { cell = h; h = OBJ_NONE; }
        else if (obj_name(h) == NONE && obj_name(c) != NONE) { h = c; cell = OBJ_NONE; }
    } else {   // POT
        if (obj_name(h) == NONE) {
            // soup_to_be_cooked_at_location, sim.cpp:175-178: an idle soup (tick < 0) with >= 1 ingredient starts cooking
            if (obj_name(c) == SOUP && obj_tick(c) < 0 && ((obj_onions(c) + obj_tomatoes(c)) & 0xFF) > 0)
                cell = obj_set_tick(c, 0);
        } else if (obj_name(h) == DISH) {
            if (obj_name(c) == SOUP && is_ready(L, c)) { h = c; cell = OBJ_NONE; rew = L.soup_rew; }
        } else if (obj_name(h) == ONION || obj_name(h) == TOMATO) {
            uint32_t soup = obj_name(c) == NONE ? obj_make(SOUP) : c;
            if (!(obj_tick(soup) >= 0 || ((obj_onions(soup) + obj_tomatoes(soup)) & 0xFF) == 3)) {
                soup += (obj_name(h) == ONION) ? 0x100u : 0x10000u;
                h = OBJ_NONE;
                rew = L.placement_rew;
            }
            cell = soup;
        }
    }
    return rew;
}

// Interactions under the Simplecooked rules (src/overcooked2_env/sim.cpp:174-286): ONE sequential pass over the
// players in id order — every kind of interaction, not only counters and pots, sees what the lower-numbered players
// did in the same step.  Lane (w, q) takes turn q; what the later turns need from the earlier ones travels through
// shared memory (object slots) and shuffles (held objects of players 0 and 1, the world's dishes-on-counters count).
// Must be called by all 32 lanes.  Returns this lane's share of the reward.
template <int G>
__device__ __forceinline__ int simplecooked_interactions(const Layout &L, const int w, const int a, const bool agent,
                                                         const int act, const int pos, const int ori, uint32_t &held,
                                                         SmemObjsT<32 / G> objs, int &dishes_out)
{
    constexpr int WPW = 32 / G;
    constexpr unsigned FULL = 0xffffffffu;
    // get_pot_states (:174-185), BEFORE the pass: pots holding something that is cooking / cooked or not yet full
    int open_pots = 0;
    for (int k = 0; k < L.npots; k++) {
        const uint32_t o = objs[L.pot_slot[k]];
        open_pots += (obj_name(o) != NONE && (obj_tick(o) >= 0 || ((obj_onions(o) + obj_tomatoes(o)) & 0xFF) < 3)) ? 1 : 0;
    }
    int rew = 0;
    for (int q = 0; q < L.P; q++) {
        // is_dish_pickup_useful (:187-196) looks at players 0 and 1, as they are at this point of the pass
        const uint32_t h0 = __shfl_sync(FULL, held, w), h1 = __shfl_sync(FULL, held, (WPW + w) & 31);
        if (agent && a == q && act == INTERACT) {
            const int face = move_in_direction(pos, ori, L.W);
            const uint32_t ci = L.cell[face];
            const int t = (int)(ci & 0xFFu);
            if (t == COUNTER) {
                uint32_t &cell = objs[(ci >> 8) & 0xFFu];
                const uint32_t c = cell;
                if (obj_name(held) != NONE && obj_name(c) == NONE) {
                    cell = held;
                    if (obj_name(held) == DISH) dishes_out = (dishes_out + 1) & 0xFF;
                    held = OBJ_NONE;
                } else if (obj_name(held) == NONE && obj_name(c) != NONE) {
                    if (obj_name(c) == DISH) dishes_out = (dishes_out - 1) & 0xFF;
                    held = c;
                    cell = OBJ_NONE;
                }
            } else if (t == ONION_SOURCE || t == TOMATO_SOURCE) {
                if (obj_name(held) == NONE) held = obj_make(t == ONION_SOURCE ? ONION : TOMATO);
            } else if (t == DISH_SOURCE) {
                if (obj_name(held) == NONE) {
                    const int dishes = (obj_name(h0) == DISH ? 1 : 0) + (L.P > 1 && obj_name(h1) == DISH ? 1 : 0);
                    if (dishes_out == 0 && dishes < open_pots) rew += L.dish_rew;
                    held = obj_make(DISH);
                }
            } else if (t == POT) {
                uint32_t &cell = objs[(ci >> 8) & 0xFFu];
                uint32_t soup = cell;
                if (obj_name(held) == DISH && obj_name(soup) == SOUP && is_ready(L, soup)) {
                    held = soup;
                    cell = OBJ_NONE;
                    rew += L.soup_rew;
                } else if (obj_name(held) == ONION || obj_name(held) == TOMATO) {
                    if (obj_name(soup) == NONE) soup = obj_make(SOUP);
                    if (!(obj_tick(soup) >= 0 || ((obj_onions(soup) + obj_tomatoes(soup)) & 0xFF) == 3)) {
                        soup += (obj_name(held) == ONION) ? 0x100u : 0x10000u;
                        held = OBJ_NONE;
                        rew += L.placement_rew;
                    }
                    // a full idle pot starts cooking by itself (:270-272)
                    if (obj_tick(soup) < 0 && ((obj_onions(soup) + obj_tomatoes(soup)) & 0xFF) == 3) soup = obj_set_tick(soup, 0);
                    cell = soup;
                }
            } else if (t == SERVING) {
                if (obj_name(held) == SOUP) {   // deliver_soup, :144-148
                    rew += L.recipe_values[obj_recipe(held)];
                    held = OBJ_NONE;
                }
            }
        }
        __syncwarp();
        dishes_out = __shfl_sync(FULL, dishes_out, (q * WPW + w) & 31);
    }
    return rew;
}

// One step of the world logic for the lane that owns agent `a` of world `w` (see step_tile).  Must be called by all 32
// lanes (shuffles / ballots inside).  In/out: the agent's pose and held object, the world's timestep; `objs` are the
// world's object slots in shared memory, shared by its G lanes.  Out: team reward and done.
template <int G>
__device__ __forceinline__ void world_logic(const Layout &L, const int lane, const int w, const int a, const bool live,
                                            const bool agent, const int act, int &pos, int &ori, uint32_t &held,
                                            SmemObjsT<32 / G> objs, int &timestep, int &reward, int &done)
{
    constexpr int WPW = 32 / G;
    constexpr unsigned FULL = 0xffffffffu;
    // N / S / E / W as signed byte offsets on the flat cell index (move_in_direction, sim.cpp:185-197)
    const uint32_t dpack = ((uint32_t)(-L.W) & 0xFFu) | ((uint32_t)L.W << 8) | (1u << 16) | (0xFFu << 24);

    // Simplecooked keeps the world's count of dishes lying on counters (num_dishes_out, uint8) in the top byte of the
    // timestep word
    int dishes_out = 0;
    if (L.rules) {
        dishes_out = (timestep >> 24) & 0xFF;
        timestep &= 0xFFFFFF;
    }

    // ---- interactions (sim.cpp:208-358), on the pre-move pose ----
    int my_rew = 0, face = 0, face_terrain = AIR, face_slot = 0;
    bool cp = false;   // this agent has a counter / pot interaction to perform
    if (!L.rules && agent && act == INTERACT) {
        face = pos + (int)(int8_t)(dpack >> (8 * ori));
        const uint32_t ci = L.cell[face];
        face_terrain = (int)(ci & 0xFFu);
        face_slot = (int)((ci >> 8) & 0xFFu);
        if (face_terrain == ONION_SOURCE || face_terrain == TOMATO_SOURCE || face_terrain == DISH_SOURCE) {
            if (obj_name(held) == NONE)
                held = obj_make(face_terrain == ONION_SOURCE ? ONION : face_terrain == TOMATO_SOURCE ? TOMATO : DISH);
        } else if (face_terrain == SERVING) {
            if (obj_name(held) == SOUP) {   // deliver_soup, sim.cpp:169-173: pays whatever the recipe table says
                my_rew = L.recipe_values[obj_recipe(held)];
                held = OBJ_NONE;
            }
        } else {
            cp = (face_terrain == POT || face_terrain == COUNTER);
        }
    }
    if (L.rules) {
        my_rew = simplecooked_interactions<G>(L, w, a, agent, act, pos, ori, held, objs, dishes_out);
        cp = false;
    }
    if (G > 1) {
        // rank among the agents of this world facing the same counter / pot (interaction_index, sim.cpp:259-282)
        const int key = cp ? face : 0x100 + lane;   // non-interactors never match anybody
        int rank = 0;
#pragma unroll
        for (int q = 0; q < G - 1; q++) {
            const int other = __shfl_sync(FULL, key, q * WPW + w);
            rank += (q < a && other == key) ? 1 : 0;
        }
        unsigned pending = __ballot_sync(FULL, cp);
        for (int r = 0; pending; r++) {
            if (cp && rank == r) {
                my_rew += counter_pot_interaction(L, face_terrain, objs[face_slot], held);
                cp = false;
            }
            __syncwarp();
            pending = __ballot_sync(FULL, cp);
        }
        reward = my_rew;   // team reward, pre_resolve_interacts / calculated_reward (sim.cpp:201,472)
#pragma unroll
        for (int o = WPW; o < 32; o <<= 1) reward += __shfl_xor_sync(FULL, reward, o);
    } else {
        if (cp) my_rew += counter_pot_interaction(L, face_terrain, objs[face_slot], held);
        reward = my_rew;
    }

    // ---- movement (sim.cpp:363-426) ----
    int npos = pos, nori = ori;
    if (agent && (unsigned)act < 4u) {
        const int cand = pos + (int)(int8_t)(dpack >> (8 * act));
        nori = act;
        npos = ((L.cell[cand] & 0xFFu) == AIR) ? cand : pos;
    }
    bool blocked = false;
    if (G > 1) {
        const int mine = agent ? (npos | (pos << 8)) : 0x10000;
        bool hit = false;
#pragma unroll
        for (int d = 1; d < G; d++) {
            const int other = __shfl_sync(FULL, mine, (lane + d * WPW) & 31);
            const int on = other & 0xFF, op = (other >> 8) & 0xFF;
            hit |= agent && other < 0x10000 && (npos == on || (npos == op && on == pos));
        }
        unsigned world_lanes = 0;
#pragma unroll
        for (int q = 0; q < G; q++) world_lanes |= 1u << (q * WPW);
        blocked = (__ballot_sync(FULL, hit) & (world_lanes << w)) != 0;   // any conflict freezes the whole world
    }
    if (!blocked) pos = npos;
    ori = nori;

    // ---- pots tick (sim.cpp:430-438), after the interactions of this step ----
    if (live)
        for (int k = a; k < L.npots; k += G) {
            uint32_t &cell = objs[L.pot_slot[k]];
            const uint32_t o = cell;
            if (obj_name(o) == SOUP && is_cooking(L, o)) cell = obj_set_tick(o, obj_tick(o) + 1);
        }

    // ---- horizon / auto-reset (sim.cpp:441-489) ----
    timestep += 1;
    done = (live && timestep >= L.horizon) ? 1 : 0;
    __syncwarp();
    if (done) {
        timestep = 0;
        dishes_out = 0;
        for (int k = a; k < L.K; k += G) objs[k] = OBJ_NONE;
        if (agent) {
            pos = L.start_pos[a];
            ori = NORTH;
            held = OBJ_NONE;
        }
    }
    timestep |= dishes_out << 24;   // 0 under the Overcooked rules
    __syncwarp();
}

// Warp-tile kernel: lane = (world, agent).  One warp (= one CTA) owns a tile of WPW = 32/G consecutive worlds and
// gives every world G lanes, lane = a * WPW + w; lane a < P owns agent a: its pose, held object, action, its
// interaction and movement proposal, and its bytes of the observation.  What couples the agents of a world is
// resolved with warp shuffles and ballots:
//   * counter / pot interactions on the SAME cell are ordered by agent id (setup_interact_time + 4 rounds,
//     sim.cpp:259-358): every lane counts the lower-numbered agents facing the same cell (shuffles) and acts in the
//     round of that rank; rounds continue while a ballot says some agent is still waiting (usually one round);
//   * the collision rule is world-global (sim.cpp:383-396,417-423): each lane tests its proposal against every other
//     agent's (shuffles), a ballot masked to the world's lanes decides whether anybody moves;
//   * the team reward is a butterfly sum over the world's lanes.
// World-level work is shared out by index mod G: object slots (load, pot tick, reset, write-back, encoding) and the
// urgency cells.  Lanes a >= P only do that shared work, so G may exceed P to put more warps on an SM.
//
// Observation path: the staging tile [WPW][W][H][C] in shared memory arrives as the template (terrain one-hot) by one
// bulk copy, receives the observer-independent dynamic bytes once, then per observer p each agent lane patches its
// two bytes and the tile (WPW * S * C contiguous bytes of obs[p]) is handed to the copy engine.
// L2 eviction policy of the observation stream: written once, read by somebody else much later
__device__ __forceinline__ uint64_t stream_out_policy()
{
    uint64_t p;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p));
    return p;
}

template <int G, class E>
__global__ void __launch_bounds__(32) step_tile(const uint8_t *__restrict__ gconst, const TileDims D, const Buffers B,
                                                const int init_only)
{
    constexpr int WPW = 32 / G;
    constexpr unsigned FULL = 0xffffffffu;
    // shared: [Layout][stage][objs u32 [K][WPW]][mbarrier]
    extern __shared__ __align__(128) uint8_t smem[];
    const int lane = threadIdx.x;
    const int w = lane % WPW, a = lane / WPW;
    const Layout &L = *reinterpret_cast<const Layout *>(smem);
    uint8_t *stage = smem + sizeof(Layout);
    uint32_t *sobj = reinterpret_cast<uint32_t *>(smem + D.const_bytes);
    uint64_t *mbar = reinterpret_cast<uint64_t *>(sobj + (size_t)D.K * WPW);
    SmemObjsT<WPW> objs{sobj + w};

    // ---- prologue: nothing here depends on the previous step, so under programmatic dependent launch it overlaps
    // the tail of the preceding kernel ----
    trace_stamp(B, 0);
    if (lane == 0) {
        mbar_init(mbar, 1);
        fence_mbar_init();
        mbar_expect_tx(mbar, D.const_bytes);
        bulk_load_g2s(smem, gconst + (size_t)(blockIdx.x % (unsigned)TILE_CONST_COPIES) * D.const_stride, D.const_bytes, mbar);
    }
    __syncwarp();
    const uint32_t SC = D.SC;
    const uint32_t tile0 = blockIdx.x * (uint32_t)WPW;
    const uint32_t n = tile0 + w;
    const bool live = n < B.N;
    const uint32_t tile_worlds = min((uint32_t)WPW, B.N - tile0);
    const int P = B.P;
    const bool agent = live && a < P;
    const size_t an = (size_t)a * B.N + n;

    grid_dependency_wait();        // the previous step's state / the caller's actions are complete and visible
    grid_launch_dependents();      // let the next kernel in the stream start its own prologue
    trace_stamp(B, 1);

    uint32_t held = OBJ_NONE;
    int pos = 0, ori = NORTH, act = STAY, timestep = 0, ep_prev = 0;
    if (!init_only) {
        if (agent) {
            const uint2 pl = B.players[an];
            act = B.actions[an];
            held = pl.x;
            pos = (int)(pl.y & 0xFF);
            ori = (int)((pl.y >> 8) & 0xFF);
        }
        if (live) {
            // object slots go global -> shared without passing through registers (LDGSTS), split over the G lanes
            const uint32_t *src = B.objs + an;
            for (int k = a; k < D.K; k += G, src += (size_t)G * B.N)
                asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(smem_u32(&objs[k])), "l"(src) : "memory");
            timestep = B.timestep[n];
            if (B.ep_return && a == 0) ep_prev = B.ep_return[n];
        }
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    mbar_wait(mbar, 0);            // Layout + template tile have landed (each lane observes the phase itself)
    __syncwarp();                  // ... and the object slots fetched by the other lanes of this world
    trace_stamp(B, 2);

    int reward = 0, done = 0;
    if (init_only) {
        // Sim::Sim (sim.cpp:556-659): reset state, then the initial observation
        if (live)
            for (int k = a; k < L.K; k += G) objs[k] = OBJ_NONE;
        if (agent) pos = L.start_pos[a];
        __syncwarp();
    } else {
        world_logic<G>(L, lane, w, a, live, agent, act, pos, ori, held, objs, timestep, reward, done);
    }

    // ---- state + step outputs go back to HBM AFTER observer 0's tile has been handed to the copy engine: the
    // observation is the critical path (34 of the step's 41 MB), the write-back hides behind the engine's read ----
    auto write_back = [&]() {
        if (agent) {
            B.players[an] = make_uint2(held, (uint32_t)pos | ((uint32_t)ori << 8));
            B.rewards[an] = reward;   // team reward to every agent, sim.cpp:472
        }
        if (live) {
            uint32_t *dst = B.objs + an;
            for (int k = a; k < L.K; k += G, dst += (size_t)G * B.N) *dst = objs[k];
            if (a == 0) {
                B.timestep[n] = timestep;
                B.done[n] = done;
            }
        }
        if (B.ep_return) {   // episode statistics, off the parity path: one atomic pair per warp and only on episode ends
            const bool owner = live && a == 0;
            const int ret = ep_prev + reward;
            if (owner) B.ep_return[n] = done ? 0 : ret;
            const unsigned dmask = __ballot_sync(FULL, owner && done);
            if (dmask) {
                long long sum = (owner && done) ? (long long)ret : 0ll;
    #pragma unroll
                for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(FULL, sum, o);
                if (lane == 0) {
                    atomicAdd(&B.stats[0], (unsigned long long)__popc(dmask));
                    atomicAdd(&B.stats[1], (unsigned long long)sum);
                }
            }
        }
    };

    // ---- observer-independent dynamic bytes: urgency, objects on counters / pots, held objects ----
    E *slab = reinterpret_cast<E *>(stage + (size_t)w * SC);
    if (live) {
        if (!L.rules && (L.horizon - timestep) < 40) {      // sim.cpp:79-83, post-increment / post-reset timestep
            const int ch = 5 * P + 15;
            for (int c = a; c < L.S; c += G) slab[L.cell_off(c) + ch] = obs_val<E>(1);
        }
        for (int k = a; k < L.K; k += G) {
            const uint32_t o = objs[k];
            if (obj_name(o) != NONE) encode_cell_object(L, slab, L.cell_of_slot[k], o);
        }
        if (agent) encode_held_object(L, slab, pos, held);
    }

    E *px = slab + (agent ? L.cell_off(pos) : 0);
    const int po = P + ori;
    for (int p = 0; p < P; p++) {
        if (agent) encode_player_at(px, p, a, po, 1);
        fence_proxy_async_smem();
        __syncwarp();
        slab_store<true>(B.obs + ((size_t)p * B.N + tile0) * SC, stage, tile_worlds * SC, lane, 32, stream_out_policy());
        if (p == 0) {
            trace_stamp(B, 4);
            write_back();
            trace_stamp(B, 3);
        }
        if (p + 1 < P) {
            if (lane == 0) bulk_wait_read<0>();   // the copy engine has finished reading the staging tile
            __syncwarp();
            if (p == 0) trace_stamp(B, 5);
            if (agent) encode_player_at(px, p, a, po, 0);
        }
    }
    if (lane == 0) bulk_wait_read<0>();
    trace_stamp(B, 6);
}

// ---------------------------------------------------------------------------------------------------------------
// Fused rollout: T steps in ONE launch (SURVEY 8(f) row 1 — the caller side of the path when the actions of all T
// steps are known up front: random-action rollouts, scripted agents, replay).  Same lane = (world, agent) layout as
// step_tile, but each warp keeps its tile's worlds ON CHIP for the whole rollout: poses / held objects / timestep in
// registers, object slots in shared memory, and the staging tile stays resident too — per step only the bytes that
// went stale are erased and the new ones written (a shadow copy records what the tile currently encodes).  Worlds
// never exchange anything, so warps drift freely: one warp's copy-engine drain overlaps other warps' logic, and the
// launch / dependency latency that bounds a 6 us single-step kernel disappears.  HBM traffic per env-step drops to
// obs + actions + rewards + done (state is read and written once per launch).
// ---------------------------------------------------------------------------------------------------------------
constexpr int ACT_AHEAD = 7;   // steps between the fetch of an action and its use in rollout_tile
struct RolloutArgs {
    const int32_t *actions;   // [A][P][N]
    uint8_t *obs;             // [O][P][N][W][H][C]
    int32_t *rewards;         // [O][P][N]
    int32_t *done;            // [O][N]
    int32_t T, A;
    int32_t O_obs, O_rew, O_done;   // slices of each output ring (1 = the simulator's own tensor)
    // b200env_random_rollout: no action ring; every agent draws policy::sample_kernel's uniformly random move for
    // (seed, first_step + t, flat action index) itself, and the last step's draw is left in the action tensor
    int32_t sample;
    uint32_t seed, first_step;
};

template <int G, class E>
__global__ void __launch_bounds__(32) rollout_tile(const uint8_t *__restrict__ gconst, const TileDims D, const Buffers B,
                                                   const RolloutArgs R)
{
    constexpr int WPW = 32 / G;
    constexpr unsigned FULL = 0xffffffffu;
    // shared: [Layout][stage][objs u32 [K][WPW]][encoded objs u32 [K][WPW]][mbarrier][actions i32 [ACT_AHEAD + 1][32]]
    extern __shared__ __align__(128) uint8_t smem[];
    const int lane = threadIdx.x;
    const int w = lane % WPW, a = lane / WPW;
    const Layout &L = *reinterpret_cast<const Layout *>(smem);
    uint8_t *stage = smem + sizeof(Layout);
    uint32_t *sobj = reinterpret_cast<uint32_t *>(smem + D.const_bytes);
    uint32_t *senc = sobj + (size_t)D.K * WPW;
    uint64_t *mbar = reinterpret_cast<uint64_t *>(senc + (size_t)D.K * WPW);
    int32_t *sact = reinterpret_cast<int32_t *>(mbar + 1) + lane;   // this lane's column of the action queue
    SmemObjsT<WPW> objs{sobj + w}, enc{senc + w};

    if (lane == 0) {
        mbar_init(mbar, 1);
        fence_mbar_init();
        mbar_expect_tx(mbar, D.const_bytes);
        bulk_load_g2s(smem, gconst + (size_t)(blockIdx.x % (unsigned)TILE_CONST_COPIES) * D.const_stride, D.const_bytes, mbar);
    }
    __syncwarp();
    const uint32_t SC = D.SC;
    const uint32_t tile0 = blockIdx.x * (uint32_t)WPW;
    const uint32_t n = tile0 + w;
    const bool live = n < B.N;
    const uint32_t tile_worlds = min((uint32_t)WPW, B.N - tile0);
    const int P = B.P;
    const bool agent = live && a < P;
    const size_t PN = (size_t)P * B.N, an = (size_t)a * B.N + n;

    grid_dependency_wait();
    // No early launch_dependents here: a following rollout launch would park 3 CTAs per SM (what shared memory leaves
    // free) for the whole 2 ms and then fill slots in the order this grid's warps happen to exit — up to 17 CTAs on
    // some SMs and 10 on others instead of an even 13.8, which made 50 back-to-back 400-step launches 9 % slower
    // than one (5.95 vs 5.47 us per step).  The dependent grid starts when this one has exited.

    uint32_t held = OBJ_NONE;
    int pos = 0, ori = NORTH, timestep = 0, ep_ret = 0;
    if (agent) {
        const uint2 pl = B.players[an];
        held = pl.x;
        pos = (int)(pl.y & 0xFF);
        ori = (int)((pl.y >> 8) & 0xFF);
    }
    if (live) {
        const uint32_t *src = B.objs + an;
        for (int k = a; k < D.K; k += G, src += (size_t)G * B.N) {
            asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(smem_u32(&objs[k])), "l"(src) : "memory");
            enc[k] = OBJ_NONE;
        }
        timestep = B.timestep[n];
        if (B.ep_return && a == 0) ep_ret = B.ep_return[n];
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    mbar_wait(mbar, 0);
    __syncwarp();

    // what the staging tile currently encodes for this lane's agent / for the world
    uint32_t enc_held = OBJ_NONE;
    int enc_pos = pos, enc_ori = ori;
    E *enc_x = reinterpret_cast<E *>(stage);   // C-vector of the cell the agent's player bytes were last written to
    bool enc_urgent = false;
    long long stat_return = 0;
    int stat_episodes = 0;
    E *slab = reinterpret_cast<E *>(stage + (size_t)w * SC);

    // ring cursors advance by pointer increments (an integer modulo per step and output costs ~25 instructions each)
    // Actions are fetched ACT_AHEAD steps before they are used, into a shared-memory queue (cp.async, one group per
    // step): with the observation stream saturating HBM a load takes several microseconds — longer than a step — and
    // a one-step register prefetch put that latency on every warp's critical path (6.26 us per step; 5.5 without it).
    const int32_t *act_p = R.actions + an;
    int act_i = 0, act_slot = 0;
    // L2 policies: the observation stream is written once and read by somebody else much later (evict first); the
    // action ring is small and read again every A steps (evict last).  Without them the 8 B of actions per env-step
    // miss in L2 every time, and those few reads sprinkled into a saturated write stream cost 12 % of the HBM rate
    // (bus turnarounds): 6.26 -> 5.52 us per step, the rate of a store-only probe (tools/probes/bulk_store_probe.cu).
    uint64_t keep_in_l2;
    const uint64_t stream_out = stream_out_policy();
    asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(keep_in_l2));
    auto fetch_action = [&](int step) {
        if (agent && step < R.T)
            asm volatile("cp.async.ca.shared.global.L2::cache_hint [%0], [%1], 4, %2;" ::"r"(smem_u32(sact + 32 * act_slot)),
                         "l"(act_p), "l"(keep_in_l2) : "memory");
        asm volatile("cp.async.commit_group;" ::: "memory");
        if (++act_i == R.A) { act_i = 0; act_p = R.actions + an; } else act_p += PN;
        if (++act_slot == ACT_AHEAD + 1) act_slot = 0;
    };
    int act_use = 0;
    if (!R.sample)
        for (int k = 0; k < ACT_AHEAD; k++) fetch_action(k);
    int32_t *rew_p = R.rewards + an, *done_p = R.done + n;
    uint8_t *obs_p = R.obs + (size_t)tile0 * SC;
    int rew_i = 0, done_i = 0, obs_i = 0;
    const size_t obs_observer = (size_t)B.N * SC, obs_slice = PN * SC;

    // (starting the warps spread over one step period instead of together was tried: no difference)
    for (int t = 0; t < R.T; t++) {
        int act = STAY;
        if (R.sample) {
            if (agent) {
                act = (int)(((uint64_t)policy::mix(R.seed, R.first_step + (uint32_t)t, (uint32_t)an) * 6u) >> 32);
                if (t + 1 == R.T) const_cast<int32_t *>(B.actions)[an] = act;
            }
        } else {
            fetch_action(t + ACT_AHEAD);
            asm volatile("cp.async.wait_group %0;" ::"n"(ACT_AHEAD) : "memory");   // step t's action has landed
            if (agent) act = sact[32 * act_use];
            if (++act_use == ACT_AHEAD + 1) act_use = 0;
        }
        int reward = 0, done = 0;
        world_logic<G>(L, lane, w, a, live, agent, act, pos, ori, held, objs, timestep, reward, done);

        if (agent) *rew_p = reward;
        if (live && a == 0) {
            *done_p = done;
            ep_ret += reward;
            if (done) { stat_return += ep_ret; stat_episodes++; ep_ret = 0; }
        }
        if (++rew_i == R.O_rew) { rew_i = 0; rew_p = R.rewards + an; } else rew_p += PN;
        if (++done_i == R.O_done) { done_i = 0; done_p = R.done + n; } else done_p += B.N;

        // the copy engine must have read the previous step's last tile before the staging tile changes
        if (t > 0) {
            if (lane == 0) bulk_wait_read<0>();
            __syncwarp();
        }
        // ---- erase what went stale ----
        if (agent && t > 0) encode_player_at(enc_x, P - 1, a, P + enc_ori, 0);
        const bool held_changed = agent && (held != enc_held || pos != enc_pos);
        if (held_changed) encode_held_object<false>(L, slab, enc_pos, enc_held);
        if (live) {
            const bool urgent = !L.rules && (L.horizon - timestep) < 40;
            if (urgent != enc_urgent) {
                const int ch = 5 * P + 15;
                for (int c = a; c < L.S; c += G) slab[L.cell_off(c) + ch] = obs_val<E>(urgent ? 1 : 0);
                enc_urgent = urgent;
            }
            for (int k = a; k < L.K; k += G) {   // a slot's cell belongs to one lane: erase + rewrite in place
                const uint32_t now = objs[k], was = enc[k];
                if (now != was) {
                    const int cell = L.cell_of_slot[k];
                    encode_cell_object<false>(L, slab, cell, was);
                    encode_cell_object<true>(L, slab, cell, now);
                    enc[k] = now;
                }
            }
        }
        __syncwarp();   // erases of one agent's old cell precede writes of another agent entering it
        // ---- write the new bytes ----
        if (held_changed) {
            encode_held_object<true>(L, slab, pos, held);
            enc_held = held;
        }
        enc_pos = pos;
        enc_ori = ori;
        enc_x = slab + (agent ? L.cell_off(pos) : 0);
        const int po = P + ori;
        uint8_t *dst = obs_p;
        if (++obs_i == R.O_obs) { obs_i = 0; obs_p = R.obs + (size_t)tile0 * SC; } else obs_p += obs_slice;
        for (int p = 0; p < P; p++, dst += obs_observer) {
            if (agent) encode_player_at(enc_x, p, a, po, 1);
            fence_proxy_async_smem();
            __syncwarp();
            slab_store<true>(dst, stage, tile_worlds * SC, lane, 32, stream_out);
            if (p + 1 < P) {
                if (lane == 0) bulk_wait_read<0>();   // register stores were tried here: slower than the copy engine
                __syncwarp();
                if (agent) encode_player_at(enc_x, p, a, po, 0);
            }
        }
    }
    if (lane == 0) bulk_wait_read<0>();

    // ---- state back to HBM, once per launch ----
    if (agent) B.players[an] = make_uint2(held, (uint32_t)pos | ((uint32_t)ori << 8));
    if (live) {
        uint32_t *dst = B.objs + an;
        for (int k = a; k < L.K; k += G, dst += (size_t)G * B.N) *dst = objs[k];
        if (a == 0) {
            B.timestep[n] = timestep;
            if (B.ep_return) B.ep_return[n] = ep_ret;
        }
    }
    if (B.ep_return) {
        const unsigned any = __ballot_sync(FULL, stat_episodes != 0);
        if (any) {
#pragma unroll
            for (int o2 = 16; o2 > 0; o2 >>= 1) {
                stat_return += __shfl_xor_sync(FULL, stat_return, o2);
                stat_episodes += __shfl_xor_sync(FULL, stat_episodes, o2);
            }
            if (lane == 0) {
                atomicAdd(&B.stats[0], (unsigned long long)stat_episodes);
                atomicAdd(&B.stats[1], (unsigned long long)stat_return);
            }
        }
    }
}

// ---------------------------------------------------------------------------------------------------------------
// CTA-per-world kernel for many-agent / large-grid layouts.
// ---------------------------------------------------------------------------------------------------------------
constexpr int WORLD_CTA_THREADS = 256;

// cooperative copy of the layout into shared memory (sizeof(Layout) is a multiple of 16)
__device__ __forceinline__ void stage_layout(Layout *dst, const Layout *__restrict__ src, int tid, int nthreads)
{
    const uint4 *s = reinterpret_cast<const uint4 *>(src);
    uint4 *d = reinterpret_cast<uint4 *>(dst);
    for (int i = tid; i < (int)(sizeof(Layout) / 16); i += nthreads) d[i] = __ldg(s + i);
}

// E = element type of the observation tensor (uint8_t: the reference's int8; uint16_t: bfloat16 bits; float), NBUF = slab
// buffers in shared memory: two when they fit (the copy of observer p drains while observer p+1 is patched), else one
template <class E, int NBUF>
__global__ void __launch_bounds__(WORLD_CTA_THREADS) step_world_cta(const Layout *__restrict__ gL, const Buffers B,
                                                                   const int init_only)
{
    extern __shared__ __align__(128) uint8_t smem[];
    __shared__ Layout sL;
    __shared__ int s_act[64];
    __shared__ uint8_t s_pos[64], s_ori[64];
    __shared__ uint32_t s_held[64];
    __shared__ uint32_t s_obj[256];
    __shared__ int s_misc[4];   // timestep, reward, done
    __shared__ int16_t s_face[64];   // the counter / pot cell each agent interacts with this step (-1 = none)
    __shared__ uint8_t s_npos[64];   // proposed positions

    const int tid = threadIdx.x;
    stage_layout(&sL, gL, tid, WORLD_CTA_THREADS);
    __syncthreads();
    const Layout &L = sL;
    const int P = L.P;
    const uint32_t SC = (uint32_t)L.S * L.C * (uint32_t)sizeof(E);   // bytes of one slab
    const uint32_t n = blockIdx.x;

    // slab buffers; buffer b serves observers p == b (mod NBUF) and is placed congruent (mod 16) to its first
    // destination so that the interior of every slab can go through the copy engine
    const uint32_t buf_stride = (SC + 31u) & ~15u;
    uint8_t *obs_w = B.obs + (size_t)n * SC;                 // observer 0 slab of this world
    const size_t obs_pstride = (size_t)B.N * SC;             // distance between observers
    uint8_t *buf[2];
    buf[0] = smem + (reinterpret_cast<uintptr_t>(obs_w) & 15u);
    buf[1] = NBUF > 1 ? smem + buf_stride + (reinterpret_cast<uintptr_t>(obs_w + obs_pstride) & 15u) : buf[0];

    if (!init_only) {
        if (tid < P) {
            const uint2 pl = B.players[(size_t)tid * B.N + n];
            s_held[tid] = pl.x;
            s_pos[tid] = (uint8_t)(pl.y & 0xFF);
            s_ori[tid] = (uint8_t)((pl.y >> 8) & 0xFF);
            const int act = B.actions[(size_t)tid * B.N + n];
            s_act[tid] = (unsigned)act <= (unsigned)INTERACT ? act : STAY;   // out-of-range moves (undefined in the reference) stay put, as in step_tile
        }
        for (int k = tid; k < L.K; k += WORLD_CTA_THREADS) s_obj[k] = B.objs[(size_t)k * B.N + n];
        if (tid == 0) s_misc[0] = B.timestep[n];
    } else {
        if (tid < P) { s_held[tid] = OBJ_NONE; s_pos[tid] = L.start_pos[tid]; s_ori[tid] = NORTH; }
        for (int k = tid; k < L.K; k += WORLD_CTA_THREADS) s_obj[k] = OBJ_NONE;
        if (tid == 0) { s_misc[0] = 0; s_misc[1] = 0; s_misc[2] = 0; }
    }
    {
        uint4 *s4 = reinterpret_cast<uint4 *>(smem);
        const uint32_t chunks = ((uint32_t)NBUF * buf_stride) / 16u;
        for (uint32_t i = tid; i < chunks; i += WORLD_CTA_THREADS) s4[i] = make_uint4(0, 0, 0, 0);
    }
    __syncthreads();

    // ---- world logic, one thread per agent (thread q < P owns agent q); what couples the agents goes through shared
    // memory and block-wide votes.  Same equivalences as world_logic<G>: source / serving interactions touch only the
    // agent's own hands; counter / pot interactions on one cell are applied in ascending agent id, one rank per round;
    // the collision rule is a world-wide vote. ----
    if (!init_only) {
        const bool agent = tid < P;
        // interactions (sim.cpp:208-358), on the pre-move pose
        uint32_t h = OBJ_NONE;
        int act = STAY, pos = 0, ori = NORTH, face = -1, face_terrain = AIR, face_slot = 0, my_rew = 0;
        bool cp = false;   // this agent has a counter / pot interaction to perform
        if (agent) {
            h = s_held[tid]; act = s_act[tid]; pos = s_pos[tid]; ori = s_ori[tid];
            if (act == INTERACT) {
                face = move_in_direction(pos, ori, L.W);
                const uint32_t ci = L.cell[face];
                face_terrain = (int)(ci & 0xFFu);
                face_slot = (int)((ci >> 8) & 0xFFu);
                if (face_terrain == ONION_SOURCE || face_terrain == TOMATO_SOURCE || face_terrain == DISH_SOURCE) {
                    if (obj_name(h) == NONE)
                        h = obj_make(face_terrain == ONION_SOURCE ? ONION : face_terrain == TOMATO_SOURCE ? TOMATO : DISH);
                } else if (face_terrain == SERVING) {
                    if (obj_name(h) == SOUP) {   // deliver_soup, sim.cpp:169-173
                        my_rew = L.recipe_values[obj_recipe(h)];
                        h = OBJ_NONE;
                    }
                } else {
                    cp = face_terrain == POT || face_terrain == COUNTER;
                }
            }
            s_face[tid] = (int16_t)(cp ? face : -1);
        }
        if (tid == 0) s_misc[1] = 0;
        __syncthreads();
        // rank among the agents facing the same counter / pot (interaction_index, sim.cpp:259-282): lower ids go first
        int rank = 0;
        if (cp)
            for (int q = 0; q < tid; q++) rank += (s_face[q] == face) ? 1 : 0;
        for (int r = 0; __syncthreads_or(cp && rank >= r); r++)   // (at most four agents can stand next to one cell)
            if (cp && rank == r) my_rew += counter_pot_interaction(L, face_terrain, s_obj[face_slot], h);
        if (my_rew) atomicAdd(&s_misc[1], my_rew);   // team reward, pre_resolve_interacts / calculated_reward (sim.cpp:201,472)

        // movement (sim.cpp:363-426): propose, then nobody moves if any two agents collide or swap
        int npos = pos, nori = ori;
        if (agent && act != INTERACT) {
            const int cand = move_in_direction(pos, act, L.W);
            nori = act == STAY ? ori : act;
            npos = L.terrain(cand) != AIR ? pos : cand;
        }
        if (agent) { s_held[tid] = h; s_npos[tid] = (uint8_t)npos; }
        __syncthreads();
        bool hit = false;
        if (agent)
            for (int q = 0; q < P; q++)
                hit |= q != tid && (npos == (int)s_npos[q] || (npos == (int)s_pos[q] && (int)s_npos[q] == pos));
        const int collide = __syncthreads_or(hit);   // (every read of s_pos above precedes the writes below)
        if (agent) {
            if (!collide) s_pos[tid] = (uint8_t)npos;
            s_ori[tid] = (uint8_t)nori;
        }
        // pots tick (sim.cpp:430-438), after the interactions of this step
        for (int k = tid; k < L.npots; k += WORLD_CTA_THREADS) {
            uint32_t &cell = s_obj[L.pot_slot[k]];
            const uint32_t o = cell;
            if (obj_name(o) == SOUP && is_cooking(L, o)) cell = obj_set_tick(o, obj_tick(o) + 1);
        }
        // horizon / auto-reset (sim.cpp:441-489)
        const int t1 = s_misc[0] + 1;
        const int done_now = t1 >= L.horizon ? 1 : 0;
        __syncthreads();
        if (done_now) {
            for (int k = tid; k < L.K; k += WORLD_CTA_THREADS) s_obj[k] = OBJ_NONE;
            if (agent) { s_pos[tid] = L.start_pos[tid]; s_ori[tid] = NORTH; s_held[tid] = OBJ_NONE; }
        }
        if (tid == 0) { s_misc[0] = done_now ? 0 : t1; s_misc[2] = done_now; }
    }
    __syncthreads();

    const int timestep = s_misc[0], reward = s_misc[1], done = s_misc[2];
    if (tid < P) {
        B.players[(size_t)tid * B.N + n] = make_uint2(s_held[tid], (uint32_t)s_pos[tid] | ((uint32_t)s_ori[tid] << 8));
        B.rewards[(size_t)tid * B.N + n] = reward;
    }
    for (int k = tid; k < L.K; k += WORLD_CTA_THREADS) B.objs[(size_t)k * B.N + n] = s_obj[k];
    if (tid == 0) {
        B.timestep[n] = timestep;
        B.done[n] = done;
        if (B.ep_return) {
            const int ret = (init_only ? 0 : B.ep_return[n]) + reward;
            B.ep_return[n] = done ? 0 : ret;
            if (done) {
                atomicAdd(&B.stats[0], 1ull);
                atomicAdd(&B.stats[1], (unsigned long long)(long long)ret);
            }
        }
    }

    // observer-independent slab content, into every buffer
    const int nbuf = P > 1 ? NBUF : 1;
    const bool urgent = (L.horizon - timestep) < 40;
    for (int b = 0; b < nbuf; b++) {
        E *slab = reinterpret_cast<E *>(buf[b]);
        for (int c = tid; c < L.S; c += WORLD_CTA_THREADS) {
            E *x = slab + L.cell_off(c) + 5 * P;
            const int t = L.terrain(c);
            if (t) x[t - 1] = obs_val<E>(1);
            if (urgent) x[15] = obs_val<E>(1);
        }
        for (int k = tid; k < L.K; k += WORLD_CTA_THREADS) encode_cell_object(L, slab, L.cell_of_slot[k], s_obj[k]);
        if (tid < P) encode_held_object(L, slab, s_pos[tid], s_held[tid]);
    }
    __syncthreads();

    for (int p = 0; p < P; p++) {
        E *slab = reinterpret_cast<E *>(buf[p % NBUF]);
        if (p >= NBUF) {
            if (tid == 0) bulk_wait_read<NBUF - 1>();   // the copy of observer p - NBUF has released this buffer
            __syncthreads();
            if (tid < P) encode_player(L, slab, p - NBUF, tid, s_pos[tid], s_ori[tid], 0);
        }
        if (tid < P) encode_player(L, slab, p, tid, s_pos[tid], s_ori[tid], 1);
        fence_proxy_async_smem();
        __syncthreads();
        slab_store<true>(obs_w + (size_t)p * obs_pstride, reinterpret_cast<const uint8_t *>(slab), SC, tid, WORLD_CTA_THREADS, stream_out_policy());
    }
    if (tid == 0) bulk_wait_read<0>();
}

// ---------------------------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------------------------
static TileDims tile_dims(const Layout &L, int G, int es)
{
    TileDims D;
    D.SC = (uint32_t)L.S * L.C * (uint32_t)es;
    D.stage_bytes = ((32u / G) * D.SC + 15u) & ~15u;
    D.const_bytes = (uint32_t)sizeof(Layout) + D.stage_bytes;
    D.const_stride = (D.const_bytes + 255u) & ~255u;
    D.K = L.K;
    return D;
}
static size_t tile_smem_bytes(const Layout &L, int G, int es)
{
    const TileDims D = tile_dims(L, G, es);
    return (size_t)D.const_bytes + (size_t)L.K * (32 / G) * 4 + 16;
}
static size_t world_cta_smem_bytes(const Layout &L, int es, int nbuf)
{
    const size_t SC = (size_t)L.S * L.C * (size_t)es;
    return (size_t)nbuf * ((SC + 31) & ~(size_t)15) + 16;
}
// slab buffers of the CTA-per-world kernel for this layout and element size: 2 if they fit next to its static shared
// memory (~4 KB), else 1, else 0 = the layout's slab does not fit at all
static size_t world_cta_smem_most()   // dynamic shared memory the kernel may ask for (its static part is ~6 KB)
{
    return (size_t)device_info().smem_optin > 16 * 1024 ? (size_t)device_info().smem_optin - 8 * 1024 : 0;
}
static int world_cta_buffers(const Layout &L, int es)
{
    const size_t most = world_cta_smem_most();
    return world_cta_smem_bytes(L, es, 2) <= most ? 2 : world_cta_smem_bytes(L, es, 1) <= most ? 1 : 0;
}
bool world_cta_fits(const Layout &L, int es) { return world_cta_buffers(L, es) > 0; }

static int min_lanes(const Layout &L) { return L.P <= 1 ? 1 : L.P <= 2 ? 2 : L.P <= 4 ? 4 : 8; }

bool tile_fits(const Layout &L, int G, int es)
{
    return L.P <= 8 && G >= min_lanes(L) && tile_smem_bytes(L, G, es) <= 200 * 1024;
}

// lanes per world: at least one per agent; beyond that, enough to put ~12+ warps on every SM (measured best at 32 K cramped_room worlds: 2 lanes) (a lone warp is
// latency-bound), preferring tiles whose byte size keeps the bulk-copy path (multiple of 16)
int pick_lanes(const Layout &L, uint32_t N, int es)
{
    int G = min_lanes(L);
    while (G < 8 && ((double)N * G / 32.0 / (double)device_info().sms < 12.0 || !tile_fits(L, G, es))) G *= 2;
    const uint32_t SC = (uint32_t)L.S * L.C * (uint32_t)es;
    // bulk stores of ~8 KB per warp stream best (profiles/r01x_obs_dtype_sweep.txt: bfloat16 cramped_room 9.97 us per
    // fused step with 4 lanes per world against 11.08 with 2; more than 4 lanes lose again to idle lanes)
    while (G < 4 && (32u / G) * SC > 12u * 1024u) G *= 2;
    while (G > min_lanes(L) && ((32u / G) * SC) % 16u != 0 && tile_fits(L, G / 2, es)) G /= 2;
    return G;
}

int pick_kernel(const Layout &L, uint32_t N, int es)
{
    // the warp-tile kernel needs P <= 8 and a staging tile small enough to keep >= 4 CTAs per SM
    if (L.P > 8) return 0;
    const int G = pick_lanes(L, N, es);
    return (tile_fits(L, G, es) && tile_smem_bytes(L, G, es) <= 56 * 1024) ? 1 : 0;
}

size_t tile_const_bytes(const Layout &L, int G, int es) { return tile_dims(L, G, es).const_bytes; }
size_t tile_const_stride(const Layout &L, int G, int es) { return tile_dims(L, G, es).const_stride; }

// the value 1 as an element of `es` bytes: int8, bfloat16 bits or float
static void put_one(uint8_t *tile, size_t elem, int es)
{
    if (es == 1) tile[elem] = 1;
    else if (es == 2) { const uint16_t one = 0x3F80; memcpy(tile + 2 * elem, &one, 2); }
    else { const float one = 1.0f; memcpy(tile + 4 * elem, &one, 4); }
}

void build_tile_const(const Layout &L, int G, int es, uint8_t *dst)
{
    const TileDims D = tile_dims(L, G, es);
    const size_t SCe = (size_t)L.S * L.C;   // elements of one slab
    memcpy(dst, &L, sizeof(Layout));
    uint8_t *tile = dst + sizeof(Layout);
    memset(tile, 0, D.stage_bytes);
    for (int w = 0; w < 32 / G; w++)
        for (int c = 0; c < L.S; c++) {
            const int t = (int)(L.cell[c] & 0xFFu);
            if (!t) continue;
            int ch = t - 1;                                                            // sim.cpp:642-645
            if (L.rules) {
                // Simplecooked numbers its terrain AIR, POT, COUNTER, ONION_SOURCE, DISH_SOURCE, SERVING, TOMATO_SOURCE
                // (src/overcooked2_env/sim.hpp:40) and sets channel 5P + that - 1 (sim.cpp:555-560); the tomato
                // source's bit would land on 5P+5, which observationSystem clears on every pass (:79) — never visible
                static const int simple_value[7] = {0, 1, 2, 3, 6, 4, 5};
                ch = simple_value[t] - 1;
                if (ch == 5) continue;
            }
            put_one(tile, (size_t)w * SCe + (size_t)L.cell_off(c) + 5 * L.P + ch, es);
        }
}

// every step kernel is launched with the programmatic-stream-serialization attribute (see common.cuh)
template <class... KArgs, class... Args>
static cudaError_t launch_pdl(void (*kernel)(KArgs...), unsigned grid, unsigned block, size_t smem, cudaStream_t st,
                              Args... args)
{
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(block);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    return cudaLaunchKernelEx(&cfg, kernel, KArgs(args)...);
}

template <int G, class E>
static cudaError_t launch_tile_ge(const Layout &hL, const LaunchArgs &a, const Buffers &B, int init_only, cudaStream_t st)
{
    if (cudaError_t e = ensure_dynamic_smem(step_tile<G, E>, 200 * 1024)) return e;
    constexpr unsigned WPW = 32 / G;
    const unsigned grid = (B.N + WPW - 1) / WPW;
    const size_t smem = even_wave_smem(tile_smem_bytes(hL, G, (int)sizeof(E)), 0, grid);   // common.cuh
    return launch_pdl(step_tile<G, E>, grid, 32, smem, st, a.tile_const[lanes_index(G)],
                      tile_dims(hL, G, (int)sizeof(E)), B, init_only);
}
template <int G>
static cudaError_t launch_tile_g(const Layout &hL, const LaunchArgs &a, const Buffers &B, int init_only, cudaStream_t st)
{
    switch (a.obs_elem) {
    case 1: return launch_tile_ge<G, uint8_t>(hL, a, B, init_only, st);
    case 2: return launch_tile_ge<G, uint16_t>(hL, a, B, init_only, st);
    case 4: return launch_tile_ge<G, float>(hL, a, B, init_only, st);
    default: return cudaErrorInvalidValue;
    }
}

static cudaError_t launch_tile(const Layout &hL, const LaunchArgs &a, const Buffers &B, int init_only, cudaStream_t st)
{
    if (a.lanes < 1 || a.lanes > 8 || !a.tile_const[lanes_index(a.lanes)]) return cudaErrorInvalidValue;
    switch (a.lanes) {
    case 1: return launch_tile_g<1>(hL, a, B, init_only, st);
    case 2: return launch_tile_g<2>(hL, a, B, init_only, st);
    case 4: return launch_tile_g<4>(hL, a, B, init_only, st);
    default: return launch_tile_g<8>(hL, a, B, init_only, st);
    }
}

static size_t rollout_smem_bytes(const Layout &L, int G, int es)
{
    return tile_smem_bytes(L, G, es) + (size_t)L.K * (32 / G) * 4 + (size_t)(ACT_AHEAD + 1) * 32 * 4;
}

template <int G, class E>
static cudaError_t launch_rollout_ge(const Layout &hL, const LaunchArgs &a, const Buffers &B, const RolloutArgs &R,
                                     cudaStream_t st)
{
    const size_t smem = rollout_smem_bytes(hL, G, (int)sizeof(E));
    if (cudaError_t e = ensure_dynamic_smem(rollout_tile<G, E>, 200 * 1024)) return e;
    constexpr unsigned WPW = 32 / G;
    const unsigned grid = (B.N + WPW - 1) / WPW;
    return launch_pdl(rollout_tile<G, E>, grid, 32, smem, st, a.tile_const[lanes_index(G)],
                      tile_dims(hL, G, (int)sizeof(E)), B, R);
}
template <int G>
static cudaError_t launch_rollout_g(const Layout &hL, const LaunchArgs &a, const Buffers &B, const RolloutArgs &R,
                                    cudaStream_t st)
{
    switch (a.obs_elem) {
    case 1: return launch_rollout_ge<G, uint8_t>(hL, a, B, R, st);
    case 2: return launch_rollout_ge<G, uint16_t>(hL, a, B, R, st);
    case 4: return launch_rollout_ge<G, float>(hL, a, B, R, st);
    default: return cudaErrorInvalidValue;
    }
}

bool rollout_available(const Layout &hL, const LaunchArgs &a)
{
    return (a.force_kernel >= 0 ? a.force_kernel : pick_kernel(hL, a.N, a.obs_elem)) == 1 && a.lanes >= 1 &&
           a.lanes <= 8 && a.tile_const[lanes_index(a.lanes)] && rollout_smem_bytes(hL, a.lanes, a.obs_elem) <= 200 * 1024;
}

cudaError_t launch_rollout(const Layout &hL, const LaunchArgs &a, int T, const int32_t *actions, int A, uint8_t *obs,
                           int32_t *rewards, int32_t *done, int O, cudaStream_t st, uint32_t seed, uint32_t first_step)
{
    if (!rollout_available(hL, a)) return cudaErrorInvalidValue;
    Buffers B;
    B.players = (uint2 *)a.players; B.objs = a.objs; B.timestep = a.timestep; B.actions = a.actions;
    B.rewards = a.rewards; B.done = a.done; B.obs = a.obs; B.ep_return = a.ep_return; B.stats = a.stats; B.N = a.N;
    B.P = hL.P;
    B.trace = nullptr;
    RolloutArgs R;
    R.actions = actions; R.T = T; R.A = A;
    R.sample = actions ? 0 : 1; R.seed = seed; R.first_step = first_step;
    R.obs = obs ? obs : a.obs;            R.O_obs = obs ? O : 1;
    R.rewards = rewards ? rewards : a.rewards; R.O_rew = rewards ? O : 1;
    R.done = done ? done : a.done;        R.O_done = done ? O : 1;
    switch (a.lanes) {
    case 1: return launch_rollout_g<1>(hL, a, B, R, st);
    case 2: return launch_rollout_g<2>(hL, a, B, R, st);
    case 4: return launch_rollout_g<4>(hL, a, B, R, st);
    default: return launch_rollout_g<8>(hL, a, B, R, st);
    }
}

cudaError_t launch_step(const Layout &hL, const Layout *dL, const LaunchArgs &a, int init_only, cudaStream_t st)
{
    Buffers B;
    B.players = (uint2 *)a.players; B.objs = a.objs; B.timestep = a.timestep; B.actions = a.actions;
    B.rewards = a.rewards; B.done = a.done; B.obs = a.obs; B.ep_return = a.ep_return; B.stats = a.stats; B.N = a.N;
    B.P = hL.P;
    B.trace = a.trace;
    if ((a.force_kernel >= 0 ? a.force_kernel : pick_kernel(hL, a.N, a.obs_elem)) == 1)
        return launch_tile(hL, a, B, init_only, st);
    const int nbuf = world_cta_buffers(hL, a.obs_elem);
    if (!nbuf) return cudaErrorInvalidValue;   // (refused at create)
    const size_t smem = world_cta_smem_bytes(hL, a.obs_elem, nbuf);
#define B200_WORLD_CTA(E, NBUF)                                                                      \
    do {                                                                                             \
        if (cudaError_t e = ensure_dynamic_smem(step_world_cta<E, NBUF>, (int)world_cta_smem_most())) return e; \
        step_world_cta<E, NBUF><<<a.N, WORLD_CTA_THREADS, smem, st>>>(dL, B, init_only);             \
    } while (0)
    if (a.obs_elem == 1) { if (nbuf == 2) B200_WORLD_CTA(uint8_t, 2); else B200_WORLD_CTA(uint8_t, 1); }
    else if (a.obs_elem == 2) { if (nbuf == 2) B200_WORLD_CTA(uint16_t, 2); else B200_WORLD_CTA(uint16_t, 1); }
    else if (a.obs_elem == 4) { if (nbuf == 2) B200_WORLD_CTA(float, 2); else B200_WORLD_CTA(float, 1); }
    else return cudaErrorInvalidValue;
#undef B200_WORLD_CTA
    return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------------------------
// Compat export: the reference's padded observation rows from the native tensor.  One warp per
// (world, cell): its P rows of 336 bytes are contiguous in the destination (4 B per lane and pass, coalesced); the
// sources are P runs of C bytes.  Off the hot path by design.
// ---------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) export_raw_kernel(const uint8_t *__restrict__ obs, uint8_t *__restrict__ dst,
                                                         const uint32_t N, const int W, const int H, const int P,
                                                         const int C)
{
    const int lane = threadIdx.x & 31;
    const size_t S = (size_t)W * H;
    const size_t pair = (size_t)blockIdx.x * 8 + (threadIdx.x >> 5);   // world * S + cell
    if (pair >= (size_t)N * S) return;
    const size_t n = pair / S;
    const int cell = (int)(pair % S), x = cell % W, y = cell / W;
    uint32_t *out = reinterpret_cast<uint32_t *>(dst + pair * (size_t)P * RAW_ROW_BYTES);
    const int words = P * (RAW_ROW_BYTES / 4);
    for (int i = lane; i < words; i += 32) {
        const int p = i / (RAW_ROW_BYTES / 4), c0 = 4 * (i % (RAW_ROW_BYTES / 4));
        const uint8_t *src = obs + ((((size_t)p * N + n) * W + x) * H + y) * C;
        uint32_t v = 0;
#pragma unroll
        for (int b = 0; b < 4; b++)
            if (c0 + b < C) v |= (uint32_t)src[c0 + b] << (8 * b);
        out[i] = v;
    }
}

cudaError_t launch_export_raw(const Layout &hL, const LaunchArgs &a, uint8_t *dst, cudaStream_t st)
{
    if (a.obs_elem != 1) return cudaErrorInvalidValue;
    const size_t pairs = (size_t)a.N * hL.S;
    export_raw_kernel<<<(unsigned)((pairs + 7) / 8), 256, 0, st>>>(a.obs, dst, a.N, hL.W, hL.H, hL.P, hL.C);
    return cudaGetLastError();
}

}   // namespace oc
}   // namespace b200
