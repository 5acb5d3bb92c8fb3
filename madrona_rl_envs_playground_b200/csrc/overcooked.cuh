// Overcooked world step + lossless observation featurisation — device code shared by the step kernels.
//
// Follows the semantics of the reference's 19-node task graph (src/overcooked_env/sim.cpp:498-537) but is NOT a
// translation of it: every agent is a lane / thread that computes its part of the post-step state directly (no ECS
// tables, no per-cell scratch, no atomics on world state), and the observation is re-encoded from that state instead of
// being patched incrementally.  Equivalences used (each proved by the parity tests against the reference CPU build):
//   * counter/pot interactions on one cell are applied in ascending agent id and interactions on different cells
//     commute, so "rank among the agents facing the same cell, one rank per round" replaces setup_interact_time + 4
//     ordered rounds (sim.cpp:259-358); source / serving interactions touch only the agent's own hands;
//   * the collision rule is world-global: nobody moves if any two agents propose the same cell or two agents swap
//     (sim.cpp:383-396,417-423), independent of the order future_player was written;
//   * obs == from-scratch encode of the post-step (post-reset) state (sim.cpp:68-167).
#pragma once

#include "common.cuh"

namespace b200 {
namespace oc {

enum : int { NORTH = 0, SOUTH = 1, EAST = 2, WEST = 3, STAY = 4, INTERACT = 5 };                      // sim.hpp:40
enum : int { AIR = 0, POT, COUNTER, ONION_SOURCE, TOMATO_SOURCE, DISH_SOURCE, SERVING };              // sim.hpp:41
enum : int { NONE = 0, TOMATO, ONION, DISH, SOUP };                                                   // sim.hpp:42

// Object{name,num_onions,num_tomatoes,cooking_tick} (sim.hpp:59-74) packed little-endian into one 32-bit word.
constexpr uint32_t OBJ_NONE = 0xFF000000u;   // {NONE,0,0,-1}
__device__ __forceinline__ uint32_t obj_make(int name) { return OBJ_NONE | (uint32_t)name; }
__device__ __forceinline__ int obj_name(uint32_t o) { return o & 0xFF; }
__device__ __forceinline__ int obj_onions(uint32_t o) { return (o >> 8) & 0xFF; }
__device__ __forceinline__ int obj_tomatoes(uint32_t o) { return (o >> 16) & 0xFF; }
__device__ __forceinline__ int obj_tick(uint32_t o) { return (int)(int8_t)(o >> 24); }
__device__ __forceinline__ uint32_t obj_set_tick(uint32_t o, int t) { return (o & 0x00FFFFFFu) | ((uint32_t)(t & 0xFF) << 24); }
__device__ __forceinline__ int obj_recipe(uint32_t o) { return 4 * obj_onions(o) + obj_tomatoes(o); }  // sim.hpp:70-73

// Static layout, identical for every world (reference: Overcooked::Config, sim.hpp:44-57).  One copy per simulator in
// global memory; every CTA stages it into shared memory first, because the tables are indexed with per-lane (divergent)
// cell numbers — a constant-bank lookup would serialise those, shared memory serves them in one pass.
struct alignas(16) Layout {
    int32_t W, H, S, P, C;          // C = 5P + 16 channels (5P + 10 under the Simplecooked rules)
    int32_t K;                      // object-capable cells (COUNTER or POT): the only cells that can hold an Object
    int32_t npots;
    int32_t horizon;
    int32_t placement_rew, soup_rew;   // already narrowed to uint8 like WorldState (sim.hpp:93-95)
    // 0 = src/overcooked_env; 1 = src/overcooked2_env ("Simplecooked", the JS-compatible rules: 5P + 10 channels,
    // one sequential interaction pass, pots start by themselves when full, dish-pickup shaping).  Terrain values are
    // stored in THIS file's enum order for both (overcooked2's own order, sim.hpp:40, is translated at create time).
    int32_t rules;
    int32_t dish_rew;                  // Simplecooked only (overcooked_env accepts and ignores it, sim.cpp:240-243)
    // per cell (row-major y*W+x), one word so that a divergent lookup is a single LDS:
    //   bits 0-7 TerrainT, bits 8-15 object slot (0xFF = cell cannot hold an object),
    //   bits 16-31 index of the cell's C-vector inside a [W][H][C] slab = x*H + y (the element offset (x*H + y) * C
    //   reaches 254 * 336 = 85 344 with 64 players on a 255-cell grid and would not fit in 16 bits)
    uint32_t cell[256];
    uint8_t cell_of_slot[256];      // object slot -> cell
    uint8_t pot_slot[256];          // slots whose terrain is POT, in ascending cell order (sim.cpp:588-595)
    uint8_t start_pos[64];          // start_y*W + start_x (sim.cpp:474)
    uint8_t recipe_values[16], recipe_times[16];

    __device__ __forceinline__ int terrain(int c) const { return (int)(cell[c] & 0xFFu); }
    __device__ __forceinline__ int slot_of_cell(int c) const { return (int)((cell[c] >> 8) & 0xFFu); }
    __host__ __device__ __forceinline__ int cell_off(int c) const { return (int)(cell[c] >> 16) * C; }
};
static_assert(sizeof(Layout) % 16 == 0, "Layout is moved with 16-byte bulk copies");

__device__ __forceinline__ int move_in_direction(int point, int dir, int W)   // sim.cpp:185-197
{
    return dir == NORTH ? point - W : dir == SOUTH ? point + W : dir == EAST ? point + 1 : dir == WEST ? point - 1 : point;
}
__device__ __forceinline__ bool is_cooking(const Layout &L, uint32_t o)        // sim.cpp:58-61
{
    const int t = obj_tick(o);
    return t >= 0 && t < (int)L.recipe_times[obj_recipe(o)];
}
__device__ __forceinline__ bool is_ready(const Layout &L, uint32_t o)          // sim.cpp:63-66
{
    const int t = obj_tick(o);
    return t >= 0 && t >= (int)L.recipe_times[obj_recipe(o)];
}
__device__ __forceinline__ int relative_player(int other, int observer)         // sim.cpp:122-145
{
    return other == observer ? 0 : (other < observer ? other + 1 : other);
}

// ---- observation bytes (sim.cpp:68-167), written into one [W][H][C] slab in shared memory ----
//
// The slab's element type E is the observation tensor's: uint8_t (the reference's int8 export), uint16_t (bfloat16 bits)
// or float — "policy-ready" observations: the value the consumer's `.float()` of the int8 tensor
// would hold (pantheonrl_extension/vectoragent.py:353), written by the step kernel itself.  E is deduced from `slab`.
template <class E> __device__ __forceinline__ E obs_val(int v);
template <> __device__ __forceinline__ uint8_t obs_val<uint8_t>(int v) { return (uint8_t)v; }
template <> __device__ __forceinline__ uint16_t obs_val<uint16_t>(int v)
{
    return (uint16_t)(__float_as_uint((float)(int)(int8_t)v) >> 16);   // bfloat16 holds every int8 value exactly
}
template <> __device__ __forceinline__ float obs_val<float>(int v) { return (float)(int)(int8_t)v; }

// channels 5P+6 .. 5P+14 contributed by the Object lying on cell `cell` (terrain POT or COUNTER).  SET = false writes
// zeros to exactly the bytes SET = true would write (the fused rollout kernel keeps its staging tile across steps and
// erases what went stale).
// Simplecooked (src/overcooked2_env/sim.cpp:79-103): 5P+5 onions of the soup in a pot, 5P+6 its cooking tick (0 while
// idle), 5P+7 a soup anywhere else, 5P+8 dish, 5P+9 onion; tomatoes have no channel.
template <bool SET = true, class E>
__device__ __forceinline__ void encode_cell_object(const Layout &L, E *slab, int cell, uint32_t o)
{
    const int name = obj_name(o);
    if (name == NONE) return;
    E *x = slab + L.cell_off(cell) + 5 * L.P;
    if (L.rules) {
        if (name == SOUP) {
            if (L.terrain(cell) == POT) {
                x[5] = SET ? obs_val<E>(obj_onions(o)) : obs_val<E>(0);
                x[6] = SET ? obs_val<E>(max(obj_tick(o), 0)) : obs_val<E>(0);
            } else {
                x[7] = obs_val<E>(SET ? 1 : 0);
            }
        } else if (name != TOMATO) {
            x[name == DISH ? 8 : 9] = obs_val<E>(SET ? 1 : 0);
        }
        return;
    }
    if (name == SOUP) {
        if (L.terrain(cell) == POT) {
            if (obj_tick(o) < 0) {
                x[6] = SET ? obs_val<E>(obj_onions(o)) : obs_val<E>(0);
                x[7] = SET ? obs_val<E>(obj_tomatoes(o)) : obs_val<E>(0);
            } else {
                x[8] = SET ? obs_val<E>(obj_onions(o)) : obs_val<E>(0);
                x[9] = SET ? obs_val<E>(obj_tomatoes(o)) : obs_val<E>(0);
                x[10] = SET ? obs_val<E>(((int)L.recipe_times[obj_recipe(o)] - obj_tick(o))) : obs_val<E>(0);
                if (is_ready(L, o)) x[11] = obs_val<E>(SET ? 1 : 0);
            }
        } else {
            x[8] = SET ? obs_val<E>(obj_onions(o)) : obs_val<E>(0);
            x[9] = SET ? obs_val<E>(obj_tomatoes(o)) : obs_val<E>(0);
            x[10] = obs_val<E>(0);
            x[11] = obs_val<E>(SET ? 1 : 0);
        }
    } else {
        x[name == DISH ? 12 : name == ONION ? 13 : 14] = obs_val<E>(SET ? 1 : 0);
    }
}

// channels contributed by the object held by the player standing on `cell` (observer independent)
template <bool SET = true, class E>
__device__ __forceinline__ void encode_held_object(const Layout &L, E *slab, int cell, uint32_t o)
{
    const int name = obj_name(o);
    if (name == NONE) return;
    E *x = slab + L.cell_off(cell) + 5 * L.P;
    if (L.rules) {   // src/overcooked2_env/sim.cpp:131-140
        if (name != TOMATO) x[name == SOUP ? 7 : name == DISH ? 8 : 9] = obs_val<E>(SET ? 1 : 0);
        return;
    }
    if (name == SOUP) {
        x[8] = SET ? obs_val<E>(obj_onions(o)) : obs_val<E>(0);
        x[9] = SET ? obs_val<E>(obj_tomatoes(o)) : obs_val<E>(0);
        x[10] = obs_val<E>(0);
        x[11] = obs_val<E>(SET ? 1 : 0);
    } else {
        x[name == DISH ? 12 : name == ONION ? 13 : 14] = obs_val<E>(SET ? 1 : 0);
    }
}

// the two observer-relative bytes of player q as seen by `observer`: presence one-hot and orientation one-hot
template <class E>
__device__ __forceinline__ void encode_player(const Layout &L, E *slab, int observer, int q, int cell, int orient,
                                              int value)
{
    const int r = relative_player(q, observer);
    E *x = slab + L.cell_off(cell);
    x[r] = obs_val<E>(value);
    x[L.P + 4 * r + orient] = obs_val<E>(value);
}
// same with the cell's C-vector already located (x = slab + cell_off) and `po` = P + orient
template <class E>
__device__ __forceinline__ void encode_player_at(E *x, int observer, int q, int po, int value)
{
    const int r = relative_player(q, observer);
    x[r] = obs_val<E>(value);
    x[po + 4 * r] = obs_val<E>(value);
}

}   // namespace oc
}   // namespace b200
