/* TEST INFRASTRUCTURE — CPU oracle for the batched env step path.  NOT a product path.
 *
 * Plain-C restatement of the reference's per-step world update + observation featurisation
 * (willwng/madrona_rl_envs_playground, src/<env>/sim.cpp) for Overcooked, Hanabi, Balance Beam and
 * Cartpole.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may load this
 * library; the product (libb200envs.so) never links, imports or calls it.
 *
 * Parity status: PINNED.  Every function here is checked bit-for-bit (Cartpole: bit-for-bit on x86 with
 * -ffp-contract=off) against the reference's own Madrona CPU exec mode compiled from /root/reference
 * (oracle/_ref, see oracle/Makefile) by tests/test_oracle.py (the *_vs_reference_live tests), and against the committed golden
 * vectors under tests/golden/ that were generated from that reference build (tests/golden/generate.py).
 *
 * Canonical array layouts used by the oracle (and by the CUDA library), all C-contiguous:
 *   Overcooked  actions i32 [P][N]; obs u8 [P][N][W][H][C] (C = 5P+16; element [p][n][x][y][c] is cell y*W+x,
 *               the layout envs/overcooked_env.py:96 hands to the policy); rewards i32 [P][N]; done i32 [N]
 *   Hanabi      actions i32 [2][N]; obs u8 [2][N][658]; state u8 [2][N][783]; mask i32 [2][N][20];
 *               active i32 [2][N]; rewards f32 [2][N]; done i32 [N]
 *   Balance     actions i32 [2][N]; obs i32 [2][N][7]; rewards f32 [2][N]; done i32 [N]
 *   Cartpole    actions i32 [N]; state f32 [N][4]; rewards f32 [N]; done i32 [N]
 * Episode indices for the RNG envs follow the reference with ONE worker thread: a single counter, worlds
 * visited in ascending order inside a step (external/madrona/src/mw/cpu_exec.cpp:211-236 with numWorkers=1).
 */
#ifndef B200ENVS_ORACLE_H
#define B200ENVS_ORACLE_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- Overcooked (src/overcooked_env/sim.hpp:44-57 Config) ---- */
typedef struct {
    int32_t height, width, num_players;
    uint8_t terrain[256];                 /* TerrainT per cell, row-major y*W+x; S = H*W <= 255 (sim.hpp:86) */
    uint8_t start_x[64], start_y[64];
    int32_t placement_in_pot_rew, dish_pickup_rew, soup_pickup_rew;
    uint8_t recipe_values[16], recipe_times[16];
    int32_t horizon;
} OracleOvercookedCfg;

void *oracle_overcooked_create(const OracleOvercookedCfg *cfg, uint32_t num_worlds);
void oracle_overcooked_destroy(void *h);
/* one step for all worlds; obs may be NULL (skips featurisation output copy) */
void oracle_overcooked_step(void *h, const int32_t *actions, uint8_t *obs, int32_t *rewards, int32_t *done);
/* featurise the current state (used for the initial observation) */
void oracle_overcooked_observe(void *h, uint8_t *obs);

/* ---- Overcooked2 / "Simplecooked" (src/overcooked2_env: same Config; terrain values in ITS enum order, sim.hpp:40;
 *      obs u8 [P][N][W][H][5P+10]) ---- */
void *oracle_overcooked2_create(const OracleOvercookedCfg *cfg, uint32_t num_worlds);
void oracle_overcooked2_destroy(void *h);
void oracle_overcooked2_step(void *h, const int32_t *actions, uint8_t *obs, int32_t *rewards, int32_t *done);
void oracle_overcooked2_observe(void *h, uint8_t *obs);

/* ---- Hanabi (2 players, src/hanabi_env/sim.hpp:13-30) ---- */
#define ORACLE_HANABI_OBS 658
#define ORACLE_HANABI_STATE 783
#define ORACLE_HANABI_MOVES 20
void *oracle_hanabi_create(uint32_t num_worlds, uint32_t colors, uint32_t ranks, uint32_t players,
                           uint32_t max_information_tokens, uint32_t max_life_tokens);
void oracle_hanabi_destroy(void *h);
int oracle_hanabi_max_info_encoded(void *h);   /* test aid: > max_information_tokens once a long (spilling) row was encoded */
void oracle_hanabi_step(void *h, const int32_t *actions);
/* copy out the exported tensors in canonical layout; any pointer may be NULL */
void oracle_hanabi_read(void *h, uint8_t *obs, uint8_t *state, int32_t *mask, int32_t *active,
                        float *rewards, int32_t *done);

/* ---- Balance Beam (src/balance_beam_env/sim.cpp) ---- */
void *oracle_balance_create(uint32_t num_worlds);
void oracle_balance_destroy(void *h);
void oracle_balance_step(void *h, const int32_t *actions);
void oracle_balance_read(void *h, int32_t *obs, float *rewards, int32_t *done);

/* ---- Cartpole (src/cartpole_env/sim.cpp) ---- */
void *oracle_cartpole_create(uint32_t num_worlds);
void oracle_cartpole_destroy(void *h);
void oracle_cartpole_step(void *h, const int32_t *actions);
void oracle_cartpole_read(void *h, float *state, float *rewards, int32_t *done);
/* overwrite the dynamic state (parity tests re-seed the oracle from the GPU state each step, as the
 * reference's own validate_step does, envs/cartpole_env.py:246-288) */
void oracle_cartpole_write_state(void *h, const float *state);

/* ---- Acrobat (src/acrobat_env/sim.cpp) ---- state f32 [N][4] = theta1, theta2, omega1, omega2 */
void *oracle_acrobat_create(uint32_t num_worlds);
void oracle_acrobat_destroy(void *h);
void oracle_acrobat_step(void *h, const int32_t *actions);
void oracle_acrobat_read(void *h, float *state, float *rewards, int32_t *done);
void oracle_acrobat_write_state(void *h, const float *state);
uint32_t oracle_acrobat_episode_length(void *h);   /* the Manager-wide counter, see oracle_acrobat.c */

#ifdef __cplusplus
}
#endif
#endif
