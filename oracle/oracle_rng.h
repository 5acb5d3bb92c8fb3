/* TEST INFRASTRUCTURE — per-world RNG of the reference, restated.
 * Follows src/cartpole_env/rng.hpp:7-36 (identical copies in balance_beam_env/ and hanabi_env/):
 * 8 TEA rounds hash the episode index into the seed; rand() is a 32-bit LCG, low 24 bits scaled to [0,1). */
#ifndef B200ENVS_ORACLE_RNG_H
#define B200ENVS_ORACLE_RNG_H
#include <stdint.h>

static inline uint32_t oracle_rng_make(uint32_t episode_idx)
{
    uint32_t v0 = episode_idx, v1 = 0, s0 = 0;
    for (int n = 0; n < 8; n++) {
        s0 += 0x9e3779b9u;
        v0 += ((v1 << 4) + 0xa341316cu) ^ (v1 + s0) ^ ((v1 >> 5) + 0xc8013ea4u);
        v1 += ((v0 << 4) + 0xad90777du) ^ (v0 + s0) ^ ((v0 >> 5) + 0x7e95761eu);
    }
    return v0;
}

static inline float oracle_rng_rand(uint32_t *v)
{
    *v = 1664525u * (*v) + 1013904223u;
    return (float)(*v & 0x00FFFFFFu) / (float)0x01000000;
}

#endif
