// Per-SM store rates, measured once per device (sm_rates.cu).
//
// Not every SM stores equally fast: on a B200 32 of the 148 SMs (ids 8-15, 24-31, 72-79, 128-135, the same on every
// box) move ~43 GB/s into memory when all SMs store, the others 52-62 (tools/probes/sm_store_rate_probe.cu).  A kernel
// that gives every SM the same share of a store-bound job ends when the slowest SM does: 148 x 43 = 6.4 TB/s, where
// the SMs together could take 7.6.  The fused rollouts therefore deal tiles to SMs in proportion to these rates.
#pragma once
#include <cstddef>
#include <cstdint>
#include <cuda_runtime.h>

namespace b200 {

// rate[%smid] relative to the mean of the device's SMs (0 = no such SM), or nullptr when nothing (sensible) was
// measured.  The first call per device measures, with `scratch` — device memory the caller is about to overwrite
// anyway, left zero-filled — as the target (at least 128 KB per SM, else nothing is measured); synchronises the device.
const float *sm_store_rates(void *scratch, size_t scratch_bytes, cudaError_t *err);

// shares of `tiles` in proportion to rate ^ exponent, by cumulative rounding over the SM ids (they sum to `tiles`
// before `cap` cuts; callers hand out what the cap cut off some other way).  Returns false when rates is nullptr.
bool sm_quotas(const float *rates, uint32_t tiles, uint32_t cap, double exponent, uint32_t quota[256]);

}   // namespace b200
