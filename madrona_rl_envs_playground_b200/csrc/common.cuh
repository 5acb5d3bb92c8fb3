// Shared device helpers for the env step kernels (sm_100a).
#pragma once

#include <cstdint>
#include <cstdlib>
#include <mutex>
#include <cuda_runtime.h>

namespace b200 {

// ---------------------------------------------------------------------------------------------------------------
// Per-world RNG.  Bit-for-bit the reference generator (src/cartpole_env/rng.hpp:7-36, shared by balance_beam_env and
// hanabi_env): the episode index is hashed by 8 TEA rounds into the LCG state; rand() keeps the low 24 bits.
// Counter-based at the episode level: seed = f(episode index), so a world can be re-seeded without any stored stream.
// ---------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t rng_make(uint32_t episode_idx)
{
    uint32_t v0 = episode_idx, v1 = 0, s0 = 0;
#pragma unroll
    for (int n = 0; n < 8; n++) {
        s0 += 0x9e3779b9u;
        v0 += ((v1 << 4) + 0xa341316cu) ^ (v1 + s0) ^ ((v1 >> 5) + 0xc8013ea4u);
        v1 += ((v0 << 4) + 0xad90777du) ^ (v0 + s0) ^ ((v0 >> 5) + 0x7e95761eu);
    }
    return v0;
}

__device__ __forceinline__ float rng_rand(uint32_t &v)
{
    v = 1664525u * v + 1013904223u;
    return __uint2float_rn(v & 0x00FFFFFFu) / 16777216.0f;   // exact: 24-bit integer scaled by 2^-24
}

// ---------------------------------------------------------------------------------------------------------------
// Ordered episode numbering.
//
// The reference seeds every new episode with a Manager-wide fetch_add (src/cartpole_env/sim.cpp:51-53).  Its
// deterministic order (one CPU worker) numbers the worlds that reset inside one step in ascending world id, after all
// episodes of earlier steps.  On the GPU that is: episode = base + exclusive prefix count of resetting worlds.
// Single pass over CTA tiles, two levels, and only the tiles that NEED a number wait:
//   * a tile takes a ticket (atomicAdd) so that every lower-numbered tile is already running or done;
//   * it publishes its reset count (one 32-bit status word, epoch:16 | count:16) and carries on with whatever does
//     not depend on episode numbers;
//   * a tile with resets sums, with one 128-bit load per lane and level, the counts of the earlier tiles of its group
//     of 128 and the sums of the earlier groups (published by whichever tile completes a group's atomic
//     accumulator).  All loads of a tile are independent, so the wait is one memory round trip once the predecessors
//     have published (a look-back chain over 2 048 tiles cost up to 64 dependent round trips on EVERY tile);
//   * status words carry the step epoch, so nothing is cleared between launches (each word is rewritten every
//     launch, so a 16-bit tag cannot alias);
//   * the tile holding the last ticket folds the step total into the global episode counter, bumps the epoch and
//     rewinds the ticket counter — every other tile has read both values before it published, and this tile waits
//     on all of them.
// ---------------------------------------------------------------------------------------------------------------
struct EpisodeScan {
    unsigned long long *status;   // 8 * (num_tiles + 16) bytes: u32 tile status [num_tiles], group sums, 2 x accumulators
    uint32_t *ticket;             // next tile index of the running launch
    uint32_t *epoch;              // launch counter (device resident so CUDA-graph replays advance it)
    uint32_t *cur_episode;        // Manager-wide episode counter (init.hpp EpisodeManager::curEpisode)
};

constexpr uint32_t SCAN_GROUP = 128;   // tiles per group = 4 status words per lane
#ifndef B200_POLL_SLEEP_NS
#define B200_POLL_SLEEP_NS 0   // back-off between two polls of a status word that has not been published yet
#endif

// sum of the counts in words [first, end) of a status array, 128 words per warp round; spins until each carries `tag`
__device__ __forceinline__ uint32_t scan_sum_lane(const uint32_t *words, uint32_t first, uint32_t end, uint32_t tag,
                                                  unsigned lane)
{
    uint32_t sum = 0;
    for (uint32_t i0 = first + 4u * lane; i0 < end; i0 += SCAN_GROUP) {   // first is a multiple of 4: aligned loads
        uint32_t v[4];
        bool ok;
        do {
            asm volatile("ld.volatile.global.v4.u32 {%0, %1, %2, %3}, [%4];"
                         : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3])
                         : "l"(words + i0)
                         : "memory");
            ok = true;
#pragma unroll
            for (int c = 0; c < 4; c++) ok = ok && (i0 + c >= end || (v[c] >> 16) == tag);
#if B200_POLL_SLEEP_NS > 0
            if (!ok) __nanosleep(B200_POLL_SLEEP_NS);
#endif
        } while (!ok);
#pragma unroll
        for (int c = 0; c < 4; c++) sum += (i0 + c < end) ? (v[c] & 0xFFFFu) : 0u;
    }
    return sum;
}

// Block-level context for one tile.  Call begin() first thing in the kernel (all threads), then
// exclusive_base(tile_count) from ALL threads once the tile's reset count is known (tile_count must be block-uniform
// and at most 256).
struct EpisodeScanTile {
    uint32_t tile, epoch, base;

    __device__ __forceinline__ void begin(const EpisodeScan &s, uint32_t *smem3)
    {
        if (threadIdx.x == 0) {
            smem3[0] = atomicAdd(s.ticket, 1u);
            smem3[1] = *(volatile uint32_t *)s.epoch;
            smem3[2] = *(volatile uint32_t *)s.cur_episode;
        }
        __syncthreads();
        tile = smem3[0];
        epoch = smem3[1];
        base = smem3[2];
    }

    // Publishes the tile's reset count (all threads call; tile_count block-uniform): its own status word for the
    // later tiles of its group, and one atomicAdd (arrivals:16 | count:16) into the group's accumulator — the tile
    // whose add completes the group publishes the group sum, so nobody polls for it.  Accumulators alternate with
    // the launch parity; the completing tile zeroes the other parity's for the next launch.
    __device__ __forceinline__ void publish(const EpisodeScan &s, uint32_t tile_count, uint32_t num_tiles)
    {
        if (threadIdx.x == 0) publish_count(s, tile_count, num_tiles);
    }
    // the same from ONE thread of the caller's choice
    __device__ __forceinline__ void publish_count(const EpisodeScan &s, uint32_t tile_count, uint32_t num_tiles)
    {
        uint32_t *status = reinterpret_cast<uint32_t *>(s.status);
        const uint32_t tag = epoch & 0xFFFFu, g = tile / SCAN_GROUP, groups = (num_tiles + SCAN_GROUP - 1) / SCAN_GROUP;
        uint32_t *gsum = status + ((num_tiles + 3u) & ~3u), *acc = gsum + ((groups + 3u) & ~3u);
        *(volatile uint32_t *)(status + tile) = (tag << 16) | tile_count;
        const uint32_t old = atomicAdd(acc + (epoch & 1u) * groups + g, (1u << 16) | tile_count);
        if ((old >> 16) + 1u == min(SCAN_GROUP, num_tiles - g * SCAN_GROUP)) {
            *(volatile uint32_t *)(gsum + g) = (tag << 16) | ((old & 0xFFFFu) + tile_count);
            *(volatile uint32_t *)(acc + ((epoch + 1u) & 1u) * groups + g) = 0u;
        }
    }

    // After publish(), ONE WARP only (all its lanes): the episode index of the first resetting world of this tile, in every lane
    // (returns at once when the tile has no resets).  A kernel with other work to do calls this as late as it can:
    // every predecessor has published by then and the sum is one memory round trip with no polling.
    __device__ __forceinline__ uint32_t exclusive_base_warp(const EpisodeScan &s, uint32_t tile_count, uint32_t num_tiles)
    {
        const bool last = tile == num_tiles - 1;
        if (!tile_count && !last) return base;
        const unsigned lane = threadIdx.x & 31u;
        const uint32_t *status = reinterpret_cast<const uint32_t *>(s.status);
        const uint32_t tag = epoch & 0xFFFFu, g = tile / SCAN_GROUP;
        uint32_t mine = scan_sum_lane(status, g * SCAN_GROUP, tile, tag, lane) +
                        scan_sum_lane(status + ((num_tiles + 3u) & ~3u), 0, g, tag, lane);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) mine += __shfl_xor_sync(0xffffffffu, mine, o);
        if (last && lane == 0) {
            *(volatile uint32_t *)s.cur_episode = base + mine + tile_count;
            *(volatile uint32_t *)s.epoch = epoch + 1;
            *(volatile uint32_t *)s.ticket = 0;
            __threadfence();
        }
        return base + mine;
    }

    // The same for a whole block (all threads call; tile_count block-uniform).
    __device__ __forceinline__ uint32_t exclusive_base(const EpisodeScan &s, uint32_t tile_count, uint32_t num_tiles,
                                                       uint32_t *smem1)
    {
        if (!tile_count && tile != num_tiles - 1) return base;
        if (threadIdx.x < 32) {
            const uint32_t e = exclusive_base_warp(s, tile_count, num_tiles);
            if (threadIdx.x == 0) smem1[0] = e;
        }
        __syncthreads();
        return smem1[0];
    }
};

// ---------------------------------------------------------------------------------------------------------------
// Episode numbering inside a fused multi-step rollout (one cooperative launch, all tiles resident, tile = blockIdx.x):
// number = launch base + resets of all tiles in earlier steps + resets of lower tiles in this step.  Per step a tile
// publishes its count (status word + group accumulator as in EpisodeScan, tagged with the global step index, in a ring
// of ROLL_RING step slots) and, before it resets, reads the previous step's group sums (=> every tile has finished
// step s-1: tiles never drift more than one step apart, which is what makes a ring of 4 slots safe) and this step's
// lower tiles / groups.
// ---------------------------------------------------------------------------------------------------------------
constexpr uint32_t ROLL_RING = 4;
struct RolloutScan {
    uint32_t *ring;          // [ROLL_RING][tiles4 + 2 * groups4] words: status per tile, sum per group, accumulator per group
    uint32_t *step_base;     // global step index of the next rollout step (tags never repeat within the ring)
    uint32_t *cur_episode;   // Manager-wide episode counter (shared with EpisodeScan)
};
struct RollSlot { uint32_t *status, *gsum, *acc; };
__device__ __forceinline__ RollSlot roll_slot(const RolloutScan &S, uint32_t num_tiles, uint32_t global_step)
{
    const uint32_t tiles4 = (num_tiles + 3u) & ~3u;
    const uint32_t groups4 = ((num_tiles + SCAN_GROUP - 1) / SCAN_GROUP + 3u) & ~3u;
    uint32_t *base = S.ring + (size_t)(global_step % ROLL_RING) * (tiles4 + 2 * groups4);
    return {base, base + tiles4, base + tiles4 + groups4};
}
inline size_t roll_ring_words(uint32_t tiles)
{
    const uint32_t tiles4 = (tiles + 3u) & ~3u;
    const uint32_t groups4 = ((tiles + SCAN_GROUP - 1) / SCAN_GROUP + 3u) & ~3u;
    return (size_t)ROLL_RING * (tiles4 + 2 * groups4) + 16;   // + slack for the 128-bit loads past the last word
}
// ONE thread: publish this tile's reset count of global step `gstep`
__device__ __forceinline__ void roll_publish(const RolloutScan &S, uint32_t num_tiles, uint32_t tile, uint32_t gstep,
                                             uint32_t count)
{
    const uint32_t tag = (gstep + 1u) & 0xFFFFu, group = tile / SCAN_GROUP;
    const uint32_t group_size = min(SCAN_GROUP, num_tiles - group * SCAN_GROUP);
    const RollSlot now = roll_slot(S, num_tiles, gstep);
    *(volatile uint32_t *)(now.status + tile) = (tag << 16) | count;
    const uint32_t old = atomicAdd(now.acc + group, (1u << 16) | count);
    if ((old >> 16) + 1u == group_size) {   // this tile completes its group
        // the accumulator of two steps ahead (last used two steps ago) is zeroed BEFORE anybody can learn from the
        // group sum that this step is complete
        *(volatile uint32_t *)(roll_slot(S, num_tiles, gstep + 2u).acc + group) = 0u;
        __threadfence();
        *(volatile uint32_t *)(now.gsum + group) = (tag << 16) | ((old & 0xFFFFu) + count);
    }
}
// ONE warp (all 32 lanes): resets of all tiles in global step `gstep` (spins until every group has published)
__device__ __forceinline__ uint32_t roll_step_total(const RolloutScan &S, uint32_t num_tiles, uint32_t gstep, unsigned lane)
{
    const uint32_t groups = (num_tiles + SCAN_GROUP - 1) / SCAN_GROUP;
    uint32_t total = scan_sum_lane(roll_slot(S, num_tiles, gstep).gsum, 0, groups, (gstep + 1u) & 0xFFFFu, lane);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) total += __shfl_xor_sync(0xffffffffu, total, o);
    return total;
}
// ONE warp: resets of the tiles below `tile` in global step `gstep`
__device__ __forceinline__ uint32_t roll_lower_tiles(const RolloutScan &S, uint32_t num_tiles, uint32_t tile, uint32_t gstep,
                                                     unsigned lane)
{
    const uint32_t tag = (gstep + 1u) & 0xFFFFu, group = tile / SCAN_GROUP;
    const RollSlot now = roll_slot(S, num_tiles, gstep);
    uint32_t lower = scan_sum_lane(now.status, group * SCAN_GROUP, tile, tag, lane) + scan_sum_lane(now.gsum, 0, group, tag, lane);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) lower += __shfl_xor_sync(0xffffffffu, lower, o);
    return lower;
}

// Named barriers (producer / consumer between warps of a CTA): the producer arrives without waiting, the consumer
// syncs; the barrier orders the producer's earlier shared-memory writes before the consumer's later reads (no fence:
// a membar would also wait for the producer's outstanding GLOBAL stores).
__device__ __forceinline__ void bar_arrive(int id, int threads)
{
    asm volatile("bar.arrive %0, %1;" ::"r"(id), "r"(threads) : "memory");
}
__device__ __forceinline__ void bar_sync(int id, int threads)
{
    asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(threads) : "memory");
}

// ---------------------------------------------------------------------------------------------------------------
// Bulk shared->global stores (TMA engine, `cp.async.bulk`, SASS UBLKCP): the observation slabs are built in shared
// memory and drained by the copy engine, so no LSU instructions are spent streaming them out.
// ---------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void fence_proxy_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

__device__ __forceinline__ void bulk_store_s2g(void *dst, const void *src_smem, uint32_t bytes)
{
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(smem_u32(src_smem)),
                 "r"(bytes)
                 : "memory");
}
// the same with an L2 eviction policy (createpolicy) for the lines written
__device__ __forceinline__ void bulk_store_s2g_hint(void *dst, const void *src_smem, uint32_t bytes, uint64_t policy)
{
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group.L2::cache_hint [%0], [%1], %2, %3;" ::"l"(dst),
                 "r"(smem_u32(src_smem)), "r"(bytes), "l"(policy)
                 : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void bulk_wait_read()
{
    asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}

// ---------------------------------------------------------------------------------------------------------------
// Bulk global->shared loads completing on an mbarrier (TMA engine, SASS UBLKCP.S.G + SYNCS).
// ---------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t arrivals)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(arrivals) : "memory");
}
__device__ __forceinline__ void fence_mbar_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
// dst (shared), src (global) 16-byte aligned, bytes a multiple of 16
__device__ __forceinline__ void bulk_load_g2s(void *dst_smem, const void *src, uint32_t bytes, uint64_t *bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(dst_smem)),
                 "l"(src), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
// every calling thread arrives once (release at CTA scope: its earlier shared-memory writes are visible to whoever
// observes the phase complete)
__device__ __forceinline__ void mbar_arrive(uint64_t *bar)
{
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity)
{
    uint32_t ok;
    do {
        asm volatile(
            "{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(ok)
            : "r"(smem_u32(bar)), "r"(parity)
            : "memory");
    } while (!ok);
}

// ---------------------------------------------------------------------------------------------------------------
// Programmatic dependent launch: a step kernel launched with cudaLaunchAttributeProgrammaticStreamSerialization may
// begin (and run its state-independent prologue) while the previous kernel of the stream drains.
// grid_dependency_wait() returns once that kernel has completed and its writes are visible; it is a no-op when the
// launch carries no programmatic dependency.
// ---------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ void grid_dependency_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void grid_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

// Store `bytes` from shared `src` to global `dst` with a group of `nthreads` threads (`tid` in [0,nthreads)); thread 0 of
// the group issues and commits the bulk copy.  When dst and src are congruent mod 16 the 16-byte-aligned interior goes
// through the copy engine and the <=15-byte head / tail through byte stores; otherwise (odd world counts) everything
// is copied by byte.  Callers must have made their shared-memory writes visible to the group (barrier) beforehand and
// must wait (bulk_wait_read) before overwriting src.
template <bool HINT = false>
__device__ __forceinline__ void slab_store(uint8_t *dst, const uint8_t *src, uint32_t bytes, int tid, int nthreads,
                                           uint64_t policy = 0)
{
    const uint32_t dmis = (uint32_t)(reinterpret_cast<uintptr_t>(dst) & 15u);
    const uint32_t smis = smem_u32(src) & 15u;
    if (dmis != smis) {
        for (uint32_t i = tid; i < bytes; i += nthreads) dst[i] = src[i];
        if (tid == 0) bulk_commit();   // keep group accounting uniform
        return;
    }
    uint32_t head = (16u - dmis) & 15u;
    if (head > bytes) head = bytes;
    const uint32_t mid = (bytes - head) & ~15u;
    const uint32_t tail = bytes - head - mid;
    if ((uint32_t)tid < head) dst[tid] = src[tid];
    if ((uint32_t)tid < tail) dst[head + mid + tid] = src[head + mid + tid];
    if (tid == 0) {
        if (mid) {
            fence_proxy_async_smem();
            // one copy: splitting it in 2 / 4 pieces changes nothing
            if (HINT) bulk_store_s2g_hint(dst + head, src + head, mid, policy);
            else bulk_store_s2g(dst + head, src + head, mid);
        }
        bulk_commit();
    }
}

// Host: per-device facts the launch heuristics need (SM count, shared memory per SM and the per-CTA reservation),
// queried once per device; one process may drive simulators on several GPUs from several host threads.
struct DeviceInfo {
    int sms = 0;                 // cudaDevAttrMultiProcessorCount
    int smem_per_sm = 0;         // cudaDevAttrMaxSharedMemoryPerMultiprocessor (bytes)
    int smem_reserved = 1024;    // cudaDevAttrReservedSharedMemoryPerBlock
    int smem_optin = 0;          // cudaDevAttrMaxSharedMemoryPerBlockOptin
};
inline std::mutex &host_cache_mutex()
{
    static std::mutex m;
    return m;
}
inline const DeviceInfo &device_info()
{
    static DeviceInfo info[64];
    static bool have[64] = {};
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 0 || dev >= 64) dev = 0;
    std::lock_guard<std::mutex> lock(host_cache_mutex());
    if (!have[dev]) {
        DeviceInfo d;
        cudaDeviceGetAttribute(&d.sms, cudaDevAttrMultiProcessorCount, dev);
        cudaDeviceGetAttribute(&d.smem_per_sm, cudaDevAttrMaxSharedMemoryPerMultiprocessor, dev);
        cudaDeviceGetAttribute(&d.smem_reserved, cudaDevAttrReservedSharedMemoryPerBlock, dev);
        cudaDeviceGetAttribute(&d.smem_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
        if (d.sms <= 0) d.sms = 1;
        info[dev] = d;
        have[dev] = true;
    }
    return info[dev];
}

// Shared memory to ASK for so that at most ceil(grid / SMs) CTAs fit on an SM, when the grid fits on the GPU in one
// wave: under programmatic dependent launch the next grid's CTAs take whatever slots this grid's exiting warps free,
// and with room for 17 per SM some SMs ended up with 17 and others with 10 instead of an even 13.8.  `need` is the
// kernel's dynamic request, `fixed` its static shared memory.
inline size_t even_wave_smem(size_t need, size_t fixed, unsigned grid)
{
    const DeviceInfo &d = device_info();
    if (getenv("B200ENV_NO_EVEN_WAVE") || d.smem_per_sm <= 0) return need;
    const size_t per_cta = need + fixed + (size_t)d.smem_reserved;
    const unsigned per_sm = (grid + d.sms - 1) / d.sms;
    const size_t fit = (size_t)d.smem_per_sm / per_cta;          // CTAs per SM as it is
    if (per_sm < 4 || per_sm >= fit) return need;                // a handful of CTAs per SM: placement does not matter (and
                                                                 // a huge request would flip the SM's carve-out); more
                                                                 // than one wave; or already capped
    // per_sm + 1 CTAs of this size no longer fit
    const size_t cap = (size_t)d.smem_per_sm / (per_sm + 1) - (size_t)d.smem_reserved - fixed + 16;
    const size_t most = d.smem_optin > 0 ? (size_t)d.smem_optin - fixed : 200 * 1024;
    return cap > need ? (cap < most && cap < 200 * 1024 ? cap : need) : need;
}

// Host: opt a kernel into `bytes` of dynamic shared memory once per (kernel, device) — the attribute is per device,
// and one process may drive simulators on several GPUs.  (Kernels share pointer TYPES, so the cache is keyed on the
// pointer value.)
template <class Kernel>
inline cudaError_t ensure_dynamic_smem(Kernel kernel, int bytes)
{
    struct Seen { const void *fn; unsigned long long devices; };
    static Seen seen[64];
    static int nseen = 0;
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    const void *fn = reinterpret_cast<const void *>(kernel);
    std::lock_guard<std::mutex> lock(host_cache_mutex());
    Seen *slot = nullptr;
    for (int i = 0; i < nseen; i++)
        if (seen[i].fn == fn) slot = &seen[i];
    if (slot && dev < 64 && ((slot->devices >> dev) & 1ull)) return cudaSuccess;
    e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
    if (e != cudaSuccess) return e;
    if (!slot && nseen < 64) { slot = &seen[nseen++]; slot->fn = fn; slot->devices = 0; }
    if (slot && dev < 64) slot->devices |= 1ull << dev;
    return cudaSuccess;
}

}   // namespace b200
