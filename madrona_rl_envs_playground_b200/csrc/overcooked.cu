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
// Two kernels:
//   step_tile<PMAX>  one warp = one tile of 32 consecutive worlds, lane = world.  Per-player state lives in registers,
//                    object slots in shared memory [K][32] (conflict free).  The lane writes its world's observation
//                    slab for observer 0 into a shared staging tile laid out like the output, the warp hands the whole
//                    tile (32*S*C contiguous bytes) to the copy engine (cp.async.bulk), then only the 2P observer-
//                    relative bytes are re-patched for the next observer.  Used when P <= 8 and 32*S*C fits.
//   step_world_cta   one CTA = one world, for many-agent / large-grid layouts (P up to 64, S up to 255).  The
//                    observer-independent slab is built once per world in two shared buffers; per observer only the
//                    player bytes are patched before the slab (S*C contiguous bytes) is bulk-stored.
#include "overcooked.cuh"
#include "overcooked_host.h"

namespace b200 {
namespace oc {

struct SmemObjs {   // object slots of one world inside a [K][32] shared tile (lane = world)
    uint32_t *base;
    __device__ __forceinline__ uint32_t &operator[](int k) { return base[k * 32]; }
};
struct SmemObjsDense {   // object slots of one world, contiguous
    uint32_t *base;
    __device__ __forceinline__ uint32_t &operator[](int k) { return base[k]; }
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
};

// observer-independent part of one world's slab: terrain one-hot, urgency, cell objects, held objects
template <int PMAX>
__device__ __forceinline__ void encode_static(const Layout &L, uint8_t *slab, int P, const uint8_t (&pos)[PMAX],
                                              const uint32_t (&held)[PMAX], SmemObjs objs, int timestep)
{
    const int shift = 5 * L.P;
    const bool urgent = (L.horizon - timestep) < 40;   // sim.cpp:79-83, post-increment / post-reset timestep
    for (int c = 0; c < L.S; c++) {
        uint8_t *x = slab + (int)L.xmajor[c] * L.C + shift;
        const int t = L.terrain[c];
        if (t) x[t - 1] = 1;                            // sim.cpp:642-645
        if (urgent) x[15] = 1;
    }
    for (int k = 0; k < L.K; k++) encode_cell_object(L, slab, L.cell_of_slot[k], objs[k]);
#pragma unroll
    for (int q = 0; q < PMAX; q++)
        if (q < P) encode_held_object(L, slab, pos[q], held[q]);
}

// cooperative copy of the layout into shared memory (sizeof(Layout) is a multiple of 16)
__device__ __forceinline__ void stage_layout(Layout *dst, const Layout *__restrict__ src, int tid, int nthreads)
{
    const uint4 *s = reinterpret_cast<const uint4 *>(src);
    uint4 *d = reinterpret_cast<uint4 *>(dst);
    for (int i = tid; i < (int)(sizeof(Layout) / 16); i += nthreads) d[i] = __ldg(s + i);
}

template <int PMAX>
__global__ void __launch_bounds__(32) step_tile(const Layout *__restrict__ gL, const Buffers B, const int init_only)
{
    extern __shared__ __align__(16) uint8_t smem[];
    __shared__ Layout sL;
    const int lane = threadIdx.x;
    stage_layout(&sL, gL, lane, 32);
    __syncwarp();
    const Layout &L = sL;
    const int P = L.P;
    const uint32_t SC = (uint32_t)L.S * L.C;
    const uint32_t tile0 = blockIdx.x * 32u;
    const uint32_t n = tile0 + lane;
    const bool live = n < B.N;
    const uint32_t tile_worlds = min(32u, B.N - tile0);

    uint8_t *stage = smem;                                            // [32][SC], same layout as the output tile
    uint32_t *sobj = reinterpret_cast<uint32_t *>(smem + ((32u * SC + 15u) & ~15u));   // [K][32]
    SmemObjs objs{sobj + lane};

    // issue the state loads first, zero the staging tile while they are in flight
    int act[PMAX];
    uint8_t pos[PMAX], ori[PMAX];
    uint32_t held[PMAX];
    int timestep = 0;
    if (live && !init_only) {
#pragma unroll
        for (int q = 0; q < PMAX; q++) {
            if (q < P) {
                const uint2 pl = B.players[(size_t)q * B.N + n];
                held[q] = pl.x;
                pos[q] = (uint8_t)(pl.y & 0xFF);
                ori[q] = (uint8_t)((pl.y >> 8) & 0xFF);
                act[q] = B.actions[(size_t)q * B.N + n];
            }
        }
        // object slots go global -> shared without passing through registers (LDGSTS), so the lane is not stalled
        // on them while it zeroes the staging tile
        for (int k = 0; k < L.K; k++)
            asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(smem_u32(&objs[k])),
                         "l"(B.objs + (size_t)k * B.N + n)
                         : "memory");
        asm volatile("cp.async.commit_group;" ::: "memory");
        timestep = B.timestep[n];
    }
    {
        uint4 *s4 = reinterpret_cast<uint4 *>(stage);
        const uint32_t chunks = (32u * SC + 15u) / 16u;
        for (uint32_t i = lane; i < chunks; i += 32) s4[i] = make_uint4(0, 0, 0, 0);
    }

    asm volatile("cp.async.wait_group 0;" ::: "memory");
    int reward = 0, done = 0;
    if (live) {
        if (init_only) {
            // Sim::Sim (sim.cpp:556-659): reset state, then the initial observation
            for (int k = 0; k < L.K; k++) objs[k] = OBJ_NONE;
#pragma unroll
            for (int q = 0; q < PMAX; q++) {
                pos[q] = (q < P) ? L.start_pos[q] : 0;
                ori[q] = NORTH;
                held[q] = OBJ_NONE;
            }
        } else {
            reward = step_world<PMAX>(L, P, act, pos, ori, held, objs, timestep, done);
        }
        // write back state + step outputs
#pragma unroll
        for (int q = 0; q < PMAX; q++) {
            if (q < P) {
                B.players[(size_t)q * B.N + n] = make_uint2(held[q], (uint32_t)pos[q] | ((uint32_t)ori[q] << 8));
                B.rewards[(size_t)q * B.N + n] = reward;   // team reward to every agent, sim.cpp:472
            }
        }
        for (int k = 0; k < L.K; k++) B.objs[(size_t)k * B.N + n] = objs[k];
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
    __syncwarp();   // staging tile zeroed by all lanes before anyone patches it

    uint8_t *slab = stage + (size_t)lane * SC;
    if (live) encode_static<PMAX>(L, slab, P, pos, held, objs, timestep);

    for (int p = 0; p < P; p++) {
        if (live) {
#pragma unroll
            for (int q = 0; q < PMAX; q++)
                if (q < P) encode_player(L, slab, p, q, pos[q], ori[q], 1);
        }
        fence_proxy_async_smem();
        __syncwarp();
        slab_store(B.obs + ((size_t)p * B.N + tile0) * SC, stage, tile_worlds * SC, lane, 32);
        if (p + 1 < P) {
            if (lane == 0) bulk_wait_read<0>();   // the copy engine has finished reading the staging tile
            __syncwarp();
            if (live) {
#pragma unroll
                for (int q = 0; q < PMAX; q++)
                    if (q < P) encode_player(L, slab, p, q, pos[q], ori[q], 0);
            }
        }
    }
    if (lane == 0) bulk_wait_read<0>();
}

// ---------------------------------------------------------------------------------------------------------------
// CTA-per-world kernel for many-agent / large-grid layouts.
// ---------------------------------------------------------------------------------------------------------------
constexpr int WORLD_CTA_THREADS = 256;

__global__ void __launch_bounds__(WORLD_CTA_THREADS) step_world_cta(const Layout *__restrict__ gL, const Buffers B,
                                                                   const int init_only)
{
    extern __shared__ __align__(16) uint8_t smem[];
    __shared__ Layout sL;
    __shared__ int s_act[64];
    __shared__ uint8_t s_pos[64], s_ori[64];
    __shared__ uint32_t s_held[64];
    __shared__ uint32_t s_obj[256];
    __shared__ int s_misc[4];   // timestep, reward, done

    const int tid = threadIdx.x;
    stage_layout(&sL, gL, tid, WORLD_CTA_THREADS);
    __syncthreads();
    const Layout &L = sL;
    const int P = L.P;
    const uint32_t SC = (uint32_t)L.S * L.C;
    const uint32_t n = blockIdx.x;

    // two slab buffers; buffer b serves observers p == b (mod 2) and is placed congruent (mod 16) to its first
    // destination so that the interior of every slab can go through the copy engine
    const uint32_t buf_stride = (SC + 31u) & ~15u;
    uint8_t *obs_w = B.obs + (size_t)n * SC;                 // observer 0 slab of this world
    const size_t obs_pstride = (size_t)B.N * SC;             // distance between observers
    uint8_t *buf[2];
    buf[0] = smem + (reinterpret_cast<uintptr_t>(obs_w) & 15u);
    buf[1] = smem + buf_stride + (reinterpret_cast<uintptr_t>(obs_w + obs_pstride) & 15u);

    if (!init_only) {
        if (tid < P) {
            const uint2 pl = B.players[(size_t)tid * B.N + n];
            s_held[tid] = pl.x;
            s_pos[tid] = (uint8_t)(pl.y & 0xFF);
            s_ori[tid] = (uint8_t)((pl.y >> 8) & 0xFF);
            s_act[tid] = B.actions[(size_t)tid * B.N + n];
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
        const uint32_t chunks = (2u * buf_stride) / 16u;
        for (uint32_t i = tid; i < chunks; i += WORLD_CTA_THREADS) s4[i] = make_uint4(0, 0, 0, 0);
    }
    __syncthreads();

    if (tid == 0 && !init_only) {
        SmemObjsDense objs{s_obj};
        int timestep = s_misc[0], done = 0;
        const int reward = step_world<64>(L, P, s_act, s_pos, s_ori, s_held, objs, timestep, done);
        s_misc[0] = timestep; s_misc[1] = reward; s_misc[2] = done;
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

    // observer-independent slab content, into both buffers
    const int nbuf = P > 1 ? 2 : 1;
    const bool urgent = (L.horizon - timestep) < 40;
    for (int b = 0; b < nbuf; b++) {
        uint8_t *slab = buf[b];
        for (int c = tid; c < L.S; c += WORLD_CTA_THREADS) {
            uint8_t *x = slab + (int)L.xmajor[c] * L.C + 5 * P;
            const int t = L.terrain[c];
            if (t) x[t - 1] = 1;
            if (urgent) x[15] = 1;
        }
        for (int k = tid; k < L.K; k += WORLD_CTA_THREADS) encode_cell_object(L, slab, L.cell_of_slot[k], s_obj[k]);
        if (tid < P) encode_held_object(L, slab, s_pos[tid], s_held[tid]);
    }
    __syncthreads();

    for (int p = 0; p < P; p++) {
        uint8_t *slab = buf[p & 1];
        if (p >= 2) {
            if (tid == 0) bulk_wait_read<1>();   // the copy of observer p-2 has released this buffer
            __syncthreads();
            if (tid < P) encode_player(L, slab, p - 2, tid, s_pos[tid], s_ori[tid], 0);
        }
        if (tid < P) encode_player(L, slab, p, tid, s_pos[tid], s_ori[tid], 1);
        fence_proxy_async_smem();
        __syncthreads();
        slab_store(obs_w + (size_t)p * obs_pstride, slab, SC, tid, WORLD_CTA_THREADS);
    }
    if (tid == 0) bulk_wait_read<0>();
}

// ---------------------------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------------------------
static size_t tile_smem_bytes(const Layout &L)
{
    const size_t SC = (size_t)L.S * L.C;
    return ((32 * SC + 15) & ~(size_t)15) + (size_t)L.K * 32 * 4;
}
static size_t world_cta_smem_bytes(const Layout &L)
{
    const size_t SC = (size_t)L.S * L.C;
    return 2 * ((SC + 31) & ~(size_t)15) + 16;
}

int pick_kernel(const Layout &L)
{
    // the warp-tile kernel needs P <= 8 and a 32-world staging tile small enough to keep >= 4 CTAs per SM
    if (L.P <= 8 && tile_smem_bytes(L) + sizeof(Layout) <= 56 * 1024) return L.P <= 2 ? 2 : (L.P <= 4 ? 4 : 8);
    return 0;
}

template <int PMAX>
static cudaError_t launch_tile(const Layout &hL, const Layout *dL, const Buffers &B, int init_only, cudaStream_t st)
{
    const size_t smem = tile_smem_bytes(hL);
    static bool attr_set = false;
    if (!attr_set) {
        cudaError_t e = cudaFuncSetAttribute(step_tile<PMAX>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
        if (e != cudaSuccess) return e;
        attr_set = true;
    }
    const unsigned grid = (B.N + 31) / 32;
    step_tile<PMAX><<<grid, 32, smem, st>>>(dL, B, init_only);
    return cudaGetLastError();
}

cudaError_t launch_step(const Layout &hL, const Layout *dL, const LaunchArgs &a, int init_only, cudaStream_t st)
{
    Buffers B;
    B.players = (uint2 *)a.players; B.objs = a.objs; B.timestep = a.timestep; B.actions = a.actions;
    B.rewards = a.rewards; B.done = a.done; B.obs = a.obs; B.ep_return = a.ep_return; B.stats = a.stats; B.N = a.N;
    switch (a.force_kernel >= 0 ? a.force_kernel : pick_kernel(hL)) {
    case 2: return launch_tile<2>(hL, dL, B, init_only, st);
    case 4: return launch_tile<4>(hL, dL, B, init_only, st);
    case 8: return launch_tile<8>(hL, dL, B, init_only, st);
    default: break;
    }
    const size_t smem = world_cta_smem_bytes(hL);
    static bool attr_set = false;
    if (!attr_set) {
        cudaError_t e = cudaFuncSetAttribute(step_world_cta, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
        if (e != cudaSuccess) return e;
        attr_set = true;
    }
    step_world_cta<<<a.N, WORLD_CTA_THREADS, smem, st>>>(dL, B, init_only);
    return cudaGetLastError();
}

}   // namespace oc
}   // namespace b200
