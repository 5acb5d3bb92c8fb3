// TEST INFRASTRUCTURE — not part of the product path.
//
// Thin C harness around the REFERENCE's own Madrona CPU exec mode.  It is compiled
// (by oracle/Makefile, target `ref`) together with the unmodified reference sources
// where they lie under /root/reference:
//     src/<env>/sim.cpp, external/madrona/src/{common,core}/*.cpp, external/madrona/src/mw/cpu_exec.cpp
// and produces oracle/_ref/libref_<env>.so.  No reference source is copied into this repo.
//
// It stands in for src/<env>/mgr.cpp (CPUImpl, e.g. src/overcooked_env/mgr.cpp:44-77,138-188)
// without nanobind / py::Tensor: the executor is driven directly through
// madrona::TaskGraphExecutor (external/madrona/include/madrona/mw_cpu.hpp:72-105).
//
// One translation unit per environment, selected with -DREF_ENV_<NAME>.
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <new>
#include <vector>
#include <algorithm>
#include <dirent.h>
#include <fcntl.h>
#include <sched.h>
#include <unistd.h>
#include <cstdio>

#if defined(REF_ENV_OVERCOOKED)
#include "sim.hpp"
namespace EnvNS = Overcooked;
#elif defined(REF_ENV_OVERCOOKED2)
#include "sim.hpp"
namespace EnvNS = Simplecooked;   // src/overcooked2_env: the JS-compatible rules (2 players, 5P+10 channels)
#elif defined(REF_ENV_HANABI)
#include "sim.hpp"
namespace EnvNS = Hanabi;
#elif defined(REF_ENV_BALANCE)
#include "sim.hpp"
namespace EnvNS = Balance;
#elif defined(REF_ENV_CARTPOLE)
#include "sim.hpp"
namespace EnvNS = Cartpole;
#elif defined(REF_ENV_ACROBAT)
#include "sim.hpp"
namespace EnvNS = Acrobat;
#else
#error "define REF_ENV_<NAME>"
#endif

#include <madrona/mw_cpu.hpp>

using namespace madrona;

namespace {

using Exec = TaskGraphExecutor<EnvNS::Engine, EnvNS::Sim, EnvNS::Config, EnvNS::WorldInit>;

struct RefSim {
    EnvNS::EpisodeManager *episodes;
    HeapArray<EnvNS::WorldInit> inits;
    Exec *exec;
    uint32_t num_worlds;
    std::vector<int> worker_stat_fds;   // /proc/self/task/<tid>/stat of the executor's worker threads
};

// The reference's thread pool has a lost-wakeup window (external/madrona/src/mw/cpu_exec.cpp:205-236): the worker
// that draws job index == numJobs stores workerWakeup = 0 AFTER the main thread may already have been released, so
// a run() started right away can have its wake-up flag overwritten and both sides wait forever.  With the
// reference's own Manager the window is hidden behind Python time between steps; a tight test loop on a busy host
// hits it.  The harness therefore lets every worker reach its futex wait before ref_step returns (thread state 'S'
// in /proc; the workers make no other blocking call).  This changes nothing the reference computes.
std::vector<int> list_tids()
{
    std::vector<int> tids;
    if (DIR *d = opendir("/proc/self/task")) {
        while (dirent *e = readdir(d)) {
            if (e->d_name[0] >= '0' && e->d_name[0] <= '9') tids.push_back(atoi(e->d_name));
        }
        closedir(d);
    }
    std::sort(tids.begin(), tids.end());
    return tids;
}

void quiesce(RefSim *s)
{
    for (int fd : s->worker_stat_fds) {
        for (int spin = 0; spin < 2000000; spin++) {
            char buf[256];
            ssize_t n = pread(fd, buf, sizeof(buf) - 1, 0);
            if (n <= 0) break;                       // thread gone / procfs unavailable: nothing to wait for
            buf[n] = 0;
            const char *p = strrchr(buf, ')');       // "<tid> (<comm>) <state> ..."
            if (!p || !p[1] || p[2] == 'S') break;
            sched_yield();
        }
    }
}

RefSim *make(uint32_t num_worlds, uint32_t num_workers, const EnvNS::Config &cfg,
             const EnvNS::WorldInit &proto)
{
    auto *episodes = new EnvNS::EpisodeManager { 0 };
    HeapArray<EnvNS::WorldInit> inits(num_worlds);
    for (uint32_t i = 0; i < num_worlds; i++) {
        inits[i] = proto;
        inits[i].episodeMgr = episodes;
    }
    std::vector<int> before = list_tids();
    auto *exec = new Exec(ThreadPoolExecutor::Config {
        .numWorlds = num_worlds,
        .numExportedBuffers = (uint32_t)EnvNS::ExportID::NumExports,
        .numWorkers = num_workers,
    }, cfg, inits.data());
    std::vector<int> fds;
    for (int tid : list_tids()) {
        if (!std::binary_search(before.begin(), before.end(), tid)) {
            char path[64];
            snprintf(path, sizeof(path), "/proc/self/task/%d/stat", tid);
            int fd = open(path, O_RDONLY | O_CLOEXEC);
            if (fd >= 0) fds.push_back(fd);
        }
    }
    auto *s = new RefSim { episodes, std::move(inits), exec, num_worlds, std::move(fds) };
    quiesce(s);
    return s;
}

}

extern "C" {

#if defined(REF_ENV_OVERCOOKED) || defined(REF_ENV_OVERCOOKED2)
// Mirrors Manager::Impl::init, src/overcooked_env/mgr.cpp:138-167 (src/overcooked2_env/mgr.cpp has the same Config).
void *ref_create(uint32_t num_worlds, uint32_t num_workers,
                 const int64_t *terrain, int64_t height, int64_t width, int64_t num_players,
                 const int64_t *start_x, const int64_t *start_y,
                 int64_t placement_in_pot_rew, int64_t dish_pickup_rew, int64_t soup_pickup_rew,
                 const int64_t *recipe_values, const int64_t *recipe_times, int64_t horizon)
{
    // The reference Config is ~10 KB of POD; value-initialise then fill like mgr.cpp does.
    auto *cfg = new EnvNS::Config {};
    cfg->height = height;
    cfg->width = width;
    cfg->num_players = num_players;
    cfg->placement_in_pot_rew = placement_in_pot_rew;
    cfg->dish_pickup_rew = dish_pickup_rew;
    cfg->soup_pickup_rew = soup_pickup_rew;
    cfg->horizon = horizon;
    for (int r = 0; r < NUM_RECIPES; r++) {
        cfg->recipe_values[r] = (uint8_t)recipe_values[r];
        cfg->recipe_times[r] = (uint8_t)recipe_times[r];
    }
    for (int p = 0; p < num_players; p++) {
        cfg->start_player_x[p] = (uint8_t)start_x[p];
        cfg->start_player_y[p] = (uint8_t)start_y[p];
    }
    for (int x = 0; x < height * width; x++) {
        cfg->terrain[x] = (EnvNS::TerrainT)terrain[x];
    }
    EnvNS::WorldInit proto { nullptr };
    RefSim *s = make(num_worlds, num_workers, *cfg, proto);
    delete cfg;
    return s;
}
#elif defined(REF_ENV_HANABI)
// Mirrors setupWorldInitData + init, src/hanabi_env/mgr.cpp:124-163.
void *ref_create(uint32_t num_worlds, uint32_t num_workers,
                 uint32_t colors, uint32_t ranks, uint32_t players,
                 uint32_t max_information_tokens, uint32_t max_life_tokens)
{
    EnvNS::Config cfg { players };
    EnvNS::WorldInit proto { nullptr, colors, ranks, players, max_information_tokens,
                             max_life_tokens };
    return make(num_worlds, num_workers, cfg, proto);
}
#else
void *ref_create(uint32_t num_worlds, uint32_t num_workers)
{
    EnvNS::Config cfg {};
    EnvNS::WorldInit proto { nullptr };
    return make(num_worlds, num_workers, cfg, proto);
}
#endif

void ref_step(void *h)
{
    RefSim *s = (RefSim *)h;
    s->exec->run();
    quiesce(s);
}

// Base pointer of exported column `slot` (ExportID order of the env's sim.hpp).  CPU memory is
// world-major: all rows of world 0, then world 1, ... (external/madrona/src/core/state.cpp:511-554).
void *ref_exported(void *h, int slot) { return ((RefSim *)h)->exec->getExported(slot); }

int ref_num_exports(void) { return (int)EnvNS::ExportID::NumExports; }

void ref_destroy(void *h)
{
    RefSim *s = (RefSim *)h;
    for (int fd : s->worker_stat_fds) close(fd);
    delete s->exec;
    delete s->episodes;
    delete s;
}

}
