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
};

RefSim *make(uint32_t num_worlds, uint32_t num_workers, const EnvNS::Config &cfg,
             const EnvNS::WorldInit &proto)
{
    auto *episodes = new EnvNS::EpisodeManager { 0 };
    HeapArray<EnvNS::WorldInit> inits(num_worlds);
    for (uint32_t i = 0; i < num_worlds; i++) {
        inits[i] = proto;
        inits[i].episodeMgr = episodes;
    }
    auto *exec = new Exec(ThreadPoolExecutor::Config {
        .numWorlds = num_worlds,
        .numExportedBuffers = (uint32_t)EnvNS::ExportID::NumExports,
        .numWorkers = num_workers,
    }, cfg, inits.data());
    return new RefSim { episodes, std::move(inits), exec, num_worlds };
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

void ref_step(void *h) { ((RefSim *)h)->exec->run(); }

// Base pointer of exported column `slot` (ExportID order of the env's sim.hpp).  CPU memory is
// world-major: all rows of world 0, then world 1, ... (external/madrona/src/core/state.cpp:511-554).
void *ref_exported(void *h, int slot) { return ((RefSim *)h)->exec->getExported(slot); }

int ref_num_exports(void) { return (int)EnvNS::ExportID::NumExports; }

void ref_destroy(void *h)
{
    RefSim *s = (RefSim *)h;
    delete s->exec;
    delete s->episodes;
    delete s;
}

}
