"""Generate the golden fixtures under tests/golden/ from the REFERENCE itself.  Run in the build container only
(`python tests/golden/generate.py`); needs /root/reference and the oracle/_ref build (make -C oracle ref).

Sources of truth:
  * <env>_*.npz        — the reference's own Madrona CPU exec mode (oracle/_ref, ONE worker so episode numbering is
                         deterministic), driven with seeded random actions; stores the action stream and every
                         exported tensor after every step in the canonical layouts of oracle/oracle.h.
  * layout_params.json — the reference's get_base_layout_params (envs/overcooked_env.py:249-371), executed from its
                         source text with the two overcooked_ai loader helpers it imports, for every layout shipped in
                         madrona_rl_envs_playground_b200/layouts/.
"""
import ast
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
REFERENCE = "/root/reference"

from oracle import bindings as ob   # noqa: E402
from madrona_rl_envs_playground_b200.overcooked_layouts import LAYOUTS_DIR, get_base_layout_params   # noqa: E402


def reference_layout_params():
    """exec the reference's layout-parsing functions (pure Python) without importing its gym-dependent module"""
    src = open(os.path.join(REFERENCE, "envs/overcooked_env.py")).read()
    tree = ast.parse(src)
    keep = []
    for node in tree.body:
        if isinstance(node, ast.FunctionDef) and node.name in ("get_true_orders", "get_base_layout_params"):
            keep.append(node)
        elif isinstance(node, ast.Assign) and all(isinstance(t, ast.Name) for t in node.targets):
            names = {t.id for t in node.targets}
            if names & {"BASE_REW_SHAPING_PARAMS", "TOMATO", "ONION", "AIR", "POT", "COUNTER", "ONION_SOURCE",
                        "TOMATO_SOURCE", "DISH_SOURCE", "SERVING", "TERRAIN_TYPES", "PLAYER_NUMS",
                        "MAX_NUM_INGREDIENTS"}:
                keep.append(node)
    ns = {"MAX_NUM_INGREDIENTS": 3}
    # loader helpers: overcooked_ai_py/utils.py:31-33 and :223-226, pointed at our layout directory
    ns["load_dict_from_file"] = lambda path: eval(open(path).read())
    ns["read_layout_dict"] = lambda name: ns["load_dict_from_file"](os.path.join(LAYOUTS_DIR, name + ".layout"))
    exec(compile(ast.Module(body=keep, type_ignores=[]), "reference_overcooked_env", "exec"), ns)
    return ns["get_base_layout_params"]


def gen_layout_params():
    ref_fn = reference_layout_params()
    out = {}
    for f in sorted(os.listdir(LAYOUTS_DIR)):
        name = f[:-len(".layout")]
        for npl in (None, 2, 7):
            key = f"{name}|{npl}"
            out[key] = ref_fn(name, 400, npl)
            assert out[key] == get_base_layout_params(name, 400, npl), key
    json.dump(out, open(os.path.join(HERE, "layout_params.json"), "w"), indent=0, sort_keys=True)
    print("layout_params.json", len(out))


def gen_overcooked():
    cases = [("cramped_room", None, 8, 30, 75), ("coordination_ring", None, 4, 25, 60),
             ("asymmetric_advantages_tomato", None, 4, 40, 90), ("multiplayer_schelling", None, 4, 20, 50),
             ("many_player_layout", 9, 2, 15, 40)]
    for layout, npl, N, horizon, T in cases:
        params = get_base_layout_params(layout, horizon, npl)
        P = params["num_players"]
        ref = ob.RefOvercooked(params, N, num_workers=1)
        rng = np.random.default_rng(1234)
        # bias towards INTERACT so that pots fill, cook and get served inside the short horizon
        actions = rng.choice(6, size=(T, P, N), p=[0.14, 0.14, 0.14, 0.14, 0.04, 0.40]).astype(np.int8)
        obs0 = ref.obs
        obs, rew, done = [], [], []
        for t in range(T):
            o, r, d = ref.step(actions[t])
            obs.append(o); rew.append(r); done.append(d)
        np.savez_compressed(os.path.join(HERE, f"overcooked_{layout}.npz"), layout=layout,
                            num_players=-1 if npl is None else npl, horizon=horizon, actions=actions, obs0=obs0,
                            obs=np.stack(obs), rewards=np.stack(rew).astype(np.int16), done=np.stack(done).astype(np.int8))
        print(layout, "reward sum", int(np.stack(rew).sum()), "resets", int(np.stack(done).sum()))
        ref.close()


def legal_random_actions(rng, mask):
    logits = rng.random(mask.shape)
    logits[mask == 0] = -1.0
    return logits.argmax(-1)


def gen_hanabi():
    N, T = 16, 260
    ref = ob.RefHanabi(N, num_workers=1)
    ref.read()
    rng = np.random.default_rng(99)
    first = {k: getattr(ref, k).copy() for k in ("obs", "state", "mask", "active")}
    rec = {k: [] for k in ("actions", "obs", "state", "mask", "active", "rewards", "done")}
    for t in range(T):
        a = legal_random_actions(rng, ref.mask)
        ref.step(a)
        rec["actions"].append(a.astype(np.int8))
        for k in ("obs", "state", "mask", "active", "rewards", "done"):
            rec[k].append(getattr(ref, k).copy())
    np.savez_compressed(os.path.join(HERE, "hanabi_full.npz"),
                        obs0=first["obs"], state0=first["state"], mask0=first["mask"].astype(np.int8),
                        active0=first["active"].astype(np.int8), actions=np.stack(rec["actions"]),
                        obs=np.stack(rec["obs"]), state=np.stack(rec["state"]),
                        mask=np.stack(rec["mask"]).astype(np.int8), active=np.stack(rec["active"]).astype(np.int8),
                        rewards=np.stack(rec["rewards"]), done=np.stack(rec["done"]).astype(np.int8))
    print("hanabi resets", int(np.stack(rec["done"]).sum()))
    ref.close()


def gen_balance():
    N, T = 64, 60
    ref = ob.RefBalance(N, num_workers=1)
    ref.read()
    rng = np.random.default_rng(5)
    obs0 = ref.obs.copy()
    actions = rng.integers(0, 4, (T, 2, N)).astype(np.int8)
    obs, rew, done = [], [], []
    for t in range(T):
        ref.step(actions[t])
        obs.append(ref.obs.copy()); rew.append(ref.rewards.copy()); done.append(ref.done.copy())
    np.savez_compressed(os.path.join(HERE, "balance.npz"), obs0=obs0, actions=actions, obs=np.stack(obs).astype(np.int8),
                        rewards=np.stack(rew), done=np.stack(done).astype(np.int8))
    print("balance resets", int(np.stack(done).sum()))
    ref.close()


def gen_cartpole():
    N, T = 64, 200
    ref = ob.RefCartpole(N, num_workers=1)
    ref.read()
    rng = np.random.default_rng(6)
    state0 = ref.state.copy()
    actions = rng.integers(0, 2, (T, N)).astype(np.int8)
    state, done = [], []
    for t in range(T):
        ref.step(actions[t])
        state.append(ref.state.copy()); done.append(ref.done.copy())
    np.savez_compressed(os.path.join(HERE, "cartpole.npz"), state0=state0, actions=actions, state=np.stack(state),
                        done=np.stack(done).astype(np.int8))
    print("cartpole resets", int(np.stack(done).sum()))
    ref.close()


def gen_acrobat():
    """src/acrobat_env through the reference build.  Before every step a few worlds are put near the upright position
    (the exported state column is re-imported at the start of a CPU step, core/state.cpp:484-509) so that terminations
    happen next to the truncations of the Manager-wide episode-length counter; 300 worlds = two 256-world tiles."""
    N, T = 300, 16
    ref = ob.RefAcrobat(N, num_workers=1)
    ref.read()
    rng = np.random.default_rng(8)
    state0 = ref.state.copy()
    actions = rng.integers(0, 3, (T, N)).astype(np.int8)
    state_in, state, done = [], [], []
    for t in range(T):
        s = ref.state.copy()
        up = rng.random(N) < 0.02
        s[up] = np.stack([rng.uniform(2.6, 3.1, up.sum()), rng.uniform(-0.3, 0.3, up.sum()),
                          rng.uniform(-1, 1, up.sum()), rng.uniform(-1, 1, up.sum())], axis=1).astype(np.float32)
        ref.write_state(s)
        state_in.append(s)
        ref.step(actions[t])
        state.append(ref.state.copy()); done.append(ref.done.copy())
    np.savez_compressed(os.path.join(HERE, "acrobat.npz"), state0=state0, actions=actions, state_in=np.stack(state_in),
                        state=np.stack(state), done=np.stack(done).astype(np.int8))
    print("acrobat resets", int(np.stack(done).sum()))
    ref.close()


def gen_rng():
    """known-answer vectors of the per-world RNG straight from the reference header (src/cartpole_env/rng.hpp) through
    the reference Cartpole reset: state = -0.05 + rand()*0.1 for episode indices 0..N-1 (sim.cpp:48-66)."""
    ref = ob.RefCartpole(32, num_workers=1)
    ref.read()
    np.save(os.path.join(HERE, "rng_cartpole_reset_states.npy"), ref.state.copy())
    ref.close()


if __name__ == "__main__":
    ob.build(ref=True)
    gen_layout_params()
    gen_overcooked()
    gen_hanabi()
    gen_balance()
    gen_cartpole()
    gen_acrobat()
    gen_rng()
    os.system(f"ls -la {HERE}")
