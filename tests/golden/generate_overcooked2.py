#!/usr/bin/env python
"""Golden vectors for Overcooked2 ("Simplecooked", src/overcooked2_env) from the REFERENCE ITSELF (oracle/_ref,
Madrona CPU exec mode, one worker).  Needs /root/reference; run once in the build container, commit the outputs:

  * layout2_params.json     — the reference's get_base_layout_params of envs/overcooked2_env.py:182-291, executed from
                              its source text (loader helpers pointed at our layouts2/ directory), for every layout
  * overcooked2_<layout>.npz — actions, initial obs, per-step obs / rewards / done

    python tests/golden/generate_overcooked2.py
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
from madrona_rl_envs_playground_b200.overcooked2_layouts import LAYOUTS_DIR, get_base_layout_params   # noqa: E402


def reference_layout_params():
    src = open(os.path.join(REFERENCE, "envs/overcooked2_env.py")).read()
    keep = []
    for node in ast.parse(src).body:
        if isinstance(node, ast.FunctionDef) and node.name in ("get_true_orders", "get_base_layout_params"):
            keep.append(node)
        elif isinstance(node, ast.Assign) and all(isinstance(t, ast.Name) for t in node.targets):
            if {t.id for t in node.targets} & {"BASE_REW_SHAPING_PARAMS", "TOMATO", "ONION", "AIR", "POT", "COUNTER",
                                                "ONION_SOURCE", "TOMATO_SOURCE", "DISH_SOURCE", "SERVING",
                                                "TERRAIN_TYPES", "PLAYER_NUMS", "MAX_NUM_INGREDIENTS"}:
                keep.append(node)
    ns = {}
    ns["load_dict_from_file"] = lambda path: eval(open(path).read())
    ns["read_layout_dict"] = lambda name: ns["load_dict_from_file"](os.path.join(LAYOUTS_DIR, name + ".layout"))
    exec(compile(ast.Module(body=keep, type_ignores=[]), "reference_overcooked2_env", "exec"), ns)
    return ns["get_base_layout_params"]


def gen_layout_params():
    ref_fn = reference_layout_params()
    out = {}
    for f in sorted(os.listdir(LAYOUTS_DIR)):
        name = f[:-len(".layout")]
        for npl in (None, 1):
            key = f"{name}|{npl}"
            out[key] = ref_fn(name, 200, npl)
            assert out[key] == get_base_layout_params(name, 200, npl), key
    json.dump(out, open(os.path.join(HERE, "layout2_params.json"), "w"), indent=0, sort_keys=True)
    print("layout2_params.json", len(out))


def gen_overcooked2():
    cases = [("simple", 8, 60, 150), ("unident_s", 4, 50, 120), ("random0", 4, 45, 100), ("random1", 4, 200, 100),
             ("random3", 4, 45, 100), ("simple_tomato", 4, 70, 150)]
    for layout, N, horizon, T in cases:
        params = get_base_layout_params(layout, horizon)
        P = params["num_players"]
        ref = ob.RefOvercooked2(params, N, num_workers=1)
        rng = np.random.default_rng(4321)
        actions = rng.choice(6, size=(T, P, N), p=[0.14, 0.14, 0.14, 0.14, 0.04, 0.40]).astype(np.int8)
        obs0 = ref.obs
        obs, rew, done = [], [], []
        for t in range(T):
            o, r, d = ref.step(actions[t])
            obs.append(o); rew.append(r); done.append(d)
        np.savez_compressed(os.path.join(HERE, f"overcooked2_{layout}.npz"), layout=layout, horizon=horizon,
                            actions=actions, obs0=obs0, obs=np.stack(obs), rewards=np.stack(rew).astype(np.int16),
                            done=np.stack(done).astype(np.int8))
        print(layout, "reward sum", int(np.stack(rew).sum()), "resets", int(np.stack(done).sum()))
        ref.close()


if __name__ == "__main__":
    ob.build(ref=True)
    gen_layout_params()
    gen_overcooked2()
