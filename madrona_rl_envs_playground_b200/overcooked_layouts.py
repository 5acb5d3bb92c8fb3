"""Overcooked layout parsing: layout dict -> simulator constructor kwargs.

Host-side mirror of the reference's ``get_base_layout_params`` (envs/overcooked_env.py:261-371) and of the two
overcooked_ai loader helpers it relies on (overcooked_ai_py/utils.py:31-33 ``load_dict_from_file`` and
:223-226 ``read_layout_dict``).  Same keys, same defaults, same recipe-table construction
(index ``4 * onions + tomatoes``), so that ``OvercookedSimulator(**params)`` receives exactly what the
reference's nanobind constructor would (src/overcooked_env/bindings.cpp:15-53).

Layout files live in ``madrona_rl_envs_playground_b200/layouts/*.layout`` (same dict format as overcooked_ai's
``data/layouts``); a path ending in ``.layout`` is loaded from disk instead, like the reference does.
"""
import ast
import os

MAX_NUM_INGREDIENTS = 3

BASE_REW_SHAPING_PARAMS = {
    "PLACEMENT_IN_POT_REW": 3,
    "DISH_PICKUP_REWARD": 0,
    "SOUP_PICKUP_REWARD": 5,
}

TOMATO = "tomato"
ONION = "onion"

# index in this list == TerrainT value (src/overcooked_env/sim.hpp:41)
TERRAIN_TYPES = [" ", "P", "X", "O", "T", "D", "S"]
# player start symbols, player i is PLAYER_NUMS[i] (envs/overcooked_env.py:243-246)
PLAYER_NUMS = list("1234567890!@#$%^&*()abcdefghijklmnopqrst")

LAYOUTS_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "layouts")


def load_dict_from_file(filepath):
    """Layout files are Python dict literals (triple-quoted grid, ``None``); parsed without ``eval``."""
    with open(filepath, "r") as f:
        return ast.literal_eval(f.read())


def read_layout_dict(layout_name):
    return load_dict_from_file(os.path.join(LAYOUTS_DIR, layout_name + ".layout"))


def _recipe_index(order):
    onions = sum(1 for ing in order["ingredients"] if ing == ONION)
    tomatoes = sum(1 for ing in order["ingredients"] if ing == TOMATO)
    return (MAX_NUM_INGREDIENTS + 1) * onions + tomatoes


def get_true_orders(orders):
    ans = [0] * ((MAX_NUM_INGREDIENTS + 1) ** 2)
    for order in orders:
        ans[_recipe_index(order)] = 1
    return ans


def get_base_layout_params(layout_name, horizon, max_num_players=None):
    """Return the kwargs dict of ``OvercookedSimulator`` for a layout (reference: envs/overcooked_env.py:261-371)."""
    if isinstance(layout_name, dict):
        params = dict(layout_name)
    elif layout_name.endswith(".layout"):
        params = load_dict_from_file(layout_name)
    else:
        params = read_layout_dict(layout_name)

    grid = [row.strip() for row in params.pop("grid").split("\n")]
    layout_grid = [list(row) for row in grid]

    player_positions = [None] * 64
    for y, row in enumerate(layout_grid):
        for x, c in enumerate(row):
            if c in PLAYER_NUMS:
                layout_grid[y][x] = " "
                idx = PLAYER_NUMS.index(c)
                if max_num_players is None or idx < max_num_players:
                    player_positions[idx] = (x, y)
    num_players = len([p for p in player_positions if p is not None])
    player_positions = player_positions[:num_players]

    params["height"] = len(layout_grid)
    params["width"] = len(layout_grid[0])
    params["terrain"] = [TERRAIN_TYPES.index(c) for row in layout_grid for c in row]
    params["num_players"] = len(player_positions)
    params["start_player_x"] = [p[0] for p in player_positions]
    params["start_player_y"] = [p[1] for p in player_positions]

    shaping = params.pop("rew_shaping_params", None) or BASE_REW_SHAPING_PARAMS
    params["placement_in_pot_rew"] = shaping["PLACEMENT_IN_POT_REW"]
    params["dish_pickup_rew"] = shaping["DISH_PICKUP_REWARD"]
    params["soup_pickup_rew"] = shaping["SOUP_PICKUP_REWARD"]

    all_orders_raw = params.pop("start_all_orders", None) or []
    bonus_orders_raw = params.pop("start_bonus_orders", None) or []
    all_orders = get_true_orders(all_orders_raw)
    bonus_orders = get_true_orders(bonus_orders_raw)
    order_bonus = params.pop("order_bonus", 2)

    n_recipes = (MAX_NUM_INGREDIENTS + 1) ** 2
    rng4 = range(MAX_NUM_INGREDIENTS + 1)

    times = [20] * n_recipes
    if "onion_time" in params and "tomato_time" in params:
        ot, tt = params.pop("onion_time"), params.pop("tomato_time")
        times = [o * ot + t * tt for o in rng4 for t in rng4]
    if "recipe_times" in params:
        for order, time in zip(all_orders_raw, params["recipe_times"]):
            times[_recipe_index(order)] = time
    if "cook_time" in params:
        times = [params.pop("cook_time")] * n_recipes
    params["recipe_times"] = times

    values = [20] * n_recipes
    if "onion_value" in params and "tomato_value" in params:
        ov, tv = params.pop("onion_value"), params.pop("tomato_value")
        values = [o * ov + t * tv for o in rng4 for t in rng4]
    if "recipe_values" in params:
        for order, value in zip(all_orders_raw, params["recipe_values"]):
            values[_recipe_index(order)] = value
    if "delivery_reward" in params:
        values = [params.pop("delivery_reward")] * n_recipes
    for i in range(n_recipes):
        if bonus_orders[i]:
            values[i] *= order_bonus
        if not all_orders[i]:
            values[i] = 0
    params["recipe_values"] = values

    params["horizon"] = horizon
    return params
