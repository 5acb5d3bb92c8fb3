"""Overcooked2 ("Simplecooked", the JS-compatible rules) layout parsing: layout dict -> simulator constructor kwargs.

Host-side mirror of ``get_base_layout_params`` of the reference's envs/overcooked2_env.py:182-291.  It differs from
the Overcooked one (overcooked_layouts.py) in three ways: the terrain enum order of src/overcooked2_env/sim.hpp:40
(AIR, POT, COUNTER, ONION_SOURCE, DISH_SOURCE, SERVING, TOMATO_SOURCE), a dish-pickup shaping reward of 3, and
recipe values that are neither zeroed for unlisted orders nor multiplied by the order bonus (that code is commented
out in the reference, :275-280).

Layout files live in ``madrona_rl_envs_playground_b200/layouts2/*.layout`` (the dict format of oldercooked_ai's
``data/layouts``: ``cook_time`` / ``delivery_reward`` / ``start_order_list`` / ``num_items_for_soup``).
"""
import os

from .overcooked_layouts import (MAX_NUM_INGREDIENTS, PLAYER_NUMS, _recipe_index, get_true_orders,
                                 load_dict_from_file)

BASE_REW_SHAPING_PARAMS = {
    "PLACEMENT_IN_POT_REW": 3,
    "DISH_PICKUP_REWARD": 3,
    "SOUP_PICKUP_REWARD": 5,
}

# index in this list == TerrainT value (src/overcooked2_env/sim.hpp:40, envs/overcooked2_env.py:146)
TERRAIN_TYPES = [" ", "P", "X", "O", "D", "S", "T"]

LAYOUTS_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "layouts2")


def read_layout_dict(layout_name):
    return load_dict_from_file(os.path.join(LAYOUTS_DIR, layout_name + ".layout"))


def get_base_layout_params(layout_name, horizon, max_num_players=None):
    """Return the kwargs dict of ``SimplecookedSimulator`` for a layout."""
    if isinstance(layout_name, dict):
        params = dict(layout_name)
    elif layout_name.endswith(".layout"):
        params = load_dict_from_file(layout_name)
    else:
        params = read_layout_dict(layout_name)

    grid = [row.strip() for row in params.pop("grid").split("\n")]
    params.pop("start_order_list", None)
    params.pop("num_items_for_soup", None)
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
    params.pop("start_bonus_orders", None)
    params.pop("order_bonus", None)
    get_true_orders(all_orders_raw)   # validates the order dicts like the reference does; the result is unused there too

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
    params["recipe_values"] = values

    params["horizon"] = horizon
    return params
