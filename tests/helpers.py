"""Shared test helpers: golden loading, action sampling, thin numpy views of the CUDA simulators."""
import os

import numpy as np

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")

OVERCOOKED_GOLDENS = ["cramped_room", "coordination_ring", "asymmetric_advantages_tomato", "multiplayer_schelling",
                      "many_player_layout"]


OVERCOOKED2_GOLDENS = ["simple", "unident_s", "random0", "random1", "random3", "simple_tomato"]


def load_golden(name):
    return np.load(os.path.join(GOLDEN, name + ".npz"))


def overcooked_params_from_golden(g):
    from madrona_rl_envs_playground_b200.overcooked_layouts import get_base_layout_params
    npl = int(g["num_players"])
    return get_base_layout_params(str(g["layout"]), int(g["horizon"]), None if npl < 0 else npl)


def legal_random_actions(rng, mask):
    logits = rng.random(mask.shape)
    logits[mask == 0] = -1.0
    return logits.argmax(-1)


def hanabi_informed_actions(rng, state, mask, active):
    """A policy that reads the standard game's state rows (own hand included) and plays to COMPLETE stacks while the
    information tokens are at their maximum — the only way to push them past it (sim.cpp:675-677 has no cap), which
    lengthens the rows and exercises the out-of-row spill.  Falls back to a random legal move."""
    N = state.shape[1]
    out = legal_random_actions(rng, mask).astype(np.int32)
    for n in range(N):
        p = int(active[:, n].argmax())
        st, legal = state[p, n], mask[p, n]
        own = st[658:783].reshape(5, 25)
        cards = np.where(own.any(1), own.argmax(1), -1)
        fw_rows = st[167:192].reshape(5, 5)
        fw = np.where(fw_rows.any(1), fw_rows.argmax(1) + 1, 0)
        info = int(st[192:200].sum())
        playable = [i for i in range(5) if cards[i] >= 0 and cards[i] % 5 == fw[cards[i] // 5]]
        fives = [i for i in playable if cards[i] % 5 == 4]
        others = [i for i in playable if cards[i] % 5 != 4]
        choice = None
        if others:
            choice = 5 + others[0]
        elif fives:
            if info >= 8:
                choice = 5 + fives[0]
            else:
                spare = [i for i in range(5) if i not in fives and legal[i]]
                choice = spare[0] if spare else 5 + fives[0]
        elif info < 8 and legal[0]:
            choice = 0
        if choice is not None and legal[choice]:
            out[p, n] = choice
    return out


def biased_overcooked_actions(rng, shape):
    return rng.choice(6, size=shape, p=[0.14, 0.14, 0.14, 0.14, 0.04, 0.40]).astype(np.int32)


class GpuSim:
    """numpy-facing wrapper over a package simulator: write actions, step through the C ABI, read tensors back."""

    def __init__(self, sim):
        import torch
        self.torch = torch
        self.sim = sim
        self.action = sim.action_tensor().to_torch()

    def t(self, name):
        return getattr(self.sim, name + "_tensor")().to_torch()

    def np(self, name):
        return self.t(name).cpu().numpy()

    def step(self, actions):
        a = self.torch.as_tensor(np.ascontiguousarray(actions, dtype=np.int32)).reshape(self.action.shape)
        self.action.copy_(a.to(self.action.device))
        self.sim.step()


def random_policy_hash(seed, step, idx):
    """numpy restatement of b200::policy::mix (csrc/random_policy.cu); checked against the library's own export."""
    idx = np.asarray(idx, dtype=np.uint64)
    m = np.uint64(0xFFFFFFFF)
    h = (np.uint64(seed) ^ ((np.uint64(step) * np.uint64(0x9E3779B9)) & m)
         ^ ((idx * np.uint64(0x85EBCA6B) + np.uint64(0xC2B2AE35)) & m)) & m
    h ^= h >> np.uint64(16)
    h = (h * np.uint64(0x7FEB352D)) & m
    h ^= h >> np.uint64(15)
    h = (h * np.uint64(0x846CA68B)) & m
    h ^= h >> np.uint64(16)
    return h


def random_policy_actions(seed, step, shape, mask=None, num_moves=None):
    """Actions b200env_random_rollout samples at `step`: flat element i takes legal move number
    floor(hash * legal / 2^32); `mask` is [..., M] (None = all `num_moves` legal)."""
    total = int(np.prod(shape))
    r = random_policy_hash(seed, step, np.arange(total))
    if mask is None:
        return ((r * np.uint64(num_moves)) >> np.uint64(32)).astype(np.int32).reshape(shape)
    m = (np.asarray(mask).reshape(total, -1) != 0)
    legal = m.sum(axis=1).astype(np.uint64)
    pick = ((r * legal) >> np.uint64(32)).astype(np.int64)
    order = np.cumsum(m, axis=1) - 1                     # rank of each legal move
    hit = m & (order == pick[:, None])
    choice = np.where(legal > 0, hit.argmax(axis=1), 0)
    return choice.astype(np.int32).reshape(shape)


def hanabi_row_lengths(cfg):
    """Encoded lengths of a 2-player observation / state row for a rule set (sim.cpp:54-365); what lies beyond them in a
    658 / 783-element row is never written by the reference."""
    c, r = cfg.get("colors", 5), cfg.get("ranks", 5)
    bits = c * r
    obs_len = (5 * bits + 2) + ((4 + (r - 2) * 2) * c - 10) + bits + cfg.get("max_information_tokens", 8) \
        + cfg.get("max_life_tokens", 3) + c * (3 + 2 * (r - 2) + 1) + (2 + 4 + 2 + c + r + 5 + 5 + bits + 2) \
        + 2 * 5 * (bits + c + r)
    return obs_len, obs_len + 5 * bits
