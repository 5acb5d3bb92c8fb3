from .balance_beam_env import BalanceMadronaTorch
from .cartpole_env import CartpoleMadronaTorch
from .hanabi_env import HanabiMadrona
from .overcooked_env import OvercookedMadrona

__all__ = ["OvercookedMadrona", "HanabiMadrona", "BalanceMadronaTorch", "CartpoleMadronaTorch"]
