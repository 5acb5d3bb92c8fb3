"""B200-native batched environment stepper for the Cartpole / Acrobat / Balance Beam / Hanabi / Overcooked environments of
willwng/madrona_rl_envs_playground.  The step path is a C-ABI library of hand-written sm_100a kernels
(``libb200envs.so``, include/b200envs.h); this package is the thin host side that keeps the reference's Python API."""
from . import simulators
from .simulators import (AcrobatSimulator, BalanceBeamSimulator, CartpoleSimulator, ExecMode, HanabiSimulator, OvercookedSimulator,
                         ReferenceLayoutOvercookedSimulator, SimplecookedSimulator, madrona, stream_gate)

__all__ = ["simulators", "ExecMode", "madrona", "OvercookedSimulator", "SimplecookedSimulator", "HanabiSimulator", "BalanceBeamSimulator",
           "CartpoleSimulator", "AcrobatSimulator", "ReferenceLayoutOvercookedSimulator", "stream_gate"]
