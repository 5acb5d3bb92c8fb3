from .vectoragent import RandomVectorAgent, VectorAgent
from .vectorenv import MadronaEnv, PlayerException, VectorMultiAgentEnv
from .vectorobservation import VectorObservation

__all__ = ["VectorAgent", "RandomVectorAgent", "VectorMultiAgentEnv", "MadronaEnv", "PlayerException",
           "VectorObservation"]
