"""Minimal observation / action space descriptors.

The reference uses ``gym.spaces`` (absent from this image).  When gym (or gymnasium) is importable its classes are
used, otherwise these stand-ins carry the same attributes the reference's trainers read (``shape``, ``n``, ``nvec``,
``low``/``high``, ``dtype``, ``sample()``).
"""
import numpy as np

try:   # pragma: no cover - depends on the environment
    from gym.spaces import Box, Discrete, MultiBinary, MultiDiscrete   # noqa: F401
except Exception:   # pragma: no cover
    try:
        from gymnasium.spaces import Box, Discrete, MultiBinary, MultiDiscrete   # noqa: F401
    except Exception:
        class _Space:
            shape = ()
            dtype = None

            def __repr__(self):
                return f"{type(self).__name__}{self.shape}"

        class Discrete(_Space):
            def __init__(self, n):
                self.n, self.shape, self.dtype = int(n), (), np.int64

            def sample(self):
                return int(np.random.randint(self.n))

        class MultiBinary(_Space):
            def __init__(self, n):
                self.n = n
                self.shape = tuple(int(v) for v in np.atleast_1d(n))
                self.dtype = np.int8

            def sample(self):
                return np.random.randint(0, 2, self.shape).astype(np.int8)

        class MultiDiscrete(_Space):
            def __init__(self, nvec):
                self.nvec = np.asarray(nvec, dtype=np.int64)
                self.shape, self.dtype = self.nvec.shape, np.int64

            def sample(self):
                return (np.random.random_sample(self.nvec.shape) * self.nvec).astype(np.int64)

        class Box(_Space):
            def __init__(self, low, high, dtype=np.float32):
                self.low, self.high = np.asarray(low, dtype=dtype), np.asarray(high, dtype=dtype)
                self.shape, self.dtype = self.low.shape, dtype

            def sample(self):
                return np.random.uniform(np.maximum(self.low, -1e6), np.minimum(self.high, 1e6)).astype(self.dtype)
