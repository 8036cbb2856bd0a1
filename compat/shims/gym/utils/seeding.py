import numpy as np


def np_random(seed=None):
    """gym 0.26.2: a PCG64 ``numpy.random.Generator`` seeded through ``SeedSequence``; returns (rng, seed)."""
    if seed is not None and not (isinstance(seed, (int, np.integer)) and seed >= 0):
        raise ValueError(f"Seed must be a non-negative integer or omitted, not {seed}")
    seed_seq = np.random.SeedSequence(None if seed is None else int(seed))
    return np.random.Generator(np.random.PCG64(seed_seq)), seed_seq.entropy
