import numpy as np


def np_random(seed=None):
    """gym.utils.seeding.np_random: returns (RandomState, seed). An int seed is passed through;
    None draws a fresh 31-bit seed (the reference then calls speedups.seed(seed) with it)."""
    if seed is None:
        seed = int(np.random.SeedSequence().generate_state(1)[0] & 0x7FFFFFFF)
    return np.random.RandomState(seed), seed
