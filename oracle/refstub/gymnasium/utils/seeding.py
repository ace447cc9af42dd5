"""gymnasium.utils.seeding.np_random is exactly this (gymnasium >= 0.26)."""
import numpy as np

RandomNumberGenerator = np.random.Generator


def np_random(seed=None):
    seed_seq = np.random.SeedSequence(seed)
    np_seed = seed_seq.entropy
    return np.random.Generator(np.random.PCG64(seed_seq)), np_seed
