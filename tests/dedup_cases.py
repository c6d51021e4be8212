"""Pose batches for the insertNode / de-duplication tests (shared by the CPU and GPU suites)."""
import numpy as np


def near_duplicate_batch(n, seed, n_base=None, scale=5.0):
    """Random poses drawn from a small pool, exact repeats mixed with offsets just below, at and just
    above the reference's 1e-12 tolerance, in one or in all components (chains a ~ b ~ c arise)."""
    rng = np.random.RandomState(seed)
    n_base = n_base or max(n // 10, 1)
    base = rng.uniform(-scale, scale, (n_base, 7))
    out = np.empty((n, 7))
    for i in range(n):
        b = base[rng.randint(n_base)].copy()
        k = rng.randint(4)
        if k == 1:
            b[rng.randint(7)] += rng.choice([0.6e-12, -0.6e-12, 0.99e-12, 1.01e-12, 2e-12])
        elif k == 2:
            b += rng.uniform(-0.7e-12, 0.7e-12, 7)
        out[i] = b
    return out
