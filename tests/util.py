"""Shared helpers for the parity tests: seeded input builders."""
import numpy as np

from wormhole_b200.rowblock import RowBlock


def random_block(rng, rows, max_nnz, key_space, with_value=False, empty_rows=True, dup_in_row=True):
    """Ragged block with ids drawn (skewed) from [0, key_space)."""
    lo = 0 if empty_rows else 1
    nnz = rng.integers(lo, max_nnz + 1, rows)
    offset = np.zeros(rows + 1, np.uint64)
    offset[1:] = np.cumsum(nnz)
    n = int(offset[-1])
    # skewed: half of the entries from a small hot set
    hot = rng.integers(0, max(1, min(key_space, 50)), n, dtype=np.uint64)
    cold = rng.integers(0, key_space, n, dtype=np.uint64)
    ids = np.where(rng.random(n) < 0.5, hot, cold).astype(np.uint64)
    value = (rng.standard_normal(n).astype(np.float32)) if with_value else None
    label = (rng.random(rows) < 0.4).astype(np.float32)
    return RowBlock(offset, ids, value, label)


def wide_ids(rng, n):
    """Full-width 64-bit ids (exercise every radix digit)."""
    return rng.integers(0, 2 ** 63, n, dtype=np.uint64) * np.uint64(2) + rng.integers(0, 2, n, dtype=np.uint64)
