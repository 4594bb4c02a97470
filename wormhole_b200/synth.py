"""Synthetic minibatch generators for the benchmark workloads named in BASELINE.json
(SURVEY.md section 8d).  Counter-based: minibatch m of a stream is a pure function of
(seed, m), so the CPU oracle and every GPU rank can regenerate identical inputs.

criteo_minibatch: B rows x 39 fields (13 "integer" + 26 "categorical"), one id per field,
binary features (value = None).  Raw id = (splitmix64(field, rank) >> 10) | (field << 54),
the bit layout learn/base/criteo_parser.h:70,82 produces; rank within a field is truncated
Zipf(1.1) by inverse CDF; labels come from a hidden linear model so logloss/AUC move.

kdd_minibatch: nnz/row = 1 + Poisson(10), one vocabulary of 54,686,452 ids, Zipf(1.0).
"""
import numpy as np

from .rowblock import RowBlock

M64 = np.uint64(0xFFFFFFFFFFFFFFFF)

# public Criteo-Kaggle categorical vocabulary sizes (C1..C26)
CRITEO_CAT = [1460, 583, 10131227, 2202608, 305, 24, 12517, 633, 3, 93145, 5683, 8351593, 3194,
              27, 14992, 5461306, 10, 5652, 2173, 4, 7046547, 18, 15, 286181, 105, 142572]
# 13 integer fields: geometric 64 .. 4096
CRITEO_INT = [int(round(64 * (4096 / 64) ** (i / 12.0))) for i in range(13)]
CRITEO_CARD = np.array(CRITEO_INT + CRITEO_CAT, dtype=np.float64)
CRITEO_FIELDS = 39
KDD_VOCAB = 54686452


def splitmix64(x):
    """Vectorised splitmix64 finaliser on uint64 arrays."""
    with np.errstate(over="ignore"):
        x = (x + np.uint64(0x9E3779B97F4A7C15)) & M64
        z = x
        z = ((z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)) & M64
        z = ((z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)) & M64
        return z ^ (z >> np.uint64(31))


def _uniform(seed, stream, n, salt):
    """n doubles in [0,1) from counters (seed, stream, salt, i)."""
    with np.errstate(over="ignore"):
        base = splitmix64(np.uint64(seed) * np.uint64(0x100000001B3) + np.uint64(stream))
        base = splitmix64(base ^ np.uint64(salt))
        ctr = base + np.arange(n, dtype=np.uint64)
    return (splitmix64(ctr) >> np.uint64(11)).astype(np.float64) * (1.0 / (1 << 53))


def _zipf_rank(u, card, s):
    """Inverse CDF of the continuous relaxation of a Zipf(s) law truncated to [1, card]."""
    if abs(s - 1.0) < 1e-9:
        r = np.exp(u * np.log(card))
    else:
        a = 1.0 - s
        r = ((card ** a - 1.0) * u + 1.0) ** (1.0 / a)
    r = np.floor(r)
    return np.minimum(np.maximum(r, 1), card).astype(np.uint64) - np.uint64(1)


def _hidden_weight(ids):
    """w*_k = 0.3 * N(0,1) (Box-Muller on two splitmix64 draws of the raw id)."""
    a = (splitmix64(ids ^ np.uint64(0xA5A5A5A5)) >> np.uint64(11)).astype(np.float64) / (1 << 53)
    b = (splitmix64(ids ^ np.uint64(0x5A5A5A5A5A)) >> np.uint64(11)).astype(np.float64) / (1 << 53)
    return 0.3 * np.sqrt(-2.0 * np.log(1.0 - a)) * np.cos(2 * np.pi * b)


def criteo_minibatch(m, rows=100000, seed=42, bias=-2.2):
    """Minibatch m of the Criteo-shape stream. bias gives roughly 25% positives."""
    F = CRITEO_FIELDS
    u = _uniform(seed, m, rows * F, 1).reshape(rows, F)
    rank = np.empty((rows, F), np.uint64)
    for f in range(F):
        rank[:, f] = _zipf_rank(u[:, f], CRITEO_CARD[f], 1.1)
    field = np.arange(F, dtype=np.uint64)[None, :]
    with np.errstate(over="ignore"):
        h = splitmix64((field << np.uint64(40)) ^ rank)
    ids = (h >> np.uint64(10)) | (field << np.uint64(54))
    score = bias + _hidden_weight(ids).sum(axis=1)
    p = 1.0 / (1.0 + np.exp(-score))
    label = (_uniform(seed, m, rows, 2) < p).astype(np.float32)
    offset = np.arange(rows + 1, dtype=np.uint64) * np.uint64(F)
    return RowBlock(offset, ids.reshape(-1), None, label)


def kdd_minibatch(m, rows=1000000, seed=43, bias=-2.0):
    """Minibatch m of the KDD2012-shape stream (ragged rows, one power-law vocabulary)."""
    rng = np.random.Generator(np.random.Philox(key=[seed, m]))
    nnz = 1 + rng.poisson(10.0, rows)
    offset = np.zeros(rows + 1, np.uint64)
    offset[1:] = np.cumsum(nnz)
    n = int(offset[-1])
    ids = _zipf_rank(_uniform(seed, m, n, 3), float(KDD_VOCAB), 1.0)
    hw = _hidden_weight(ids)
    score = bias + np.add.reduceat(hw, offset[:-1].astype(np.int64))
    p = 1.0 / (1.0 + np.exp(-score))
    label = (_uniform(seed, m, rows, 4) < p).astype(np.float32)
    return RowBlock(offset, ids, None, label)


def criteo_text_minibatch(m, rows=100000, seed=42):
    """Minibatch m of the Criteo-shape stream as TEXT in the dataset's own schema (learn/base/criteo_parser.h:17-21):
    the same per-field Zipf ranks as criteo_minibatch, written as decimal integers (13 fields) and 8 hex digits
    (26 fields), tab separated, `label` first; labels are Bernoulli(0.25). Used to feed the GPU reader."""
    F = CRITEO_FIELDS
    u = _uniform(seed, m, rows * F, 1).reshape(rows, F)
    cols = [np.where(_uniform(seed, m, rows, 2) < 0.25, "1", "0")]
    for f in range(F):
        r = _zipf_rank(u[:, f], CRITEO_CARD[f], 1.1).astype(np.uint64)
        cols.append(np.char.mod("%d" if f < 13 else "%08x", r))
    lines = cols[0]
    for c in cols[1:]:
        lines = np.char.add(np.char.add(lines, "\t"), c)
    return ("\n".join(lines.tolist()) + "\n").encode()
