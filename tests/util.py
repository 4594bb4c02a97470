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


def criteo_text(rng, rows, crlf=False, trailing_newline=True, short_lines=True):
    """Synthetic text in the Criteo schema (learn/base/criteo_parser.h:17-21): label, 13 integer fields, 26
    categorical fields of 8 hex digits, tab separated; fields are empty with some probability, blank lines occur,
    and (short_lines) some rows stop early after a non-empty categorical field, which the reference accepts."""
    out = []
    for _ in range(rows):
        f = [str(int(rng.integers(0, 2)))]
        for _i in range(13):
            f.append("" if rng.random() < 0.2 else str(int(rng.integers(-3, 50000) if rng.random() < .8 else rng.integers(0, 10 ** 12))))
        ncat = 26
        if short_lines and rng.random() < 0.05:
            ncat = int(rng.integers(1, 26))
        cats = [("" if rng.random() < 0.15 else "%08x" % int(rng.integers(0, 2 ** 32 if rng.random() < .5 else 40)))
                for _i in range(ncat)]
        if ncat < 26 and cats[-1] == "":
            cats[-1] = "%08x" % int(rng.integers(0, 2 ** 32))
        out.append("\t".join(f + cats))
        if rng.random() < 0.03:
            out.append("")                       # blank line
    nl = "\r\n" if crlf else "\n"
    s = nl.join(out)
    if trailing_newline:
        s += nl
    return s.encode()


def libsvm_text(rng, rows, with_value="values", key_space=2 ** 40):
    """Synthetic libsvm text: `label idx[:val] ...`, some blank lines, rows without features, mixed separators.
    with_value: "values" (general floats), "ones" (every token `idx:1`) or "bare" (`idx` alone)."""
    out = []
    for _ in range(rows):
        n = int(rng.integers(0, 12))
        toks = [rng.choice(["1", "0", "-1", "+1", "0.5", "1e0"])]
        for _i in range(n):
            k = int(rng.integers(0, key_space))
            if with_value == "values":
                v = rng.choice(["1", "0.5", "-2.25", "3e-2", "1.5E+1", "%.6g" % rng.normal(), "%d" % rng.integers(-9, 99), ".5", "7."])
                toks.append("%d:%s" % (k, v))
            else:
                toks.append("%d" % k if with_value == "bare" else "%d:1" % k)
        sep = " " if rng.random() < .9 else "\t"
        out.append(sep.join(toks) + (" " if rng.random() < .1 else ""))
        if rng.random() < 0.03:
            out.append("  " if rng.random() < .5 else "")
    return ("\n".join(out) + "\n").encode()


def agaricus_text():
    """learn/data/agaricus.txt.train rebuilt from the committed arrays (`label idx:1 ...`, one row per line)."""
    import os
    z = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "agaricus_train.npz"))
    off, idx, lab = z["offset"].astype(np.int64), z["index"], z["label"]
    lines = []
    for r in range(len(lab)):
        lines.append(" ".join([str(int(lab[r]))] + ["%d:1" % int(i) for i in idx[off[r]:off[r + 1]]]))
    return ("\n".join(lines) + "\n").encode()


def adfea_text(rng, rows):
    """Synthetic text in the adfea layout (learn/base/adfea_parser.h): `lineid count label idx:gid ...`, one row per
    line, some blank lines and rows without features."""
    out = []
    for r in range(rows):
        n = int(rng.integers(0, 15))
        toks = [str(r + 1), str(n), str(int(rng.integers(0, 2)))]
        toks += ["%d:%d" % (int(rng.integers(0, 2 ** 62)), int(rng.integers(0, 1000))) for _ in range(n)]
        out.append((" " if rng.random() < .8 else "\t").join(toks))
        if rng.random() < 0.03:
            out.append("")
    return ("\n".join(out) + "\n").encode()
