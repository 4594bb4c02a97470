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


# Shares and sizes the parity tests allow for the two counted causes in assert_state_close: every other entry meets
# rtol * |want| + 1e-7. Measured on B200 over the whole GPU suite (profiles/r2_tolerances.md): no flips at all; stragglers
# in 10 of 560 comparisons, at most 3 entries or 3.3e-4 of the entries of one comparison, the largest 6.7e-6 in size.
STRAGGLERS = 5e-4
FLIP_ABS = 1e-5


def assert_state_close(got, want, rtol=1e-5, atol=1e-7, what="", stragglers=0.0, flip_abs=0.0, loose_factor=1.0):
    """Model state against the checker: |got - want| <= rtol * |want| + atol per entry (1e-5 relative
    with an absolute floor of 1e-7). Entries outside that bar are not waved through by a wider tolerance; they are
    counted, by cause, and each cause has its own bound:
      flips       one side is exactly 0 and the other is not: the key sits on the L1 threshold of L1L2::Solve
                  (learn/linear/penalty.h:36-41) and the last-bit difference of z decides. At most `flip_abs` in size.
      stragglers  entries whose gradient sums cancel: the reference adds each X^T dual output in fp32 in row order
                  (learn/base/spmv.h:103-117) while the device adds exactly, so the two differ by the reference's own
                  rounding, which is relative to the sum of |terms|, not to the result. Bounded by
                  loose_factor * rtol * max(|want|, 1), and their number by max(3, `stragglers` x entries).
    Returns (n_flips, n_stragglers, n) for callers that want to report them."""
    a = np.asarray(got, np.float64).ravel()
    b = np.asarray(want, np.float64).ravel()
    assert a.shape == b.shape, (what, a.shape, b.shape)
    err = np.abs(a - b)
    out = err > rtol * np.abs(b) + atol
    flip = out & ((a == 0) != (b == 0))
    rest = out & ~flip
    n = max(1, a.size)
    import os
    rep = os.environ.get("WH_TOL_REPORT")   # calibration runs: log the counts instead of judging them
    if rep:
        with open(rep, "a") as f:
            f.write("%s n=%d flips=%d flip_max=%g stragglers=%d share=%.3g worst_err=%g worst_loose=%g\n" % (
                what, a.size, int(flip.sum()), float(np.maximum(np.abs(a[flip]), np.abs(b[flip])).max()) if flip.any() else 0.0,
                int(rest.sum()), rest.sum() / n, float(err[rest].max()) if rest.any() else 0.0,
                float((err[rest] / (rtol * np.maximum(np.abs(b[rest]), 1.0))).max()) if rest.any() else 0.0))
        return int(flip.sum()), int(rest.sum()), a.size
    if flip.any():
        worst = float(np.maximum(np.abs(a[flip]), np.abs(b[flip])).max())
        assert worst <= flip_abs, "%s: %d entries flipped across the L1 threshold, the largest is %g (> %g)" % (
            what, int(flip.sum()), worst, flip_abs)
    if rest.any():
        loose = loose_factor * rtol * np.maximum(np.abs(b[rest]), 1.0)
        assert (err[rest] <= loose).all(), "%s: worst %g x the loose bound" % (what, float((err[rest] / loose).max()))
        share = rest.sum() / n
        assert rest.sum() <= max(3, stragglers * n), "%s: %d of %d entries (%.3g) miss rtol=%g atol=%g, allowed share %.3g; worst err %g" % (
            what, int(rest.sum()), a.size, share, rtol, atol, stragglers, float(err[rest].max()))
    return int(flip.sum()), int(rest.sum()), a.size
