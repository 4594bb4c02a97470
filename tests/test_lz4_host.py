"""The library's LZ4 block decoder (host code, no GPU) against liblz4's own encoder and decoder (pyarrow's lz4_raw codec)."""
import ctypes as C

import numpy as np

from wormhole_b200 import capi


def test_lz4_block_decode_matches_liblz4():
    import pyarrow as pa
    L = capi.lib()
    L.wh_lz4_block_decode.restype = C.c_long
    rng = np.random.default_rng(3)
    cases = [b"", b"a", b"abc" * 1000, bytes(rng.integers(0, 256, 5000, dtype=np.uint8)),
             np.arange(0, 40000, 39, dtype=np.uint64).tobytes(),                      # offsets of a Criteo-shaped block
             (rng.random(20000) < .25).astype(np.float32).tobytes(),                  # labels
             bytes(rng.integers(0, 4, 70000, dtype=np.uint8)), b"\x00" * 300000]     # long matches, long literal runs
    for raw in cases:
        comp = pa.compress(raw, codec="lz4_raw", asbytes=True) if raw else b"\x00"
        out = C.create_string_buffer(max(1, len(raw)))
        n = L.wh_lz4_block_decode(comp, C.c_size_t(len(comp)), out, C.c_size_t(len(raw)))
        assert n == len(raw) and out.raw[:n] == raw
        assert raw == (pa.decompress(comp, decompressed_size=len(raw), codec="lz4_raw", asbytes=True) if raw else b"")
    # malformed: offset beyond the output so far, truncated input, output too small
    assert L.wh_lz4_block_decode(b"\x10a\x05\x00", C.c_size_t(4), C.create_string_buffer(64), C.c_size_t(64)) == -1
    comp = pa.compress(b"abc" * 1000, codec="lz4_raw", asbytes=True)
    assert L.wh_lz4_block_decode(comp[:-3], C.c_size_t(len(comp) - 3), C.create_string_buffer(3000), C.c_size_t(3000)) != 3000
    assert L.wh_lz4_block_decode(comp, C.c_size_t(len(comp)), C.create_string_buffer(100), C.c_size_t(100)) == -1
