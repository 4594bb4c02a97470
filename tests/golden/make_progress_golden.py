"""Writes tests/golden/golden_progress.json: progress-table rows printed by the reference's own Progress structs
(learn/linear/progress.h, learn/difacto/progress.h, through oracle/_ref) for a fixed list of merged progress
vectors. Run where /root/reference exists: python tests/golden/make_progress_golden.py"""
import ctypes as C
import json
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))


def vectors():
    rng = np.random.default_rng(99)
    lin, dif = [], []
    for i in range(12):
        ex = float(rng.choice([0, 1, 37, 1000, 6513, 100000, 12345678, 4.5e9]))
        cnt = float(rng.integers(1, 40))
        lin.append([ex * rng.uniform(0.05, 0.9), cnt * rng.uniform(.5, 1), cnt * rng.uniform(.5, 1), cnt, ex,
                    float(rng.integers(-50, 300000))])
        dif.append([ex * rng.uniform(0.05, 0.9), cnt * rng.uniform(.5, 1), ex * rng.uniform(0.05, 0.9), 0.0, cnt, ex,
                    float(rng.integers(-50, 300000)), float(rng.integers(0, 5000000))])
    return lin, dif


def main():
    lin, dif = vectors()
    L = C.CDLL(os.path.join(ROOT, "oracle", "_ref", "libwh_ref_linear.so"))
    D = C.CDLL(os.path.join(ROOT, "oracle", "_ref", "libwh_ref_difacto.so"))
    buf = C.create_string_buffer(512)
    out = {"linear": {"in": lin, "rows": []}, "difacto": {"in": dif, "rows": []}}
    L.ref_linear_progress_head(buf, 512)
    out["linear"]["head"] = buf.value.decode()
    D.ref_difacto_progress_head(buf, 512)
    out["difacto"]["head"] = buf.value.decode()
    t, n = C.c_double(0), C.c_double(0)
    for v in lin:
        L.ref_linear_progress_row((C.c_double * 6)(*v), C.byref(t), C.byref(n), buf, 512)
        out["linear"]["rows"].append(buf.value.decode())
    carry = (C.c_double * 3)(0, 0, 0)
    for v in dif:
        D.ref_difacto_progress_row((C.c_double * 8)(*v), carry, buf, 512)
        out["difacto"]["rows"].append(buf.value.decode())
    json.dump(out, open(os.path.join(HERE, "golden_progress.json"), "w"), indent=1)


if __name__ == "__main__":
    main()
