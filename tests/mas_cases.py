"""Shared helpers for the MAS parity tests (test infrastructure)."""
import importlib.util
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_spec = importlib.util.spec_from_file_location("make_mas_golden", os.path.join(_HERE, "golden", "make_mas_golden.py"))
gen = importlib.util.module_from_spec(_spec)
_spec.loader.exec_module(gen)


def golden_names(blob):
    return sorted({k.split("/")[0] for k in blob.files if k.endswith("/value")})


def golden_case(blob, name):
    v = blob[name + "/value"]
    m = blob[name + "/mask"].astype(np.float32)
    p = np.unpackbits(blob[name + "/path_bits"], axis=-1)[..., : v.shape[-1]].astype(np.int32)
    return v, m, p


def golden_full(blob):
    assert int(blob["full_16x128x800/seed"][0]) == gen.FULL_SEED
    v, m = gen.full_case()
    p = np.unpackbits(blob["full_16x128x800/path_bits"], axis=-1)[..., : v.shape[-1]].astype(np.int32)
    return v, m, p


def check_path_properties(path, t_xs, t_ys):
    """Size-independent invariants of a monotonic alignment (core.pyx:87-92)."""
    B = path.shape[0]
    for b in range(B):
        tx, ty = int(t_xs[b]), int(t_ys[b])
        p = path[b]
        assert p[tx:, :].sum() == 0 and p[:, ty:].sum() == 0
        if tx < 1 or ty < 1 or tx > ty:
            continue
        col = p[:tx, :ty].sum(0)
        assert (col == 1).all(), "each valid frame has exactly one token"
        tok = p[:tx, :ty].argmax(0)
        assert tok[0] == 0 and tok[-1] == tx - 1
        d = np.diff(tok)
        assert ((d == 0) | (d == 1)).all(), "monotone, no skips"
