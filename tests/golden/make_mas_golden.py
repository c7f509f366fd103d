"""Generates tests/golden/mas_golden.npz from the reference's OWN compiled core.pyx.

Run in the build container (needs oracle/_ref/, i.e. /root/reference at `make -C oracle`
time):   python tests/golden/make_mas_golden.py

Every case stores the inputs the reference saw (value, mask) -- or, for the full-size
case, the numpy seed that regenerates them -- and the int path it returned, bit-packed.
The wrapper semantics are /root/reference/matcha/utils/monotonic_align/__init__.py:40-55,
the core is core.pyx:16-128.  Case list follows SURVEY.md §8(c).
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import mas as omas  # noqa: E402


def ragged_mask(rng, B, T_x, T_y, lo=0.7):
    """test_monotonic_align_speed.py:17-26, with t_x<=t_y enforced (the reference's domain)."""
    mask = np.ones((B, T_x, T_y), dtype=np.float32)
    for i in range(B):
        t_x = max(1, int(T_x * (lo + (1 - lo) * rng.random())))
        t_y = max(t_x, int(T_y * (lo + (1 - lo) * rng.random())))
        mask[i, t_x:, :] = 0
        mask[i, :, t_y:] = 0
    return mask


def fixed_mask(B, T_x, T_y, t_xs, t_ys):
    mask = np.zeros((B, T_x, T_y), dtype=np.float32)
    for i, (a, b) in enumerate(zip(t_xs, t_ys)):
        mask[i, :a, :b] = 1
    return mask


def cases():
    rng = np.random.default_rng(1234)
    out = {}
    # 1 random normal, ragged
    out["randn_ragged"] = (rng.standard_normal((4, 20, 50), dtype=np.float32), ragged_mask(rng, 4, 20, 50))
    # 2 all-zero: the tie rule of core.pyx:73
    out["all_zero_4x8"] = (np.zeros((1, 4, 8), np.float32), np.ones((1, 4, 8), np.float32))
    # 3 integer-quantised: dense ties
    v = np.round(rng.standard_normal((8, 33, 97), dtype=np.float32) * 2).astype(np.float32)
    out["quantised_ties"] = (v, ragged_mask(rng, 8, 33, 97))
    # 4 t_x == t_y (pure diagonal), several sizes incl. a 32-multiple
    out["square"] = (rng.standard_normal((4, 40, 40), dtype=np.float32), fixed_mask(4, 40, 40, [40, 32, 17, 1], [40, 32, 17, 1]))
    # 5 t_x == 1
    out["single_token"] = (rng.standard_normal((2, 5, 37), dtype=np.float32), fixed_mask(2, 5, 37, [1, 1], [37, 1]))
    # 6 t_y == t_x + 1
    out["one_slack"] = (rng.standard_normal((3, 35, 36), dtype=np.float32), fixed_mask(3, 35, 36, [35, 31, 2], [36, 32, 3]))
    # 7 log-prior-scaled: -1e2..-1e4 per cell, cumulative to ~-1e6
    v = -(10.0 ** rng.uniform(2, 4, size=(4, 24, 120))).astype(np.float32)
    out["logprior_scale"] = (v, ragged_mask(rng, 4, 24, 120))
    # 8 sentinel arithmetic: cumulative scores cross the -1e9 penalty
    v = (-3e8 * (1 + rng.random((2, 12, 40)))).astype(np.float32)
    out["below_penalty"] = (v, ragged_mask(rng, 2, 12, 40))
    # 9 non-finite values: NaN compares false, +-inf propagate (core.pyx:73)
    v = rng.standard_normal((3, 16, 48), dtype=np.float32)
    v[0, 3, 10] = np.nan; v[0, 7, 20:23] = np.nan; v[1, 5, 9] = np.inf; v[1, 6, 30] = -np.inf; v[2, :, 5] = np.nan
    out["non_finite"] = (v, ragged_mask(rng, 3, 16, 48))
    # 10 sizes straddling the kernel's 32-wide tiles, T_y not a multiple of 4
    out["tile_edges"] = (rng.standard_normal((5, 65, 131), dtype=np.float32),
                         fixed_mask(5, 65, 131, [65, 64, 33, 32, 31], [131, 128, 97, 96, 33]))
    # 11 the speed script's smallest shape (test_monotonic_align_speed.py:127)
    out["speed_8x50x200"] = (rng.standard_normal((8, 50, 200), dtype=np.float32), ragged_mask(rng, 8, 50, 200))
    # 12 more rows than one warp pass covers (kernel row-block boundary at 416)
    v = np.round(rng.standard_normal((2, 450, 520), dtype=np.float32) * 4).astype(np.float32) / 4
    out["tall_450x520"] = (v, fixed_mask(2, 450, 520, [450, 417], [520, 500]))
    # 13 interior zeros in the mask (not an outer product of prefixes): value*mask matters
    m = ragged_mask(rng, 2, 18, 60); m[0, 4:7, 10:30] = 0; m[1, 1, 5:9] = 0
    out["holey_mask"] = (rng.standard_normal((2, 18, 60), dtype=np.float32), m)
    return out


FULL_SEED = 20240501  # config 1 (BASELINE.json configs[0]): 16 x 128 x 800, ragged


def full_case():
    rng = np.random.default_rng(FULL_SEED)
    v = rng.standard_normal((16, 128, 800), dtype=np.float32)
    return v, ragged_mask(rng, 16, 128, 800)


def main():
    assert omas.ref_available(), "oracle/_ref missing: run `make -C oracle` first"
    blob = {}
    for name, (v, m) in cases().items():
        p = omas.maximum_path_ref(v, m)
        blob[name + "/value"] = v
        blob[name + "/mask"] = m.astype(np.uint8)
        blob[name + "/path_bits"] = np.packbits(p.astype(np.uint8), axis=-1)
        print(f"{name:18s} value{v.shape} ones={int(p.sum())}")
    v, m = full_case()
    p = omas.maximum_path_ref(v, m)
    blob["full_16x128x800/seed"] = np.array([FULL_SEED])
    blob["full_16x128x800/path_bits"] = np.packbits(p.astype(np.uint8), axis=-1)
    print(f"full_16x128x800    ones={int(p.sum())}")
    np.savez_compressed(os.path.join(os.path.dirname(__file__), "mas_golden.npz"), **blob)


if __name__ == "__main__":
    main()
