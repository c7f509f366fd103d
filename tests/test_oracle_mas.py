"""CPU tests: pin the C restatement (oracle/mas_oracle.c) against the golden vectors made
from the reference's compiled core.pyx, and against that compiled core directly when
oracle/_ref is present."""
import numpy as np
import pytest

from oracle import mas as omas
from tests.mas_cases import check_path_properties, gen, golden_case, golden_full, golden_names


def test_golden_file_has_all_cases(mas_golden):
    assert set(golden_names(mas_golden)) == set(gen.cases().keys())


@pytest.mark.parametrize("name", sorted(gen.cases().keys()))
def test_oracle_matches_golden(mas_golden, name):
    v, m, p = golden_case(mas_golden, name)
    got = omas.maximum_path_oracle(v, m)
    assert np.array_equal(got, p)
    if name != "holey_mask":
        t_x, t_y = m.sum(1)[:, 0].astype(int), m.sum(2)[:, 0].astype(int)
        check_path_properties(got, t_x, t_y)


def test_oracle_matches_golden_full_config1(mas_golden):
    v, m, p = golden_full(mas_golden)
    got = omas.maximum_path_oracle(v, m)
    assert np.array_equal(got, p)


def test_tie_rule_is_core_pyx_not_python_fallback():
    # all-zero 4x8 -> token 0 holds frames 0..4
    p = omas.maximum_path_oracle(np.zeros((1, 4, 8), np.float32), np.ones((1, 4, 8), np.float32))[0]
    want = np.array([[1, 1, 1, 1, 1, 0, 0, 0], [0, 0, 0, 0, 0, 1, 0, 0], [0, 0, 0, 0, 0, 0, 1, 0], [0, 0, 0, 0, 0, 0, 0, 1]])
    assert np.array_equal(p, want)


def test_oracle_clobbers_scores_like_reference():
    # core.pyx:83-85 overwrites the score matrix with cumulative scores
    rng = np.random.default_rng(0)
    v = rng.standard_normal((1, 3, 5), dtype=np.float32)
    w = v.copy()
    path = np.zeros(v.shape, np.int32)
    omas.oracle_compute_batch_alignments(path, w, [3], [5])
    assert w[0, 0, 0] == v[0, 0, 0] and w[0, 0, 1] == np.float32(v[0, 0, 0] + v[0, 0, 1])
    assert w[0, 2, 0] == np.float32(-1e9)


def test_out_of_domain_lengths_leave_zero_path():
    v = np.zeros((3, 4, 4), np.float32)
    path = np.zeros(v.shape, np.int32)
    bad = omas.oracle_compute_batch_alignments(path, v, [0, 4, 2], [4, 3, 0])
    assert bad == 3 and path.sum() == 0


@pytest.mark.skipif(not omas.ref_available(), reason="oracle/_ref not built")
@pytest.mark.parametrize("seed", range(6))
def test_oracle_matches_compiled_reference_random(seed):
    rng = np.random.default_rng(100 + seed)
    B, T_x, T_y = int(rng.integers(1, 6)), int(rng.integers(1, 70)), int(rng.integers(70, 200))
    v = rng.standard_normal((B, T_x, T_y), dtype=np.float32)
    if seed % 2:
        v = np.round(v * 2).astype(np.float32)
    m = gen.ragged_mask(rng, B, T_x, T_y, lo=0.3)
    want_scores = np.ascontiguousarray((v * m).astype(np.float32))
    got_scores = want_scores.copy()
    t_x, t_y = m.sum(1)[:, 0].astype(np.int32), m.sum(2)[:, 0].astype(np.int32)
    want = np.zeros(v.shape, np.int32)
    got = np.zeros(v.shape, np.int32)
    omas.ref_compute_batch_alignments(want, want_scores, t_x, t_y)
    omas.oracle_compute_batch_alignments(got, got_scores, t_x, t_y)
    assert np.array_equal(got, want)
    assert np.array_equal(got_scores.view(np.uint32), want_scores.view(np.uint32))


@pytest.mark.parametrize("kind", ["bonus_row", "log_prior_scale"])
def test_oracle_matches_compiled_reference_band_edges(kind):
    """Paths that run along the band edges (core.pyx:58-59) with log-prior-sized scores: the shapes that
    exposed a kernel bug in round 1 (tests/test_mas_gpu.py::test_band_edge_paths_vs_oracle)."""
    if not omas.ref_available():
        pytest.skip("oracle/_ref not built (needs /root/reference)")
    rng = np.random.default_rng(7)
    v = rng.standard_normal((3, 128, 800), dtype=np.float32)
    if kind == "log_prior_scale":
        v = (v * 10 - 150).astype(np.float32)
    v[:, 40, :] += 5.0
    m = gen.ragged_mask(rng, 3, 128, 800)
    m[0] = 1
    assert np.array_equal(omas.maximum_path_oracle(v, m), omas.maximum_path_ref(v, m))
