"""Pins oracle/alignment.py against tensors recorded from the reference's own MatchaTTS.forward /
synthesise (tests/golden/alignment_golden.npz, made by tests/golden/make_alignment_golden.py)."""
import os

import numpy as np
import pytest
import torch

from oracle import alignment as oal

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def gold():
    return np.load(os.path.join(ROOT, "tests", "golden", "alignment_golden.npz"))


def _t(g, key):
    return torch.from_numpy(g[key])


@pytest.mark.parametrize("case", ["train_a", "train_b"])
def test_training_seams_match_reference(gold, case):
    mu_x, y = _t(gold, f"{case}/mu_x"), _t(gold, f"{case}/y")
    xl, yl = _t(gold, f"{case}/x_lengths"), _t(gold, f"{case}/y_lengths")
    r = oal.align(mu_x, y, xl, yl)
    # same torch build, same ops, same order: the CPU restatement is bit-identical
    assert torch.equal(r["log_prior"], _t(gold, f"{case}/log_prior"))
    assert torch.equal(r["attn"], _t(gold, f"{case}/attn"))
    assert torch.equal(r["logw_"], _t(gold, f"{case}/logw_"))
    assert torch.equal(r["mu_y"], _t(gold, f"{case}/mu_y"))
    torch.testing.assert_close(r["prior_loss"], _t(gold, f"{case}/prior_loss"), rtol=1e-6, atol=0)
    dl = oal.duration_loss(_t(gold, f"{case}/logw"), r["logw_"], xl)
    torch.testing.assert_close(dl, _t(gold, f"{case}/dur_loss"), rtol=1e-6, atol=0)


@pytest.mark.parametrize("case", ["train_a", "train_b"])
def test_c_restatement_and_compiled_reference_agree(gold, case):
    from oracle import mas as omas
    if not omas.ref_available():
        pytest.skip("oracle/_ref not built (needs /root/reference)")
    lp, m = _t(gold, f"{case}/log_prior"), _t(gold, f"{case}/attn_mask")
    assert torch.equal(oal.maximum_path(lp, m, use_ref=True), oal.maximum_path(lp, m))


@pytest.mark.parametrize("case", ["synth_a", "synth_b"])
def test_synthesis_prestep_matches_reference(gold, case):
    mu_x, logw, x_mask = (_t(gold, f"{case}/{k}") for k in ("mu_x", "logw", "x_mask"))
    ls = float(gold[f"{case}/length_scale"])
    r = oal.synthesis_prestep(mu_x, logw, x_mask, ls)
    assert torch.equal(r["y_lengths"], _t(gold, f"{case}/mel_lengths"))
    assert torch.equal(r["mu_y"], _t(gold, f"{case}/mu_y"))
    assert torch.equal(r["y_mask"], _t(gold, f"{case}/y_mask"))
    assert torch.equal(r["attn"].squeeze(1), _t(gold, f"{case}/gp_path"))
    # matcha_tts.py:241 slices attn[:, :, :y_max_length], i.e. the TOKEN axis of [B,1,T_x,T_y] (a no-op here)
    assert torch.equal(r["attn"][:, :, : r["y_max_length"]], _t(gold, f"{case}/attn"))
    # generate_path on its own, from the recorded call arguments (utils/model.py:77-114)
    assert torch.equal(oal.generate_path(_t(gold, f"{case}/gp_duration"), _t(gold, f"{case}/gp_mask")),
                       _t(gold, f"{case}/gp_path"))


def test_fix_len_compatibility():
    assert [oal.fix_len_compatibility(n) for n in (1, 4, 5, 799, 800)] == [4, 4, 8, 800, 800]
