"""CPU tests: pin the torch restatement of the text encoder (oracle/text_encoder.py) against golden outputs
produced by the reference's own TextEncoder (tests/golden/make_text_encoder_golden.py)."""
import os

import numpy as np
import pytest
import torch

from oracle import text_encoder as oenc

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CASES = ["ragged_T50", "full_T33", "one_token", "long_T160"]


@pytest.fixture(scope="module")
def golden():
    return np.load(os.path.join(ROOT, "tests", "golden", "text_encoder_golden.npz"))


@pytest.fixture(scope="module")
def weights(golden):
    return oenc.make_weights(int(golden["weight_seed"][0]), "rich", int(golden["n_vocab"][0]))


def rel_l2(a, b):
    return float(torch.linalg.norm((a - b).double()) / torch.linalg.norm(b.double()))


def test_param_count_matches_reference():
    assert sum(int(np.prod(s)) for s in oenc.param_shapes(178).values()) == 7_195_345     # 7.2 M, SURVEY §8f-4


@pytest.mark.parametrize("name", CASES)
def test_text_encoder_matches_reference(golden, weights, name):
    torch.set_num_threads(4)
    ids, lengths = torch.from_numpy(golden[name + "/ids"]), torch.from_numpy(golden[name + "/lengths"])
    mu, logw, mask = oenc.text_encoder_forward(weights, ids, lengths)
    assert torch.equal(mask, torch.from_numpy(golden[name + "/mask"]))
    assert rel_l2(mu, torch.from_numpy(golden[name + "/mu"])) < 2e-6
    assert rel_l2(logw, torch.from_numpy(golden[name + "/logw"])) < 2e-6
    # padded tokens are exactly zero in both outputs (text_encoder.py:97, 398)
    pad = mask[:, 0, :] == 0
    assert float(mu.transpose(1, 2)[pad].abs().sum()) == 0.0 and float(logw[:, 0, :][pad].abs().sum()) == 0.0


def test_fp64_restatement_is_within_fp32_noise(golden, weights):
    name = "ragged_T50"
    w64 = oenc.cast_weights(weights, torch.float64)
    mu, logw, _ = oenc.text_encoder_forward(w64, torch.from_numpy(golden[name + "/ids"]), torch.from_numpy(golden[name + "/lengths"]))
    assert rel_l2(mu.float(), torch.from_numpy(golden[name + "/mu"])) < 3e-6
    assert rel_l2(logw.float(), torch.from_numpy(golden[name + "/logw"])) < 5e-6


def test_padding_leaks_through_the_ffn_as_in_the_reference(weights):
    """Reference behaviour worth knowing: FFN.forward masks only the FIRST conv's input (text_encoder.py:259), so
    relu(conv1) is non-zero on the first padded token (bias + the last valid tokens) and the second k=3 conv reads
    it at the last valid token.  An utterance therefore encodes differently alone and padded inside a longer batch;
    the CUDA path has to compute the hidden activation on padded rows too, not skip them."""
    g = torch.Generator().manual_seed(2)
    ids = torch.randint(0, 178, (1, 20), generator=g)
    a = oenc.text_encoder_forward(weights, ids, torch.tensor([20]))
    ids2 = torch.cat([ids, torch.randint(0, 178, (1, 13), generator=g)], 1)
    b = oenc.text_encoder_forward(weights, ids2, torch.tensor([20]))
    assert rel_l2(b[0][:, :, :20], a[0]) > 1e-3
    # ... but WHICH ids sit in the padding is irrelevant (the embedding is masked before anything reads it)
    ids3 = torch.cat([ids, torch.randint(0, 178, (1, 13), generator=g)], 1)
    c = oenc.text_encoder_forward(weights, ids3, torch.tensor([20]))
    assert torch.equal(b[0], c[0]) and torch.equal(b[1], c[1])
