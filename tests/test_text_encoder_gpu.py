"""GPU parity tests of the text encoder (csrc/text_encoder.cu through matcha_b200.text_encoder.TextEncoder):
against the golden outputs of the reference's own TextEncoder (tests/golden/text_encoder_golden.npz) and against
the torch restatement in fp64 on the same seeded inputs, at the bench shape and at tile-edge shapes.
Tolerance: rel-L2 <= 1e-5 (fp32 arithmetic; the north_star's fp32 bar)."""
import os

import numpy as np
import pytest
import torch

from oracle import text_encoder as oenc

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
TOL = 1e-5


def rel_l2(a, b):
    return float(torch.linalg.norm((a.double().cpu() - b.double().cpu())) / torch.linalg.norm(b.double().cpu()))


@pytest.fixture(scope="module")
def golden():
    return np.load(os.path.join(ROOT, "tests", "golden", "text_encoder_golden.npz"))


@pytest.fixture(scope="module")
def encoder(golden):
    from matcha_b200.text_encoder import TextEncoder
    n_vocab = int(golden["n_vocab"][0])
    m = TextEncoder(n_vocab=n_vocab)
    m.load_state_dict(oenc.make_weights(int(golden["weight_seed"][0]), "rich", n_vocab))
    return m


@pytest.fixture(scope="module")
def weights64(golden):
    w = oenc.make_weights(int(golden["weight_seed"][0]), "rich", int(golden["n_vocab"][0]))
    return oenc.cast_weights({k: v.cuda() for k, v in w.items()}, torch.float64)


@pytest.mark.parametrize("name", ["ragged_T50", "full_T33", "one_token", "long_T160"])
def test_matches_reference_golden(golden, encoder, name):
    ids = torch.from_numpy(golden[name + "/ids"]).cuda()
    lengths = torch.from_numpy(golden[name + "/lengths"]).cuda()
    encoder.use_graph = False
    mu, logw, mask = encoder(ids, lengths)
    assert mu.dtype == torch.float32 and mu.is_cuda and mu.shape == golden[name + "/mu"].shape
    assert torch.equal(mask.cpu(), torch.from_numpy(golden[name + "/mask"]))
    assert rel_l2(mu, torch.from_numpy(golden[name + "/mu"])) < TOL
    assert rel_l2(logw, torch.from_numpy(golden[name + "/logw"])) < TOL
    pad = mask[:, 0, :] == 0
    assert float(mu.transpose(1, 2)[pad].abs().sum()) == 0.0 and float(logw[:, 0, :][pad].abs().sum()) == 0.0


@pytest.mark.parametrize("B,T,frac", [(32, 128, 0.7), (3, 31, 0.5), (2, 257, 0.6), (5, 64, 1.0), (1, 1, 1.0), (4, 400, 0.7)])
def test_matches_fp64_oracle(encoder, weights64, B, T, frac):
    """Bench shape (32 x 128 ragged), rows not a multiple of the 32-row tile, several attention key blocks, a single token,
    and T_text = 400 (BASELINE configs[1]'s upper end)."""
    g = torch.Generator().manual_seed(B * 1000 + T)
    ids = torch.randint(0, encoder.vocab_size, (B, T), generator=g).cuda()
    lens = torch.tensor([T] + [max(1, int(T * (frac + (1 - frac) * float(torch.rand((), generator=g))))) for _ in range(B - 1)])
    encoder.use_graph = False
    mu, logw, mask = encoder(ids, lens.cuda())
    omu, ologw, omask = oenc.text_encoder_forward(weights64, ids, lens.cuda())
    assert torch.equal(mask.double(), omask)
    assert rel_l2(mu, omu) < TOL
    assert rel_l2(logw, ologw) < TOL


def test_graph_replay_equals_plain_launches(encoder):
    g = torch.Generator().manual_seed(9)
    B, T = 6, 70
    encoder.use_graph = False
    batches = []
    for _ in range(3):
        ids = torch.randint(0, encoder.vocab_size, (B, T), generator=g).cuda()
        lens = torch.tensor([T, 60, 50, 44, 33, 7]).cuda()
        batches.append((ids, lens, encoder(ids, lens)))
    encoder.use_graph = True
    encoder._graphs.clear(); encoder._seen.clear()
    for ids, lens, want in batches:                     # 1st call plain, 2nd captures, 3rd replays
        got = encoder(ids, lens)
        assert all(torch.equal(a, b) for a, b in zip(got, want))
    assert len(encoder._graphs) == 1
    encoder.use_graph = False


def test_int32_ids_and_cpu_lengths(encoder):
    g = torch.Generator().manual_seed(4)
    ids = torch.randint(0, encoder.vocab_size, (2, 20), generator=g)
    a = encoder(ids.cuda(), torch.tensor([20, 11]).cuda())
    b = encoder(ids.int().cuda(), torch.tensor([20, 11]))          # lengths on the host, int32 ids
    assert all(torch.equal(x, y) for x, y in zip(a, b))


def test_c_abi_argument_checks(encoder):
    import ctypes
    from matcha_b200._lib import load_library
    lib = load_library()
    ws = torch.empty(int(lib.matcha_b200_text_encoder_workspace_bytes(1, 8)), dtype=torch.uint8, device="cuda")
    assert lib.matcha_b200_text_encoder_forward(None, None, None, 178, None, None, None, None, None, 1, 8, ws.data_ptr(), ws.numel(), None) == -1
    assert lib.matcha_b200_text_encoder_workspace_bytes(0, 8) == -1
    assert lib.matcha_b200_text_encoder_blob_floats(178) == sum(int(np.prod(s)) for s in oenc.param_shapes(178).values()) + 3
