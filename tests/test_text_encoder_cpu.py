"""CPU tests of the text encoder's host logic: state_dict ingest (incl. the MatchaTTS "encoder." prefix), weight
packing against the library's blob size, the RoPE tables, and that nothing silently falls back to the CPU."""
from types import SimpleNamespace

import pytest
import torch

from matcha_b200 import text_encoder as te
from oracle import text_encoder as oenc


def test_expected_keys_are_the_reference_state_dict():
    assert dict(te.expected_keys(178)) == dict(oenc.param_shapes(178))
    assert list(te.expected_keys(150)) == list(oenc.param_shapes(150))


def test_pack_weights_layout_and_size():
    w = oenc.make_weights(3, "rich", 150)
    blob = te.pack_weights(w, 150)
    assert blob.dtype == torch.float32 and blob.numel() == sum(v.numel() for v in w.values()) + 3    # dp bias padded to 4
    assert torch.equal(blob[:150 * 192], w["embedding.weight"].reshape(-1))
    o = 150 * 192
    c0 = w["prenet.convolutions.0.weight"]                                  # [192, 192, 5] -> [out][tap][cin]
    assert torch.equal(blob[o:o + 192 * 5 * 192].view(192, 5, 192)[7, 3], c0[7, :, 3])
    assert blob.numel() % 4 == 0


def test_state_dict_prefix_and_errors():
    w = oenc.make_weights(1, "init", 178)
    m = te.TextEncoder(n_vocab=178)
    m.load_state_dict({"encoder." + k: v for k, v in w.items()})            # as stored in a MatchaTTS checkpoint
    assert list(m.state_dict().keys()) == list(w.keys())
    assert torch.equal(m.state_dict()["encoder.attention_layers.2.key_conv.weight"], w["encoder.attention_layers.2.key_conv.weight"])
    with pytest.raises(KeyError):
        te.TextEncoder(n_vocab=178).load_state_dict({k: v for k, v in w.items() if k != "mean_projection.bias"})
    ep, dp = te.default_params()
    ep.n_layers = 4
    with pytest.raises(NotImplementedError):
        te.TextEncoder("RoPE Encoder", ep, dp, 178)


def test_rope_tables_match_oracle():
    cos, sin = te.rope_tables(77)
    ocos, osin = oenc.rope_tables(77)
    assert cos.shape == (77, 48) and torch.equal(cos, ocos) and torch.equal(sin, osin)
    assert torch.equal(cos[:, :24], cos[:, 24:])


def test_initialize_weights_distribution():
    m = te.TextEncoder(n_vocab=178).initialize_weights(seed=5)
    sd = m.state_dict()
    assert sd["prenet.projection.weight"].abs().max().item() == 0            # text_encoder.py:44-46
    assert abs(sd["embedding.weight"].std().item() / 192 ** -0.5 - 1) < 0.03
    q = sd["encoder.attention_layers.0.query_conv.weight"]
    assert q.abs().max().item() <= (6.0 / 384) ** 0.5                        # xavier-uniform
    again = te.TextEncoder(n_vocab=178).initialize_weights(seed=5).state_dict()
    assert all(torch.equal(sd[k], again[k]) for k in sd)


def test_no_cpu_fallback():
    m = te.TextEncoder(n_vocab=178).initialize_weights(0)
    with pytest.raises(RuntimeError):
        m(torch.zeros(1, 8, dtype=torch.long), torch.tensor([8]))


def test_matcha_tts_state_dict_round_trip_and_errors():
    """The MatchaTTS mirror ingests a checkpoint-shaped state_dict (encoder.* and decoder.estimator.*) and hands the same keys back."""
    from matcha_b200.matcha_tts import MatchaTTS
    from oracle import decoder as odec
    we, wd = oenc.make_weights(2, "init", 178), odec.make_weights(3, "init")
    sd = {"encoder." + k: v for k, v in we.items()}
    sd.update({"decoder.estimator." + k: v for k, v in wd.items()})
    m = MatchaTTS(178, precision="fp16").load_state_dict(sd)
    out = m.state_dict()
    assert set(out.keys()) == set(sd.keys())
    assert torch.equal(out["encoder.encoder.ffn_layers.5.conv_net.3.weight"], sd["encoder.encoder.ffn_layers.5.conv_net.3.weight"])
    assert torch.equal(out["decoder.estimator.final_proj.bias"], sd["decoder.estimator.final_proj.bias"])
    assert m.encoder.precision == "fp16" and MatchaTTS(178, precision="fp32").encoder.precision == "fp32"
    with pytest.raises(NotImplementedError):
        MatchaTTS(178, out_channels=100)
    with pytest.raises(RuntimeError):                                   # no CPU fallback anywhere in the flow
        m.synthesise(torch.zeros(1, 8, dtype=torch.long), torch.tensor([8]), n_timesteps=2)
