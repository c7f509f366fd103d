"""CPU tests of the decoder's host logic: state_dict ingest, weight packing, the Euler time grid, and that
nothing silently falls back to the CPU."""
import pytest
import torch

from matcha_b200 import decoder as dec
from matcha_b200.flow_matching import CFM
from oracle import decoder as odec


def test_expected_keys_are_the_reference_state_dict():
    assert dict(dec.expected_keys()) == dict(odec.param_shapes())


def test_pack_weights_layout_and_size():
    w = odec.make_weights(3, "rich")
    blob = dec.pack_weights(w)
    assert blob.dtype == torch.float32 and blob.numel() == sum(v.numel() for v in w.values()) + 80
    # frequency table first (decoder.py:26-27), then time_mlp.linear_1 row-major
    half = 80
    freq = torch.exp(torch.arange(half, dtype=torch.float32) * -(torch.log(torch.tensor(10000.0)) / (half - 1)))
    assert torch.equal(blob[:80], freq)
    assert torch.equal(blob[80:80 + 1024 * 160], w["time_mlp.linear_1.weight"].reshape(-1))


def test_state_dict_prefixes_and_errors():
    w = odec.make_weights(1, "init")
    d = dec.Decoder()
    d.load_state_dict({"decoder.estimator." + k: v for k, v in w.items()})
    assert set(d.state_dict().keys()) == set(w.keys())
    with pytest.raises(KeyError):
        dec.Decoder().load_state_dict({k: v for k, v in w.items() if k != "final_proj.bias"})
    with pytest.raises(NotImplementedError):
        dec.Decoder(channels=(128, 128))


def test_euler_time_grid_matches_oracle_loop():
    for n in (1, 2, 4, 10, 50):
        t_span = torch.linspace(0, 1, n + 1)
        ts, dts = dec.euler_time_grid(t_span)
        ots, odts = odec.euler_time_grid(t_span)
        assert torch.equal(torch.stack(ts), ots) and torch.equal(torch.stack(dts), odts)


def test_no_cpu_fallback():
    m = CFM(160, 80, type("P", (), {"solver": "euler", "sigma_min": 1e-4})(),
            dict(channels=(256, 256), dropout=0.05, attention_head_dim=64, n_blocks=1, num_mid_blocks=2, num_heads=4))
    m.load_state_dict(odec.make_weights(0, "init"))
    x = torch.zeros(1, 80, 8)
    with pytest.raises(RuntimeError):
        m.estimator(x, torch.ones(1, 1, 8), x, torch.tensor(0.5))


def test_initialize_weights_distribution():
    """C14: Kaiming-normal conv/linear weights, zero biases, unit norms; ConvTranspose keeps torch's default."""
    from matcha_b200.decoder import Decoder, expected_keys
    d = Decoder().initialize_weights(seed=3)
    sd = d.state_dict()
    assert list(sd.keys()) == list(expected_keys().keys())
    w = sd["Mid_Blocks.0.0.block1.block.0.weight"]                      # [256, 256, 3], fan_in 768
    assert abs(w.std().item() / (2.0 / 768) ** 0.5 - 1) < 0.02 and abs(w.mean().item()) < 1e-3
    assert sd["Mid_Blocks.0.0.block1.block.0.bias"].abs().max().item() == 0
    assert torch.equal(sd["final_norm.weight"], torch.ones(256)) and torch.equal(sd["Mid_Blocks.0.1.0.norm1.bias"], torch.zeros(256))
    ct = sd["Upsampling_Blocks.0.2.conv.weight"]                        # ConvTranspose1d [256, 256, 4]: U(+-1/32)
    assert ct.abs().max().item() <= 1 / 32 and ct.abs().max().item() > 0.99 / 32
    again = Decoder().initialize_weights(seed=3).state_dict()
    assert all(torch.equal(sd[k], again[k]) for k in sd)
