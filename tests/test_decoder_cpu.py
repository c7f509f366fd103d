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
