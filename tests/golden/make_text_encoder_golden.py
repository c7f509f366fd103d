"""Generates tests/golden/text_encoder_golden.npz from the reference's OWN TextEncoder.

Run in the build container:   python tests/golden/make_text_encoder_golden.py
Imports /root/reference/matcha/models/components/text_encoder.py UNMODIFIED (it needs torch and einops
only).  The module is built with the configuration MatchaTTS uses (matcha_tts.py:122-139); its weights come from
oracle.text_encoder.make_weights(seed, "rich") through load_state_dict(strict=True), so the tests can
regenerate them without the reference tree.  fp32, CPU, eval mode (dropout off).  Only ids, lengths and the
reference's three outputs are stored.
"""
import os
import sys
from types import SimpleNamespace

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, "/root/reference")

from oracle import text_encoder as oenc  # noqa: E402

WEIGHT_SEED, N_VOCAB = 23, 178


def reference_encoder():
    from matcha.models.components.text_encoder import TextEncoder
    enc = TextEncoder(
        "RoPE Encoder",
        SimpleNamespace(n_feats=80, n_channels=192, filter_channels=768, n_heads=2, n_layers=6, kernel_size=3,
                        p_dropout=0.1, prenet=True),
        SimpleNamespace(filter_channels_dp=256, kernel_size=3, p_dropout=0.1), N_VOCAB).eval()
    res = enc.load_state_dict(oenc.make_weights(WEIGHT_SEED, "rich", N_VOCAB), strict=True)
    assert not res.missing_keys and not res.unexpected_keys
    assert sum(p.numel() for p in enc.parameters()) == sum(int(np.prod(s)) for s in oenc.param_shapes(N_VOCAB).values())
    return enc


def main():
    torch.set_num_threads(4)
    enc = reference_encoder()
    g = torch.Generator().manual_seed(31)
    blob = {"weight_seed": np.array([WEIGHT_SEED]), "n_vocab": np.array([N_VOCAB])}
    cases = {"ragged_T50": (3, 50, [50, 37, 21]), "full_T33": (2, 33, [33, 33]), "one_token": (2, 9, [9, 1]),
             "long_T160": (2, 160, [160, 113])}
    with torch.inference_mode():
        for name, (B, T, lens) in cases.items():
            ids = torch.randint(0, N_VOCAB, (B, T), generator=g)
            lengths = torch.tensor(lens)
            mu, logw, mask = enc(ids, lengths)
            blob[f"{name}/ids"] = ids.numpy()
            blob[f"{name}/lengths"] = lengths.numpy()
            blob[f"{name}/mu"] = mu.numpy()
            blob[f"{name}/logw"] = logw.numpy()
            blob[f"{name}/mask"] = mask.numpy()
            print(name, tuple(mu.shape), float(mu.abs().mean()), float(logw.abs().mean()))
    out = os.path.join(os.path.dirname(__file__), "text_encoder_golden.npz")
    np.savez_compressed(out, **blob)
    print(out, os.path.getsize(out), "bytes")


if __name__ == "__main__":
    main()
