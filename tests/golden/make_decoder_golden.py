"""Generates tests/golden/decoder_golden.npz from the reference's OWN decoder modules.

Run in the build container:   python tests/golden/make_decoder_golden.py
Imports /root/reference/matcha/models/components/{decoder,transformer,flow_matching}.py
UNMODIFIED, with the diffusers stand-in of oracle/_shims first on sys.path (diffusers is a
third-party dependency the container lacks).  Weights come from
oracle.decoder.make_weights(seed, "rich") loaded through load_state_dict, so the tests can
regenerate them without the reference tree.  Everything runs in fp32 on the CPU, eval mode.
"""
import os
import sys
from types import SimpleNamespace

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, "/root/reference")
sys.path.insert(0, os.path.join(ROOT, "oracle", "_shims"))

from oracle import decoder as odec  # noqa: E402

WEIGHT_SEED = 11


def reference_cfm():
    from matcha.models.components import transformer
    from matcha.models.components.flow_matching import ConditionalFlowMatching
    assert transformer.DIFFUSERS_AVAILABLE
    cfm = ConditionalFlowMatching(
        in_channels=160, out_channel=80, cfm_params=SimpleNamespace(solver="euler", sigma_min=1e-4),
        decoder_params=dict(channels=(256, 256), dropout=0.05, attention_head_dim=64, n_blocks=1,
                            num_mid_blocks=2, num_heads=4)).eval()
    w = odec.make_weights(WEIGHT_SEED, "rich")
    missing = cfm.estimator.load_state_dict(w, strict=True)
    assert not missing.missing_keys and not missing.unexpected_keys
    return cfm


def mask_from_lengths(lengths, T):
    m = torch.zeros(len(lengths), 1, T)
    for i, n in enumerate(lengths):
        m[i, :, :n] = 1
    return m


def main():
    torch.set_num_threads(4)
    cfm = reference_cfm()
    g = torch.Generator().manual_seed(5)
    blob = {"weight_seed": np.array([WEIGHT_SEED])}

    def put(name, **kw):
        for k, v in kw.items():
            blob[f"{name}/{k}"] = v.detach().numpy() if isinstance(v, torch.Tensor) else np.asarray(v)

    with torch.inference_mode():
        # A: Euler solve, even T, ragged mask, 2 steps (solve_ode_euler, flow_matching.py:67)
        B, T = 2, 32
        mu, z = torch.randn(B, 80, T, generator=g), torch.randn(B, 80, T, generator=g)
        mask = mask_from_lengths([32, 21], T)
        t_span = torch.linspace(0, 1, 3)
        out = cfm.solve_ode_euler(z, t_span, mu, mask, None, None)
        put("solve_T32_n2", mu=mu, z=z, mask=mask, t_span=t_span, out=out)
        # B: one estimator call, odd T (nearest-interp path, decoder.py:338), per-sample t
        B, T = 2, 37
        mu, x = torch.randn(B, 80, T, generator=g), torch.randn(B, 80, T, generator=g)
        mask = mask_from_lengths([37, 30], T)
        t = torch.tensor([0.25, 0.8125])
        put("eval_T37_tvec", mu=mu, x=x, mask=mask, t=t, out=cfm.estimator(x, mask, mu, t))
        # C: one estimator call, scalar t, full mask, T multiple of 4 but not of 32
        B, T = 3, 44
        mu, x = torch.randn(B, 80, T, generator=g), torch.randn(B, 80, T, generator=g)
        mask = mask_from_lengths([44, 44, 44], T)
        t = torch.tensor(0.5)
        put("eval_T44_full", mu=mu, x=x, mask=mask, t=t, out=cfm.estimator(x, mask, mu, t))
        # D: 4-step solve with a shorter utterance whose T/2 mask becomes all ones
        B, T = 2, 40
        mu, z = torch.randn(B, 80, T, generator=g), torch.randn(B, 80, T, generator=g)
        mask = mask_from_lengths([40, 19], T)
        t_span = torch.linspace(0, 1, 5)
        put("solve_T40_n4", mu=mu, z=z, mask=mask, t_span=t_span, out=cfm.solve_ode_euler(z, t_span, mu, mask, None, None))
    # E: compute_loss (flow_matching.py:106-151): replay its RNG draws (rand for t, then randn_like)
    B, T = 2, 32
    x1, mu = torch.randn(B, 80, T, generator=g), torch.randn(B, 80, T, generator=g)
    mask = mask_from_lengths([32, 27], T)
    torch.manual_seed(99)
    with torch.no_grad():
        loss, phi = cfm.compute_loss(x1, mask, mu)
    torch.manual_seed(99)
    t = torch.rand([B, 1, 1]); z = torch.randn_like(x1)
    put("loss_T32", x1=x1, mu=mu, mask=mask, t=t, z=z, loss=loss, phi_t=phi)
    np.savez_compressed(os.path.join(os.path.dirname(__file__), "decoder_golden.npz"), **blob)
    for k, v in blob.items():
        print(k, v.shape)


if __name__ == "__main__":
    main()
