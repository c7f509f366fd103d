"""Generates tests/golden/alignment_golden.npz from the reference's OWN MatchaTTS.forward / synthesise.

Run in the build container:   python tests/golden/make_alignment_golden.py
/root/reference/matcha/models/matcha_tts.py is imported UNMODIFIED.  Three third-party packages the
container lacks are replaced by import stubs: pytorch_lightning (LightningModule
= nn.Module with no-op save_hyperparameters/log), matplotlib (never called on this path) and diffusers
(oracle/_shims).  ``matcha.utils.monotonic_align.core`` resolves to the reference's core.pyx compiled
into oracle/_ref (the reference tree itself is read-only, so its own in-place build cannot be used).
The tensors at the seams of the training step (matcha_tts.py:267-322) and of synthesis (:216-229) are
recorded by wrapping -- not replacing -- the functions the reference calls there.
"""
import glob
import importlib.util
import os
import sys
import types

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, "/root/reference")
sys.path.insert(0, os.path.join(ROOT, "oracle", "_shims"))


def install_stubs():
    pl = types.ModuleType("pytorch_lightning")

    class LightningModule(torch.nn.Module):
        def save_hyperparameters(self, *a, **k):
            pass

        def log(self, *a, **k):
            pass

        def log_dict(self, *a, **k):
            pass

    pl.LightningModule = LightningModule
    pl.LightningDataModule = object
    plu = types.ModuleType("pytorch_lightning.utilities")
    plu.grad_norm = lambda *a, **k: {}
    pl.utilities = plu
    sys.modules["pytorch_lightning"] = pl
    sys.modules["pytorch_lightning.utilities"] = plu
    for name in ("matplotlib", "matplotlib.pyplot"):
        sys.modules[name] = types.ModuleType(name)
    sys.modules["matplotlib"].pyplot = sys.modules["matplotlib.pyplot"]
    sys.modules["matplotlib"].use = lambda *a, **k: None
    hits = glob.glob(os.path.join(ROOT, "oracle", "_ref", "core*.so"))
    assert hits, "run `make -C oracle` first (compiles the reference core.pyx into oracle/_ref)"
    spec = importlib.util.spec_from_file_location("matcha.utils.monotonic_align.core", hits[0])
    core = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(core)
    sys.modules["matcha.utils.monotonic_align.core"] = core


def main():
    install_stubs()
    import matcha.models.matcha_tts as mt
    import matcha.utils.monotonic_align as ma
    assert ma.maximum_path_c.__module__ != ma.__name__, "compiled core not picked up"
    torch.manual_seed(3)
    torch.set_num_threads(4)
    model = mt.MatchaTTS(n_vocab=150, out_channels=80, hidden_channels=192).eval()
    blob = {}

    def put(name, **kw):
        for k, v in kw.items():
            blob[f"{name}/{k}"] = v.detach().cpu().numpy() if isinstance(v, torch.Tensor) else np.asarray(v)

    rec = {}
    orig_mp, orig_dl, orig_gp = ma.maximum_path, mt.duration_loss, mt.generate_path
    orig_enc, orig_closs, orig_dec_fwd = model.encoder.forward, model.decoder.compute_loss, model.decoder.forward

    def mp(value, mask):
        rec["log_prior"], rec["attn_mask"] = value.clone(), mask.clone()
        return orig_mp(value, mask)

    def dl(logw, logw_, lengths):
        rec["logw"], rec["logw_"] = logw.clone(), logw_.clone()
        return orig_dl(logw, logw_, lengths)

    def gp(duration, mask):
        rec["gp_duration"], rec["gp_mask"] = duration.clone(), mask.clone()
        out = orig_gp(duration, mask)
        rec["gp_path"] = out.clone()
        return out

    def enc(x, x_lengths, *a, **k):
        out = orig_enc(x, x_lengths, *a, **k)
        rec["mu_x"], rec["enc_logw"], rec["x_mask"] = (o.clone() for o in out)
        return out

    def closs(x1, mask, mu, spks=None, cond=None):
        rec["closs_x1"], rec["closs_mask"], rec["mu_y"] = x1.clone(), mask.clone(), mu.clone()
        return torch.zeros(()), None          # the decoder loss itself is pinned by decoder_golden.npz

    def dec_fwd(mu, mask, n_timesteps, temperature=1.0, spks=None, cond=None):
        rec["dec_mu"], rec["dec_mask"] = mu.clone(), mask.clone()
        return torch.zeros_like(mu)

    ma.maximum_path = mp
    mt.duration_loss, mt.generate_path = dl, gp
    model.encoder.forward, model.decoder.compute_loss, model.decoder.forward = enc, closs, dec_fwd

    g = torch.Generator().manual_seed(17)
    # ---- training step (matcha_tts.py:267-322), two shapes incl. one where t_x == T_x / t_y == T_y
    for name, (B, T_x, T_y) in {"train_a": (3, 24, 96), "train_b": (2, 37, 130)}.items():
        x = torch.randint(3, 150, (B, T_x), generator=g)
        x_lengths = torch.tensor([T_x] + [int(T_x * f) for f in (0.7, 0.85)][: B - 1])
        y = torch.randn(B, 80, T_y, generator=g)
        y_lengths = torch.tensor([T_y] + [int(T_y * f) for f in (0.75, 0.9)][: B - 1])
        with torch.no_grad():
            dur_loss, prior_loss, _, attn = model(x, x_lengths, y, y_lengths)
        put(name, x_lengths=x_lengths, y=y, y_lengths=y_lengths, mu_x=rec["mu_x"], logw=rec["logw"],
            x_mask=rec["x_mask"], log_prior=rec["log_prior"], attn_mask=rec["attn_mask"], attn=attn,
            logw_=rec["logw_"], mu_y=rec["mu_y"], prior_loss=prior_loss, dur_loss=dur_loss)
    # ---- synthesis pre-step (matcha_tts.py:216-229), integer and fractional length_scale
    for name, (B, T_x, ls) in {"synth_a": (3, 21, 1.0), "synth_b": (2, 30, 1.3)}.items():
        x = torch.randint(3, 150, (B, T_x), generator=g)
        x_lengths = torch.tensor([T_x] + [int(T_x * f) for f in (0.6, 0.8)][: B - 1])
        with torch.inference_mode():
            out = model.synthesise(x, x_lengths, n_timesteps=1, length_scale=ls)
        put(name, x_lengths=x_lengths, length_scale=ls, mu_x=rec["mu_x"], logw=rec["enc_logw"], x_mask=rec["x_mask"],
            gp_duration=rec["gp_duration"], gp_mask=rec["gp_mask"], gp_path=rec["gp_path"], mu_y=rec["dec_mu"],
            y_mask=rec["dec_mask"], attn=out["attn"], mel_lengths=out["mel_lengths"])
    out_path = os.path.join(ROOT, "tests", "golden", "alignment_golden.npz")
    np.savez_compressed(out_path, **blob)
    print(out_path, os.path.getsize(out_path), "bytes;", len(blob), "arrays")


if __name__ == "__main__":
    main()
