"""Condition producer without a GPU: the oracle restatement and the module's torch-op branch against
the golden outputs of the unmodified reference's FastSpeech2Acoustic (tests/golden/fs2_*.npz, oracle/make_golden.py: gen_fs2),
and the checkpoint ABI (same state_dict keys in the same order, strict loads)."""
import pytest
import torch

import diffsinger_b200 as ds
from conftest import load_fs2_golden, rel_l2
from oracle.fastspeech_oracle import fs2_condition


def _oracle(d, sd, hp):
    variances = [v for v in ("energy", "breathiness", "voicing", "tension") if hp.get(f"use_{v}_embed")]
    spk = sd["spk_embed.weight"][d["spk_embed_id"]][:, None, :] if hp["use_spk_id"] else None
    return fs2_condition(sd, d["tokens"], d["mel2ph"], d["f0"], num_heads=hp["num_heads"], variances=variances,
                         key_shift=d.get("key_shift"), speed=d.get("speed"), spk_embed=spk, **{v: d[v] for v in variances})


@pytest.mark.parametrize("tag", ["a", "b"])
def test_oracle_matches_reference_golden(tag):
    d, y, sd, hp = load_fs2_golden(tag)
    assert rel_l2(_oracle(d, sd, hp), y) < 2e-6


@pytest.mark.parametrize("tag", ["a", "b"])
def test_module_torch_branch_and_checkpoint_abi(tag):
    d, y, sd, hp = load_fs2_golden(tag)
    ds.hparams.clear()
    ds.hparams.update(hp)
    m = ds.FastSpeech2Acoustic(sd["txt_embed.weight"].shape[0]).eval()
    assert list(m.state_dict().keys()) == list(sd.keys())
    assert all(m.state_dict()[k].shape == v.shape and m.state_dict()[k].dtype == v.dtype for k, v in sd.items())
    m.load_state_dict(sd, strict=True)
    kw = {k: v for k, v in d.items() if k not in ("tokens", "mel2ph", "f0")}
    with torch.no_grad():
        out = m(d["tokens"], d["mel2ph"], d["f0"], **kw)
    assert out.shape == y.shape and rel_l2(out, y) < 2e-6
    # gradients flow through the torch branch (training path)
    m.train()
    out = m(d["tokens"], d["mel2ph"], d["f0"], **kw)
    out.square().mean().backward()
    assert m.encoder.layers[0].op.self_attn.in_proj.weight.grad.abs().sum() > 0 and m.pitch_embed.weight.grad.abs().sum() > 0


def test_unsupported_positional_embeddings_are_rejected():
    ds.hparams.clear()
    ds.hparams.update(hidden_size=64, enc_layers=1, num_heads=2, enc_ffn_kernel_size=3, ffn_act="gelu", use_rope=False, rel_pos=True,
                      use_pos_embed=True, use_spk_id=False)
    with pytest.raises(NotImplementedError, match="rotary"):
        ds.FastSpeech2Acoustic(5)
