"""TEST INFRASTRUCTURE ONLY -- CPU restatement (torch fp32, functional, from a plain state_dict) of the reference's condition
producer: FastSpeech2Acoustic.forward (modules/fastspeech/acoustic_encoder.py:82-118) over FastSpeech2Encoder
(modules/fastspeech/tts_modules.py:353-440) with rotary self-attention (modules/commons/common_layers.py:152-258,
rotary_embedding_torch.py:35-39,119,174-188), eval mode (no dropout).  Pinned against the imported reference by
tests/test_oracle_golden.py (fixtures tests/golden/fs2_*.npz).  Never imported by the product."""
from __future__ import annotations

import math

import torch
import torch.nn.functional as F


def _rope(t, freqs):
    L = t.shape[-2]
    ang = torch.arange(L, dtype=torch.float32)[:, None] * freqs[None, :]
    ang = torch.stack((ang, ang), dim=-1).flatten(-2)  # f0 f0 f1 f1 ...
    pairs = t.reshape(*t.shape[:-1], -1, 2)
    rot = torch.stack((-pairs[..., 1], pairs[..., 0]), dim=-1).reshape(t.shape)
    return t * ang.cos() + rot * ang.sin()


@torch.no_grad()
def fs2_condition(sd, txt_tokens, mel2ph, f0, *, num_heads, variances=(), key_shift=None, speed=None, spk_embed=None, languages=None,
                  **curves):
    """sd: FastSpeech2Acoustic.state_dict(); variances: names in embedding order; curves: name -> [B, T]; spk_embed [B, 1 or T, H]."""
    H = sd["txt_embed.weight"].shape[1]
    B, Tt = txt_tokens.shape
    pad = txt_tokens == 0
    keep = (~pad).float()[..., None]
    dur = mel2ph.new_zeros(B, Tt + 1).scatter_add(1, mel2ph, torch.ones_like(mel2ph))[:, 1:].float()
    lin1 = lambda name, v: v[:, :, None] * sd[name + ".weight"][:, 0] + sd[name + ".bias"]
    x = math.sqrt(H) * sd["txt_embed.weight"][txt_tokens] + lin1("dur_embed", dur)
    if languages is not None:
        x = x + sd["lang_embed.weight"][languages]
    x = x * keep
    n_layers = 1 + max(int(k.split(".")[2]) for k in sd if k.startswith("encoder.layers."))
    D = H // num_heads
    for i in range(n_layers):
        p = f"encoder.layers.{i}.op."
        h = F.layer_norm(x, (H,), sd[p + "layer_norm1.weight"], sd[p + "layer_norm1.bias"])
        q, k, v = ((h @ sd[p + "self_attn.in_proj.weight"].T).split(H, dim=-1))
        q, k, v = (t.view(B, Tt, num_heads, D).transpose(1, 2) for t in (q, k, v))
        fr = sd[p + "self_attn.rotary_embed.freqs"]
        s = (_rope(q, fr) @ _rope(k, fr).transpose(-2, -1)) / math.sqrt(D)
        s = s.masked_fill(pad[:, None, None, :], float("-inf"))
        o = (F.softmax(s, dim=-1) @ v).transpose(1, 2).reshape(B, Tt, H) @ sd[p + "self_attn.out_proj.weight"].T
        x = (x + o) * keep
        h = F.layer_norm(x, (H,), sd[p + "layer_norm2.weight"], sd[p + "layer_norm2.bias"])
        w1 = sd[p + "ffn.ffn_1.weight"]
        ks = w1.shape[2]
        h = F.conv1d(h.transpose(1, 2), w1, sd[p + "ffn.ffn_1.bias"], padding=ks // 2).transpose(1, 2) * ks ** -0.5
        h = F.gelu(h) @ sd[p + "ffn.ffn_2.weight"].T + sd[p + "ffn.ffn_2.bias"]
        x = (x + h) * keep
    x = F.layer_norm(x, (H,), sd["encoder.layer_norm.weight"], sd["encoder.layer_norm.bias"]) * keep
    enc = F.pad(x, [0, 0, 1, 0])
    cond = torch.gather(enc, 1, mel2ph[..., None].expand(-1, -1, H))
    if spk_embed is not None:
        cond = cond + spk_embed
    cond = cond + lin1("pitch_embed", (1 + f0 / 700).log())
    for name in variances:
        cond = cond + lin1(f"variance_embeds.{name}", curves[name])
    if key_shift is not None:
        cond = cond + lin1("key_shift_embed", key_shift)
    if speed is not None:
        cond = cond + lin1("speed_embed", speed)
    return cond
