"""Condition producer in front of the sampler: drop-in for the reference's ``FastSpeech2Acoustic``
(modules/fastspeech/acoustic_encoder.py:14-118) -- phoneme embedding + FastSpeech2 transformer encoder
(modules/fastspeech/tts_modules.py:353-440, commons/common_layers.py:120-258) + the ``torch.gather`` length regulation and the
per-frame pitch / variance / key-shift / speed / speaker embeddings (acoustic_encoder.py:97-116).

Same constructor (``vocab_size``; everything else from ``hparams``), same parameter names and shapes -- ``load_state_dict(strict=True)``
both ways, including the shared ``rotary_embed.freqs`` entries -- and the same ``forward(txt_tokens, mel2ph, f0, ...)`` -> condition
``[B, T, H]``.  Calls that need no gradient on a CUDA module run on libdsb200.so: LayerNorm / Linear / Conv1d through ``ds_layernorm`` /
``ds_gemm`` (GELU and residual epilogues), attention + RoPE through ``ds_attention``, the gather and every per-frame ``Linear(1, H)``
through one ``ds_length_regulate`` pass; nothing of it synchronises, so the whole ``.ds`` tensors -> condition -> sampler call can be
captured in one CUDA graph.  Calls under autograd (training) and CPU modules take the differentiable torch-op branch.

Natively supported: the configuration of configs/acoustic.yaml (``use_rope: true``, ``ffn_act: gelu`` / ``relu``).  The older positional
embeddings (``rel_pos`` / sinusoidal with nn.MultiheadAttention) are not built here: construct the reference's encoder for those.
"""
from __future__ import annotations

import math

import torch
import torch.nn as nn
import torch.nn.functional as F

from .hparams import get_hparams

PAD_INDEX = 0  # utils/phoneme_utils.py


def _linear(i, o, bias=True):
    m = nn.Linear(i, o, bias=bias)
    nn.init.xavier_uniform_(m.weight)  # XavierUniformInitLinear, common_layers.py:23-31
    if bias:
        nn.init.constant_(m.bias, 0.0)
    return m


def _embedding(n, dim, padding_idx=None):
    m = nn.Embedding(n, dim, padding_idx=padding_idx)  # NormalInitEmbedding, common_layers.py:12-20
    nn.init.normal_(m.weight, mean=0, std=dim ** -0.5)
    if padding_idx is not None:
        nn.init.constant_(m.weight[padding_idx], 0)
    return m


class RotaryEmbedding(nn.Module):
    """The one thing of rotary_embedding_torch.RotaryEmbedding a checkpoint holds: ``freqs`` (:119,131)."""

    def __init__(self, dim, theta=10000.0):
        super().__init__()
        self.theta = float(theta)
        self.freqs = nn.Parameter(1.0 / (theta ** (torch.arange(0, dim, 2)[:(dim // 2)].float() / dim)), requires_grad=False)

    def rotate(self, t):
        """t [B, heads, L, D] -> rotated (interleaved pairs; rotary_embedding_torch.py:35-39,174-188)."""
        L = t.shape[-2]
        ang = torch.arange(L, device=t.device, dtype=torch.float32)[:, None] * self.freqs.float()[None, :]   # [L, D/2]
        cos, sin = ang.cos().repeat_interleave(2, -1), ang.sin().repeat_interleave(2, -1)
        x1, x2 = t[..., 0::2], t[..., 1::2]
        rot = torch.stack((-x2, x1), dim=-1).flatten(-2)
        return t * cos + rot * sin


class SelfAttention(nn.Module):
    def __init__(self, dim, heads, rotary_embed):
        super().__init__()
        self.embed_dim, self.num_heads, self.head_dim = dim, heads, dim // heads
        self.in_proj = nn.Linear(dim, dim * 3, bias=False)
        self.out_proj = nn.Linear(dim, dim, bias=False)
        self.rotary_embed = rotary_embed

    def forward(self, x, key_padding_mask=None):  # common_layers.py:175-213 (attention dropout is 0 in the encoder)
        B, L, _ = x.shape
        q, k, v = (t.view(B, L, self.num_heads, self.head_dim).transpose(1, 2) for t in self.in_proj(x).split(self.embed_dim, dim=-1))
        q, k = self.rotary_embed.rotate(q), self.rotary_embed.rotate(k)
        s = torch.matmul(q, k.transpose(-2, -1)) / math.sqrt(self.head_dim)
        if key_padding_mask is not None:
            s = s.masked_fill(key_padding_mask[:, None, None, :], float('-inf'))
        o = torch.matmul(F.softmax(s, dim=-1), v).transpose(1, 2).reshape(B, L, self.embed_dim)
        return self.out_proj(o)


class FFN(nn.Module):
    def __init__(self, dim, filt, kernel_size, act):
        super().__init__()
        if act not in ('gelu', 'relu'):
            raise NotImplementedError(f"diffsinger_b200.FastSpeech2Acoustic: ffn_act={act!r} (gelu / relu are built)")
        self.kernel_size, self.act = kernel_size, act
        self.ffn_1 = nn.Conv1d(dim, filt, kernel_size, padding=kernel_size // 2)
        self.ffn_2 = _linear(filt, dim)

    def forward(self, x):  # common_layers.py:141-150
        x = self.ffn_1(x.transpose(1, 2)).transpose(1, 2) * self.kernel_size ** -0.5
        return self.ffn_2(F.gelu(x) if self.act == 'gelu' else F.relu(x))


class EncSALayer(nn.Module):
    def __init__(self, dim, heads, kernel_size, act, rotary_embed):
        super().__init__()
        self.layer_norm1 = nn.LayerNorm(dim)
        self.self_attn = SelfAttention(dim, heads, rotary_embed)
        self.layer_norm2 = nn.LayerNorm(dim)
        self.ffn = FFN(dim, 4 * dim, kernel_size, act)

    def forward(self, x, pad):  # common_layers.py:236-254 (dropout: eval / p applied by the caller's training loop only)
        keep = (~pad).to(x.dtype)[..., None]
        x = (x + self.self_attn(self.layer_norm1(x), key_padding_mask=pad)) * keep
        return (x + self.ffn(self.layer_norm2(x))) * keep


class TransformerEncoderLayer(nn.Module):
    def __init__(self, dim, heads, kernel_size, act, rotary_embed):
        super().__init__()
        self.op = EncSALayer(dim, heads, kernel_size, act, rotary_embed)


class FastSpeech2Encoder(nn.Module):
    def __init__(self, hidden_size, num_layers, ffn_kernel_size, ffn_act, num_heads):
        super().__init__()
        rotary = RotaryEmbedding(hidden_size // num_heads)  # one module shared by all layers (tts_modules.py:364-375)
        self.layers = nn.ModuleList([TransformerEncoderLayer(hidden_size, num_heads, ffn_kernel_size, ffn_act, rotary)
                                     for _ in range(num_layers)])
        self.layer_norm = nn.LayerNorm(hidden_size)
        self.embed_scale = math.sqrt(hidden_size)
        self.num_heads = num_heads

    def forward(self, main_embed, extra_embed, pad):  # tts_modules.py:392-430 (RoPE: no additive position embedding)
        keep = (~pad).to(main_embed.dtype)[..., None]
        x = (self.embed_scale * main_embed + extra_embed) * keep
        for layer in self.layers:
            x = layer.op(x, pad) * keep
        return self.layer_norm(x) * keep


class FastSpeech2Acoustic(nn.Module):
    """Same constructor / parameters / forward as the reference's class of this name (acoustic_encoder.py:14-118)."""

    VARIANCES = ('energy', 'breathiness', 'voicing', 'tension')  # acoustic_encoder.py:32-43: embedding order

    def __init__(self, vocab_size):
        super().__init__()
        hp = get_hparams()
        H = hp['hidden_size']
        if not (hp.get('use_rope', False) and hp.get('use_pos_embed', True)):
            raise NotImplementedError("diffsinger_b200.FastSpeech2Acoustic builds the rotary-embedding encoder (use_rope: true, "
                                      "configs/acoustic.yaml); keep the reference's encoder for rel_pos / sinusoidal configurations")
        self.hidden_size = H
        self.txt_embed = _embedding(vocab_size, H, PAD_INDEX)
        self.use_lang_id = hp.get('use_lang_id', False)
        if self.use_lang_id:
            self.lang_embed = _embedding(hp['num_lang'] + 1, H, padding_idx=0)
        self.dur_embed = _linear(1, H)
        self.encoder = FastSpeech2Encoder(H, hp['enc_layers'], hp['enc_ffn_kernel_size'], hp['ffn_act'], hp['num_heads'])
        self.pitch_embed = _linear(1, H)
        self.variance_embed_list = [v for v in self.VARIANCES if hp.get(f'use_{v}_embed', False)]
        self.use_variance_embeds = len(self.variance_embed_list) > 0
        if self.use_variance_embeds:
            self.variance_embeds = nn.ModuleDict({v: _linear(1, H) for v in self.variance_embed_list})
        self.use_key_shift_embed = hp.get('use_key_shift_embed', False)
        if self.use_key_shift_embed:
            self.key_shift_embed = _linear(1, H)
        self.use_speed_embed = hp.get('use_speed_embed', False)
        if self.use_speed_embed:
            self.speed_embed = _linear(1, H)
        self.use_spk_id = hp['use_spk_id']
        if self.use_spk_id:
            self.spk_embed = _embedding(hp['num_spk'], H)
        self._packed = None
        self._packed_key = None
        self.register_load_state_dict_post_hook(lambda module, incompatible: module.invalidate_native())

    def invalidate_native(self):
        self._packed = None

    # ------------------------------------------------------------------ shared front: token-level inputs of the encoder
    def _token_inputs(self, txt_tokens, mel2ph, languages):
        txt_embed = self.txt_embed(txt_tokens)
        B, Tt = txt_tokens.shape
        dur = mel2ph.new_zeros(B, Tt + 1).scatter_add(1, mel2ph, torch.ones_like(mel2ph))[:, 1:].float()  # mel2ph_to_dur
        extra = self.dur_embed(dur[:, :, None])
        if self.use_lang_id:
            extra = extra + self.lang_embed(languages)
        return txt_embed, extra, txt_tokens == PAD_INDEX

    def _curves(self, f0, key_shift, speed, variances):
        """(curve [B, T or 1] fp32, Linear(1, H)) per per-frame embedding, in the reference's order."""
        out = [((1 + f0 / 700).log(), self.pitch_embed)]
        for v in self.variance_embed_list:
            out.append((variances[v], self.variance_embeds[v]))
        if self.use_key_shift_embed:
            out.append((key_shift, self.key_shift_embed))
        if self.use_speed_embed:
            out.append((speed, self.speed_embed))
        return out

    def forward(self, txt_tokens, mel2ph, f0, key_shift=None, speed=None, spk_embed_id=None, languages=None, **kwargs):
        spk = None
        if self.use_spk_id:
            spk = kwargs.get('spk_mix_embed')
            if spk is None:
                spk = self.spk_embed(spk_embed_id)[:, None, :]
        curves = self._curves(f0, key_shift, speed, kwargs)
        needs_grad = torch.is_grad_enabled() and (any(p.requires_grad for p in self.parameters())
                                                  or any(c.requires_grad for c, _ in curves if torch.is_tensor(c)))
        if needs_grad or not txt_tokens.is_cuda:
            return self._forward_torch(txt_tokens, mel2ph, languages, curves, spk)
        return self._forward_native(txt_tokens, mel2ph, languages, curves, spk)

    def _forward_torch(self, txt_tokens, mel2ph, languages, curves, spk):
        txt_embed, extra, pad = self._token_inputs(txt_tokens, mel2ph, languages)
        enc = F.pad(self.encoder(txt_embed, extra, pad), [0, 0, 1, 0])
        cond = torch.gather(enc, 1, mel2ph[..., None].repeat([1, 1, enc.shape[-1]]))
        if spk is not None:
            cond = cond + spk
        for c, lin in curves:
            cond = cond + lin(c[:, :, None])
        return cond

    # ------------------------------------------------------------------ native path
    def _pack(self, dev):
        key = tuple((p.data_ptr(), p._version) for p in self.parameters())
        if self._packed is not None and self._packed_key == key and self._packed['dev'] == dev:
            return self._packed
        bf, f32 = torch.bfloat16, torch.float32
        H = self.hidden_size
        layers = []
        for lay in self.encoder.layers:
            op = lay.op
            k = op.ffn.kernel_size
            w1 = op.ffn.ffn_1.weight.detach().to(f32) * k ** -0.5                      # [4H, H, k]; the k^-0.5 of common_layers.py:144 folded
            layers.append(dict(
                g1=op.layer_norm1.weight.detach().to(f32).contiguous(), b1=op.layer_norm1.bias.detach().to(f32).contiguous(),
                g2=op.layer_norm2.weight.detach().to(f32).contiguous(), b2=op.layer_norm2.bias.detach().to(f32).contiguous(),
                w_in=op.self_attn.in_proj.weight.detach().to(bf).contiguous(),        # [3H, H]
                w_out=op.self_attn.out_proj.weight.detach().to(bf).contiguous(),      # [H, H]
                w_f1=w1.permute(0, 2, 1).reshape(w1.shape[0], k * H).to(bf).contiguous(),  # [4H, k*H], tap-major K
                b_f1=(op.ffn.ffn_1.bias.detach().to(f32) * k ** -0.5).contiguous(),
                w_f2=op.ffn.ffn_2.weight.detach().to(bf).contiguous(),                # [H, 4H]
                b_f2=op.ffn.ffn_2.bias.detach().to(f32).contiguous(),
                k=k, act=op.ffn.act, theta=op.self_attn.rotary_embed.theta))
        self._packed = dict(dev=dev, layers=layers, zeros3=torch.zeros(3 * H, device=dev), zeros=torch.zeros(H, device=dev),
                            gf=self.encoder.layer_norm.weight.detach().to(f32).contiguous(),
                            bff=self.encoder.layer_norm.bias.detach().to(f32).contiguous())
        self._packed_key = key
        return self._packed

    @torch.no_grad()
    def _forward_native(self, txt_tokens, mel2ph, languages, curves, spk):
        import ctypes as C
        from . import _native as nat
        from . import ops
        nat.require_device()
        dev = txt_tokens.device
        H, heads = self.hidden_size, self.encoder.num_heads
        pk = self._pack(dev)
        B, Tt = txt_tokens.shape
        T = mel2ph.shape[1]
        bf = torch.bfloat16
        txt_embed, extra, pad = self._token_inputs(txt_tokens, mel2ph, languages)
        pad_u8 = pad.to(torch.uint8).contiguous()
        x32 = ((self.encoder.embed_scale * txt_embed + extra) * (~pad)[..., None]).reshape(B * Tt, H).contiguous()
        x = torch.empty(B * Tt, H, dtype=bf, device=dev)
        ops.cast_bf16(x, x32)
        a = torch.empty_like(x)
        qkv = torch.empty(B * Tt, 3 * H, dtype=bf, device=dev)
        att = torch.empty_like(x)
        hbuf = torch.empty(B * Tt, 4 * H, dtype=bf, device=dev)
        L = nat.lib()
        st = ops.stream
        for lay in pk['layers']:
            ops.layernorm(a, x, lay['g1'], lay['b1'])
            ops.gemm(ops.EPI_LINEAR, ops.tile_n(3 * H), B, Tt, a, lay['w_in'], qkv, bias=pk['zeros3'])
            nat.check(L.ds_attention(qkv.data_ptr(), att.data_ptr(), pad_u8.data_ptr(), B, Tt, H, heads, 1, lay['theta'], st()))
            ops.gemm(ops.EPI_RES, ops.tile_n(H), B, Tt, att, lay['w_out'], x, res=x, vec=pk['zeros'])
            nat.check(L.ds_mask_rows(x.data_ptr(), pad_u8.data_ptr(), B * Tt, H, st()))
            ops.layernorm(a, x, lay['g2'], lay['b2'])
            ops.gemm(ops.EPI_LINEAR, ops.tile_n(4 * H), B, Tt, a, lay['w_f1'], hbuf, k_per_tap=H, n_taps=lay['k'], dil=1, bias=lay['b_f1'],
                     act=ops.ACT_GELU if lay['act'] == 'gelu' else ops.ACT_RELU)
            ops.gemm(ops.EPI_RES, ops.tile_n(H), B, Tt, hbuf, lay['w_f2'], x, res=x, vec=lay['b_f2'])
            nat.check(L.ds_mask_rows(x.data_ptr(), pad_u8.data_ptr(), B * Tt, H, st()))
        ops.layernorm(a, x, pk['gf'], pk['bff'])
        nat.check(L.ds_mask_rows(a.data_ptr(), pad_u8.data_ptr(), B * Tt, H, st()))
        # length regulation + every per-frame Linear(1, H) + the speaker embedding in one pass          acoustic_encoder.py:97-116
        keep = []
        arr = (nat.CurveEmbed * max(len(curves), 1))()
        for j, (c, lin) in enumerate(curves):
            c = c.to(dtype=torch.float32)
            w = lin.weight.detach().to(torch.float32)[:, 0].contiguous()
            b_ = lin.bias.detach().to(torch.float32).contiguous()
            keep += [c, w, b_]
            arr[j].value = c.data_ptr()
            arr[j].stride_b = c.stride(0) if c.shape[0] > 1 else 0
            arr[j].stride_t = c.stride(1) if c.shape[1] > 1 else 0
            arr[j].weight, arr[j].bias = w.data_ptr(), b_.data_ptr()
        add_ptr, add_sb, add_st = None, 0, 0
        if spk is not None:
            spk = spk.to(torch.float32)
            if spk.stride(-1) != 1:
                spk = spk.contiguous()
            keep.append(spk)
            add_ptr = spk.data_ptr()
            add_sb = spk.stride(0) if spk.shape[0] > 1 else 0
            add_st = spk.stride(1) if spk.shape[1] > 1 else 0
        cond = torch.empty(B, T, H, dtype=torch.float32, device=dev)
        mel2ph = mel2ph.to(torch.int64).contiguous()
        nat.check(L.ds_length_regulate(cond.data_ptr(), a.data_ptr(), mel2ph.data_ptr(), B, T, Tt, H, arr, len(curves), add_ptr, add_sb,
                                       add_st, st()))
        for t in keep + [a, x, qkv, att, hbuf, pad_u8, mel2ph]:
            if t.is_cuda:
                t.record_stream(torch.cuda.current_stream())
        return cond
