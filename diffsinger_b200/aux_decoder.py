"""Drop-in for the reference's ``modules/aux_decoder`` (ConvNeXt aux decoder, the shallow-diffusion front that produces
``x_start`` right before the sampling loop in the default acoustic config; SURVEY.md section 8f #2,
configs/acoustic.yaml:85-101).

Same constructor signatures and parameter names / shapes as ``modules/aux_decoder/convnext.py:58-87`` and
``modules/aux_decoder/__init__.py:30-75``.  Under ``torch.no_grad()`` on a CUDA device the decoder is composed from the
primitive ops of libdsb200.so: the k = 7 input / output convolutions run as 7-tap implicit GEMMs (time-shifted TMA tiles,
zero fill = zero padding), each ConvNeXt block as depthwise k = 7 -> LayerNorm -> GEMM + GELU -> GEMM + layer-scale +
residual (layer scale folded into the packed weights).  With autograd enabled it is plain differentiable torch.
"""
from typing import Optional

import torch
import torch.nn as nn

from . import ops
from .backbones import filter_kwargs


class ConvNeXtBlock(nn.Module):
    def __init__(self, dim: int, intermediate_dim: int, layer_scale_init_value: Optional[float] = None, drop_out: float = 0.0):
        super().__init__()
        self.dwconv = nn.Conv1d(dim, dim, kernel_size=7, padding=3, groups=dim)
        self.norm = nn.LayerNorm(dim, eps=1e-6)
        self.pwconv1 = nn.Linear(dim, intermediate_dim)
        self.act = nn.GELU()
        self.pwconv2 = nn.Linear(intermediate_dim, dim)
        self.gamma = (nn.Parameter(layer_scale_init_value * torch.ones(dim), requires_grad=True)
                      if layer_scale_init_value > 0 else None)
        self.drop_path = nn.Identity()
        self.dropout = nn.Dropout(drop_out) if drop_out > 0. else nn.Identity()

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        residual = x
        x = self.dwconv(x).transpose(1, 2)
        x = self.pwconv2(self.act(self.pwconv1(self.norm(x))))
        if self.gamma is not None:
            x = self.gamma * x
        x = self.dropout(x.transpose(1, 2))
        return residual + self.drop_path(x)


class ConvNeXtDecoder(nn.Module):
    def __init__(self, in_dims, out_dims, /, *, num_channels=512, num_layers=6, kernel_size=7, dropout_rate=0.1):
        super().__init__()
        self.in_dims, self.out_dims, self.num_channels, self.kernel_size = in_dims, out_dims, num_channels, kernel_size
        self.inconv = nn.Conv1d(in_dims, num_channels, kernel_size, stride=1, padding=(kernel_size - 1) // 2)
        self.conv = nn.ModuleList(
            ConvNeXtBlock(dim=num_channels, intermediate_dim=num_channels * 4, layer_scale_init_value=1e-6, drop_out=dropout_rate)
            for _ in range(num_layers))
        self.outconv = nn.Conv1d(num_channels, out_dims, kernel_size, stride=1, padding=(kernel_size - 1) // 2)
        self._packed = None

    def __getstate__(self):
        d = self.__dict__.copy()
        d['_packed'] = None
        return d

    # noinspection PyUnusedLocal
    def forward(self, x, infer=False, lengths=None):
        """x [B, T, in_dims] -> [B, T, out_dims].  lengths (native path only): frame counts of a ragged batch whose rows beyond an
        utterance's end are ZERO in x; the residual stream is then kept at zero there, so the k = 7 convolutions see each utterance's
        own padding and every utterance equals its batch-1 result (outputs beyond the end are undefined)."""
        if torch.is_grad_enabled() and (x.requires_grad or any(p.requires_grad for p in self.parameters())):
            if lengths is not None:
                raise RuntimeError("ragged batches are an inference feature of the native path")
            x = self.inconv(x.transpose(1, 2))
            for conv in self.conv:
                x = conv(x)
            return self.outconv(x).transpose(1, 2)
        return self._forward_native(x, lengths)

    # ------------------------------------------------------------------ native path
    def _pack(self, dev):
        key = tuple((p.data_ptr(), p._version) for p in self.parameters())
        if self._packed is not None and self._packed["key"] == key:
            return self._packed
        if dev.type != 'cuda':
            raise RuntimeError("diffsinger_b200 ConvNeXtDecoder has no CPU inference path (the torch-op branch is for autograd only)")
        if self.kernel_size % 2 == 0 or self.num_channels % 64:
            raise RuntimeError("native ConvNeXtDecoder needs an odd kernel_size and num_channels % 64 == 0")
        bf, C, k = torch.bfloat16, self.num_channels, self.kernel_size
        f32 = lambda t: t.detach().to(device=dev, dtype=torch.float32).contiguous()
        Hp, bn_o = ops.pad64(self.in_dims), (256 if self.out_dims > 128 else (128 if self.out_dims > 64 else 64))
        Or = (self.out_dims + bn_o - 1) // bn_o * bn_o
        w_in = torch.zeros(C, k, Hp, device=dev)  # K order: tap-major, [tap][c_in]
        w_in[:, :, :self.in_dims] = f32(self.inconv.weight).permute(0, 2, 1)
        w_out = torch.zeros(Or, k, C, device=dev)
        w_out[:self.out_dims] = f32(self.outconv.weight).permute(0, 2, 1)
        b_out = torch.zeros(Or, device=dev)
        b_out[:self.out_dims] = f32(self.outconv.bias)
        blocks = []
        for blk in self.conv:
            g = f32(blk.gamma) if blk.gamma is not None else torch.ones(C, device=dev)
            blocks.append(dict(dw_w=f32(blk.dwconv.weight)[:, 0, :].contiguous(), dw_b=f32(blk.dwconv.bias),
                               ln_g=f32(blk.norm.weight), ln_b=f32(blk.norm.bias),
                               w1=f32(blk.pwconv1.weight).to(bf).contiguous(), b1=f32(blk.pwconv1.bias),
                               w2=(g[:, None] * f32(blk.pwconv2.weight)).to(bf).contiguous(), b2=(g * f32(blk.pwconv2.bias)).contiguous()))
        self._packed = dict(key=key, Hp=Hp, bn_o=bn_o, w_in=w_in.reshape(C, k * Hp).to(bf).contiguous(), b_in=f32(self.inconv.bias),
                            w_out=w_out.reshape(Or, k * C).to(bf).contiguous(), b_out=b_out, blocks=blocks, bufs={})
        return self._packed

    @torch.no_grad()
    def _forward_native(self, x, lengths=None):
        B, T, _ = x.shape
        lens = None
        if lengths is not None:
            lengths = [int(v) for v in lengths]
            if len(lengths) != B or min(lengths) < 1 or max(lengths) > T:
                raise ValueError(f"ragged batch: need {B} lengths in [1, {T}], got {lengths}")
            lens = torch.tensor(lengths, dtype=torch.int32).to(x.device)
        P = self._pack(x.device)
        C, rows, bf = self.num_channels, B * T, torch.bfloat16
        bkey = (B, T, ops.stream())
        bufs = P["bufs"].get(bkey)
        if bufs is None:
            if len(P["bufs"]) >= 16:
                P["bufs"].clear()
            mk = lambda ch: torch.empty(rows, ch, dtype=bf, device=x.device)
            bufs = P["bufs"][bkey] = dict(xb=mk(P["Hp"]), h=mk(C), d=mk(C), n=mk(C), m=mk(4 * C))
        xb, h, d, n, m = bufs["xb"], bufs["h"], bufs["d"], bufs["n"], bufs["m"]
        ops.cast_bf16(xb, x.detach().to(torch.float32).contiguous().view(rows, self.in_dims))
        bn_c = ops.tile_n(C)
        ops.gemm(ops.EPI_LINEAR, bn_c, B, T, xb, P["w_in"], h, n_taps=self.kernel_size, bias=P["b_in"], lengths=lens)
        for blk in P["blocks"]:
            ops.dwconv(d, h, blk["dw_w"], blk["dw_b"], B, T, 7)
            ops.layernorm(n, d, blk["ln_g"], blk["ln_b"], 1e-6)
            ops.gemm(ops.EPI_LINEAR, ops.tile_n(4 * C), B, T, n, blk["w1"], m, bias=blk["b1"], act=ops.ACT_GELU)
            ops.gemm(ops.EPI_RES, bn_c, B, T, m, blk["w2"], h, vec=blk["b2"], vec_bstride=0, res=h, lengths=lens)
        out = torch.empty(B, T, self.out_dims, dtype=torch.float32, device=x.device)
        ops.gemm(ops.EPI_OUTF32, P["bn_o"], B, T, h, P["w_out"], out.view(rows, self.out_dims), n_taps=self.kernel_size, bias=P["b_out"],
                 n_real=self.out_dims)
        return out


AUX_DECODERS = {'convnext': ConvNeXtDecoder}
AUX_LOSSES = {'convnext': nn.L1Loss}


def build_aux_decoder(in_dims: int, out_dims: int, aux_decoder_arch: str, aux_decoder_args: dict) -> torch.nn.Module:
    decoder_cls = AUX_DECODERS[aux_decoder_arch]
    kwargs = filter_kwargs(aux_decoder_args, decoder_cls)
    return AUX_DECODERS[aux_decoder_arch](in_dims, out_dims, **kwargs)


def build_aux_loss(aux_decoder_arch):
    return AUX_LOSSES[aux_decoder_arch]()


class AuxDecoderAdaptor(nn.Module):
    """reference modules/aux_decoder/__init__.py:30-75."""

    def __init__(self, in_dims: int, out_dims: int, num_feats: int, spec_min: list, spec_max: list,
                 aux_decoder_arch: str, aux_decoder_args: dict):
        super().__init__()
        self.decoder = build_aux_decoder(in_dims=in_dims, out_dims=out_dims * num_feats,
                                         aux_decoder_arch=aux_decoder_arch, aux_decoder_args=aux_decoder_args)
        self.out_dims = out_dims
        self.n_feats = num_feats
        if spec_min is not None and spec_max is not None:
            spec_min = torch.FloatTensor(spec_min)[None, None, :].transpose(-3, -2)
            spec_max = torch.FloatTensor(spec_max)[None, None, :].transpose(-3, -2)
            self.register_buffer('spec_min', spec_min, persistent=False)
            self.register_buffer('spec_max', spec_max, persistent=False)

    def norm_spec(self, x):
        k = (self.spec_max - self.spec_min) / 2.
        b = (self.spec_max + self.spec_min) / 2.
        return (x - b) / k

    def denorm_spec(self, x):
        k = (self.spec_max - self.spec_min) / 2.
        b = (self.spec_max + self.spec_min) / 2.
        return x * k + b

    def forward(self, condition, infer=False, lengths=None):
        x = self.decoder(condition, infer=infer) if lengths is None else self.decoder(condition, infer=infer, lengths=lengths)  # [B, T, F x C]
        if self.n_feats > 1:
            x = x.reshape(-1, x.shape[1], self.n_feats, self.out_dims).transpose(1, 2)  # [B, F, T, C]
        if infer:
            x = self.denorm_spec(x)
        return x
