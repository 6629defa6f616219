"""Drop-in for the reference's ``modules/backbones/lynxnet.py`` (registry key ``'lynxnet'``, the default acoustic backbone,
configs/acoustic.yaml:73-79).

Same constructor signature, parameter names / shapes (``load_ckpt(strict=True)`` works both ways) and
``forward(spec, diffusion_step, cond)`` contract as the reference.  Under ``torch.no_grad()`` the forward is composed, on
the host, from the primitive ops of libdsb200.so (include/dsb200.h): tcgen05 GEMMs for every 1x1 convolution (GELU /
SwiGLU / residual epilogues fused), ``ds_layernorm``, ``ds_dwconv`` (depthwise k = 31 + PReLU/SiLU/ReLU) -- 34 launches per
evaluation, captured with the sampler update in one CUDA graph by core/runtime.py.  The only torch arithmetic on that path
is the diffusion-step embedding MLP (O(C^2) per step value, once per sampler configuration).  With autograd enabled the
forward is expressed with differentiable torch ops (training ``p_losses``); there is no CPU inference path.
"""
import torch
import torch.nn as nn
import torch.nn.functional as F

from ..hparams import get_hparams
from .wavenet import SinusoidalPosEmb


class _Conv1d(nn.Conv1d):
    def __init__(self, *args, **kwargs):
        super().__init__(*args, **kwargs)
        nn.init.kaiming_normal_(self.weight)


class Transpose(nn.Module):
    def __init__(self, dims):
        super().__init__()
        assert len(dims) == 2, 'dims must be a tuple of two dimensions'
        self.dims = dims

    def forward(self, x):
        return x.transpose(*self.dims)


class SwiGLU(nn.Module):
    """reference modules/commons/common_layers.py: out, gate = chunk(2); out * silu(gate)."""

    def __init__(self, dim=-1):
        super().__init__()
        self.dim = dim

    def forward(self, x):
        out, gate = x.chunk(2, dim=self.dim)
        return out * F.silu(gate)


class LYNXConvModule(nn.Module):
    """Parameter container + differentiable forward (reference lynxnet.py:29-70); ``net`` indices match the reference."""

    @staticmethod
    def calc_same_padding(kernel_size):
        pad = kernel_size // 2
        return pad, pad - (kernel_size + 1) % 2

    def __init__(self, dim, expansion_factor, kernel_size=31, activation='PReLU', dropout=0.0):
        super().__init__()
        inner_dim = dim * expansion_factor
        activation_classes = {'SiLU': nn.SiLU, 'ReLU': nn.ReLU, 'PReLU': lambda: nn.PReLU(inner_dim)}
        activation = activation if activation is not None else 'PReLU'
        if activation not in activation_classes:
            raise ValueError(f'{activation} is not a valid activation')
        padding = self.calc_same_padding(kernel_size)
        self.net = nn.Sequential(
            nn.LayerNorm(dim),
            Transpose((1, 2)),
            nn.Conv1d(dim, inner_dim * 2, 1),
            SwiGLU(dim=1),
            nn.Conv1d(inner_dim, inner_dim, kernel_size=kernel_size, padding=padding[0], groups=inner_dim),
            activation_classes[activation](),
            nn.Conv1d(inner_dim, dim, 1),
            Transpose((1, 2)),
            nn.Dropout(dropout) if float(dropout) > 0. else nn.Identity()
        )

    def forward(self, x):
        return self.net(x)


class LYNXNetResidualLayer(nn.Module):
    def __init__(self, dim_cond, dim, expansion_factor, kernel_size=31, activation='PReLU', dropout=0.0):
        super().__init__()
        self.diffusion_projection = nn.Conv1d(dim, dim, 1)
        self.conditioner_projection = nn.Conv1d(dim_cond, dim, 1)
        self.convmodule = LYNXConvModule(dim=dim, expansion_factor=expansion_factor, kernel_size=kernel_size,
                                         activation=activation, dropout=dropout)

    def forward(self, x, conditioner, diffusion_step, front_cond_inject=False):
        if front_cond_inject:
            x = x + self.conditioner_projection(conditioner)
            res_x = x
        else:
            res_x = x
            x = x + self.conditioner_projection(conditioner)
        x = x + self.diffusion_projection(diffusion_step)
        x = x.transpose(1, 2)
        x = self.convmodule(x)
        x = x.transpose(1, 2) + res_x
        return x


class LYNXNet(nn.Module):
    def __init__(self, in_dims, n_feats, *, num_layers=6, num_channels=512, expansion_factor=2, kernel_size=31,
                 activation='PReLU', dropout=0.0, strong_cond=False):
        super().__init__()
        self.in_dims = in_dims
        self.n_feats = n_feats
        self.num_layers, self.num_channels = num_layers, num_channels
        self.expansion_factor, self.kernel_size = expansion_factor, kernel_size
        self.activation = activation if activation is not None else 'PReLU'
        self.hidden_size = get_hparams()['hidden_size']  # read at construct time like lynxnet.py:111
        self.input_projection = _Conv1d(in_dims * n_feats, num_channels, 1)
        self.diffusion_embedding = nn.Sequential(
            SinusoidalPosEmb(num_channels),
            nn.Linear(num_channels, num_channels * 4),
            nn.GELU(),
            nn.Linear(num_channels * 4, num_channels),
        )
        self.residual_layers = nn.ModuleList([
            LYNXNetResidualLayer(dim_cond=self.hidden_size, dim=num_channels, expansion_factor=expansion_factor,
                                 kernel_size=kernel_size, activation=activation, dropout=dropout)
            for _ in range(num_layers)
        ])
        self.norm = nn.LayerNorm(num_channels)
        self.output_projection = _Conv1d(num_channels, in_dims * n_feats, kernel_size=1)
        self.strong_cond = strong_cond
        nn.init.zeros_(self.output_projection.weight)
        self._engine = None

    # ------------------------------------------------------------------ native engine plumbing
    def __deepcopy__(self, memo):
        import copy
        new = self.__class__.__new__(self.__class__)
        memo[id(self)] = new
        for k, v in self.__dict__.items():
            new.__dict__[k] = None if k == '_engine' else copy.deepcopy(v, memo)
        return new

    def __getstate__(self):
        d = self.__dict__.copy()
        d['_engine'] = None
        return d

    def engine(self):
        from ..lynx_engine import LynxEngine
        dev = self.input_projection.weight.device
        if dev.type != 'cuda':
            raise RuntimeError("diffsinger_b200.LYNXNet has no CPU inference path: move the module to a B200 (sm_100a) device; "
                               "the torch-op branch is for autograd only.")
        if self.kernel_size % 2 == 0:
            raise RuntimeError("the native LYNXNet path needs an odd depthwise kernel_size (3, 7 or 31)")
        if self._engine is None or self._engine.device != dev:
            self._engine = LynxEngine(self, dev)
        if self._engine.weights_key != LynxEngine.weights_fingerprint(self):
            self._engine.pack(self)
        return self._engine

    # ------------------------------------------------------------------ forward
    def forward(self, spec, diffusion_step, cond):
        """spec [B, F, M, T], diffusion_step [B] or [1], cond [B, H, T] -> [B, F, M, T]."""
        if torch.is_grad_enabled() and (spec.requires_grad or cond.requires_grad or any(p.requires_grad for p in self.parameters())):
            return self._forward_autograd(spec, diffusion_step, cond)
        return self._forward_native(spec, diffusion_step, cond)

    @torch.no_grad()
    def _forward_native(self, spec, diffusion_step, cond):
        eng = self.engine()
        B, Fn, M, T = spec.shape
        x = spec.reshape(B, Fn * M, T).transpose(1, 2).contiguous().float()
        ws = eng.workspace(B, T)
        ws.set_cond(cond.transpose(1, 2))
        per_sample = diffusion_step.numel() > 1
        vecs = eng.step_vectors(self, diffusion_step.reshape(-1), B, per_sample)  # [1, L, B or 1, C]
        out = torch.empty_like(x)
        eng.eval(ws, x, vecs[0], out)
        return out.transpose(1, 2).reshape(B, Fn, M, T)

    def _forward_autograd(self, spec, diffusion_step, cond):
        x = spec[:, 0] if self.n_feats == 1 else spec.flatten(start_dim=1, end_dim=2)
        x = self.input_projection(x)
        if not self.strong_cond:
            x = F.gelu(x)
        d = self.diffusion_embedding(diffusion_step).unsqueeze(-1)
        for layer in self.residual_layers:
            x = layer(x, cond, d, front_cond_inject=self.strong_cond)
        x = self.norm(x.transpose(1, 2)).transpose(1, 2)
        x = self.output_projection(x)
        if self.n_feats == 1:
            return x[:, None, :, :]
        return x.reshape(-1, self.n_feats, self.in_dims, x.shape[2])
