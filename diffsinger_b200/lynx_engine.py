"""Host-side composition of the LYNXNet denoiser from the primitive ops of libdsb200.so (include/dsb200.h).

One evaluation (reference ``LYNXNet.forward``, modules/backbones/lynxnet.py:133-172) on time-major activations
``[B*T, channels]`` (bf16 between ops, fp32 at the sampler boundary):

    cast x -> bf16 | GEMM input_projection (+GELU)                                             lynxnet.py:145-147
    per layer:       GEMM conditioner_projection, epilogue  + x + (b_c + W_d d + b_d)[b]       :82-89
                     LayerNorm | GEMM C -> 2I, SwiGLU epilogue | depthwise k conv + act        :52-60
                     GEMM I -> C, epilogue + residual + bias                                   :61,92
    LayerNorm | GEMM output_projection -> fp32                                                 :157-160

PyTorch supplies device memory and the current stream; the per-step embedding vectors ``W_d * mlp(emb(t)) + b_d`` are the
only torch arithmetic (O(C^2) per step value, computed once per sampler configuration).
"""
from __future__ import annotations

import ctypes as C
from typing import Dict, List, Sequence, Tuple

import torch
import torch.nn.functional as F

from . import _native as nat
from .engine import Program


def _st() -> int:
    return torch.cuda.current_stream().cuda_stream


def _bn(n: int) -> int:
    if n % 64:
        raise nat.NativeError(f"GEMM width {n} must be a multiple of 64")
    return 256 if n % 256 == 0 else (128 if n % 128 == 0 else 64)


def _pad64(n: int) -> int:
    return (n + 63) // 64 * 64


class LynxWorkspace:
    """Activation buffers of one (B, T) on the stream that created them."""

    def __init__(self, eng: "LynxEngine", B: int, T: int):
        self.eng, self.B, self.T = eng, B, T
        rows, dev, bf = B * T, eng.device, torch.bfloat16
        mk = lambda ch: torch.empty(rows, ch, dtype=bf, device=dev)
        self.xb, self.cond = mk(eng.FMp), mk(eng.H)
        self.h, self.u, self.n, self.g, self.dw = mk(eng.C), mk(eng.C), mk(eng.C), mk(eng.I), mk(eng.I)
        self._bufs: List[torch.Tensor] = []
        self.graphs: Dict[Tuple, torch.cuda.CUDAGraph] = {}
        self.last_launches = 0

    def set_cond(self, cond_bth: torch.Tensor):
        c = cond_bth.detach().to(torch.float32).contiguous()
        assert c.shape == (self.B, self.T, self.eng.H), c.shape
        nat.check(nat.lib().ds_cast_bf16(self.cond.data_ptr(), c.data_ptr(), self.B * self.T, self.eng.H, self.eng.H, _st()))

    def bufs(self, n: int) -> List[torch.Tensor]:
        while len(self._bufs) < n:
            self._bufs.append(torch.empty(self.B, self.T, self.eng.FM, dtype=torch.float32, device=self.eng.device))
        return self._bufs[:n]


class LynxEngine:
    MAX_WORKSPACES = 32

    def __init__(self, module, device: torch.device):
        nat.require_device()
        self.device = device
        self.M, self.Fn = module.in_dims, module.n_feats
        self.FM, self.FMp = self.M * self.Fn, _pad64(self.M * self.Fn)
        self.C, self.L, self.H = module.num_channels, module.num_layers, module.hidden_size
        self.I = self.C * module.expansion_factor
        self.k = module.kernel_size
        self.act = {'ReLU': 1, 'PReLU': 2, 'SiLU': 3}[module.activation]
        self.strong = bool(module.strong_cond)
        for name, v in (("num_channels", self.C), ("hidden_size", self.H), ("inner dim", self.I)):
            if v % 64:
                raise nat.NativeError(f"LYNXNet {name}={v} must be a multiple of 64")
        if self.k not in (3, 7, 31):
            raise nat.NativeError(f"depthwise kernel_size={self.k} not instantiated (3, 7, 31)")
        self.bn_c, self.bn_g = _bn(self.C), (256 if (2 * self.I) % 256 == 0 else 128)
        self.bn_o = 256 if self.FM > 128 else (128 if self.FM > 64 else 64)
        self.FMr = (self.FM + self.bn_o - 1) // self.bn_o * self.bn_o
        self.workspaces: Dict[Tuple, LynxWorkspace] = {}
        self.weights_key = None
        self._vec_cache: Dict[Tuple, torch.Tensor] = {}

    @staticmethod
    def weights_fingerprint(module) -> Tuple:
        return tuple((p.data_ptr(), p._version) for p in module.parameters())

    # ------------------------------------------------------------------ weights
    def pack(self, module):
        """bf16 [rows][K] matrices (K-major), fp32 biases / norm / depthwise parameters.  The C -> 2I matrix is permuted
        so that every BN-row pass holds BN/2 'out' rows followed by the BN/2 'gate' rows of the same channels (SwiGLU
        pairs channel c with c + I, lynxnet.py:57)."""
        bf, dev = torch.bfloat16, self.device
        f32 = lambda t: t.detach().to(device=dev, dtype=torch.float32).contiguous()
        sd = module.state_dict()
        w_in = torch.zeros(self.C, self.FMp, dtype=bf, device=dev)
        w_in[:, :self.FM] = f32(sd["input_projection.weight"])[:, :, 0].to(bf)
        w_out = torch.zeros(self.FMr, self.C, dtype=bf, device=dev)
        w_out[:self.FM] = f32(sd["output_projection.weight"])[:, :, 0].to(bf)
        b_out = torch.zeros(self.FMr, dtype=torch.float32, device=dev)
        b_out[:self.FM] = f32(sd["output_projection.bias"])
        half = self.bn_g // 2
        perm = torch.cat([torch.cat([torch.arange(p * half, (p + 1) * half), self.I + torch.arange(p * half, (p + 1) * half)])
                          for p in range(2 * self.I // self.bn_g)]).to(dev)
        layers = []
        for i in range(self.L):
            p, q = f"residual_layers.{i}.", f"residual_layers.{i}.convmodule.net."
            layers.append(dict(
                wc=f32(sd[p + "conditioner_projection.weight"])[:, :, 0].to(bf).contiguous(),
                bc=f32(sd[p + "conditioner_projection.bias"]),
                ln_g=f32(sd[q + "0.weight"]), ln_b=f32(sd[q + "0.bias"]),
                w2=f32(sd[q + "2.weight"])[:, :, 0][perm].to(bf).contiguous(), b2=f32(sd[q + "2.bias"])[perm].contiguous(),
                dw_w=f32(sd[q + "4.weight"])[:, 0, :].contiguous(), dw_b=f32(sd[q + "4.bias"]),
                slope=f32(sd[q + "5.weight"]) if self.act == 2 else None,
                w6=f32(sd[q + "6.weight"])[:, :, 0].to(bf).contiguous(), b6=f32(sd[q + "6.bias"])))
        self.w = dict(w_in=w_in, b_in=f32(sd["input_projection.bias"]), w_out=w_out, b_out=b_out,
                      norm_g=f32(sd["norm.weight"]), norm_b=f32(sd["norm.bias"]), layers=layers)
        self.weights_key = self.weights_fingerprint(module)
        self._vec_cache.clear()
        for ws in self.workspaces.values():
            ws.graphs.clear()

    @torch.no_grad()
    def step_vectors(self, module, t: torch.Tensor, B: int, per_sample: bool) -> torch.Tensor:
        """fp32 [n_times, L, Bt, C]: the per-layer step vector W_d_l * mlp(emb(t)) + b_d_l (lynxnet.py:149,89) -- plus the
        conditioner-projection bias when it is added in that GEMM's epilogue (strong_cond = False) -- one row per step
        value (per_sample False: Bt = 1) or per (step value, utterance).  torch arithmetic, O(C^2) per entry."""
        t = t.to(self.device).reshape(-1)
        Bt = B if per_sample else 1
        d = module.diffusion_embedding(t)  # [n_times * Bt, C]
        rows = []
        for l in module.residual_layers:
            v = F.linear(d, l.diffusion_projection.weight[:, :, 0], l.diffusion_projection.bias)
            rows.append(v if self.strong else v + l.conditioner_projection.bias)
        out = torch.stack(rows, dim=0)  # [L, n_times * Bt, C]
        return out.reshape(self.L, -1, Bt, self.C).permute(1, 0, 2, 3).float().contiguous()

    def workspace(self, B: int, T: int) -> LynxWorkspace:
        key = (B, T, _st())
        ws = self.workspaces.get(key)
        if ws is None:
            while len(self.workspaces) >= self.MAX_WORKSPACES:
                self.workspaces.pop(next(iter(self.workspaces)))
            ws = self.workspaces[key] = LynxWorkspace(self, B, T)
        else:
            self.workspaces[key] = self.workspaces.pop(key)
        return ws

    # ------------------------------------------------------------------ one evaluation
    def _gemm(self, epi, BN, B, T, a, a_ch, w, w_rows, w_k, out, ld_out, bias=None, vec=None, vec_bstride=0, res=None, act=0,
              n_real=0, scale=1.0):
        d = nat.GemmDesc()
        d.epi, d.BN, d.B, d.T = epi, BN, B, T
        d.a0, d.a0_channels, d.a0_c0, d.kb_a0, d.n_taps, d.dil = a.data_ptr(), a_ch, 0, w_k // 64, 1, 0
        d.a1, d.a1_channels, d.kb_a1 = None, 0, 0
        d.w, d.w_rows, d.w_k, d.b_row0, d.n_tiles = w.data_ptr(), w_rows, w_k, 0, w_rows // BN
        d.bias = bias.data_ptr() if bias is not None else None
        d.vec_sub = None
        d.vec_add, d.vec_bstride = (vec.data_ptr() if vec is not None else None), vec_bstride
        d.res_in, d.ld_res = (res.data_ptr() if res is not None else None), (res.shape[1] if res is not None else 0)
        d.out, d.ld_out, d.out_c0, d.n_real = out.data_ptr(), ld_out, 0, n_real
        d.relu, d.scale = act, scale
        nat.check(nat.lib().ds_gemm(C.byref(d), _st()))

    def eval(self, ws: LynxWorkspace, x_f32: torch.Tensor, vecs: torch.Tensor, out_f32: torch.Tensor) -> int:
        """x / out: fp32 [B, T, F*M] time-major; vecs: fp32 [L, Bt, C] step vectors of this evaluation.  Returns #launches."""
        L, w, B, T, Cc, I = nat.lib(), self.w, ws.B, ws.T, self.C, self.I
        rows = B * T
        Bt = vecs.shape[1]
        nat.check(L.ds_cast_bf16(ws.xb.data_ptr(), x_f32.data_ptr(), rows, self.FM, self.FMp, _st()))
        self._gemm(0, self.bn_c, B, T, ws.xb, self.FMp, w["w_in"], Cc, self.FMp, ws.h, Cc, bias=w["b_in"], act=0 if self.strong else 2)
        n = 2
        for i, lw in enumerate(w["layers"]):
            dvec = vecs[i]  # [Bt, C]
            if self.strong:  # res_x = x + cond projection; the step vector is added on the way into the norm
                self._gemm(2, self.bn_c, B, T, ws.cond, self.H, lw["wc"], Cc, self.H, ws.u, Cc, vec=lw["bc"], vec_bstride=0, res=ws.h)
                res, ln_vec = ws.u, dvec
            else:            # res_x = x; u = x + cond projection + (bias + step vector)
                self._gemm(2, self.bn_c, B, T, ws.cond, self.H, lw["wc"], Cc, self.H, ws.u, Cc, vec=dvec,
                           vec_bstride=Cc if Bt > 1 else 0, res=ws.h)
                res, ln_vec = ws.h, None
            nat.check(L.ds_layernorm(ws.n.data_ptr(), ws.u.data_ptr(), lw["ln_g"].data_ptr(), lw["ln_b"].data_ptr(), rows, Cc, 1e-5,
                                     ln_vec.data_ptr() if ln_vec is not None else None, T if Bt > 1 else rows, _st()))
            self._gemm(5, self.bn_g, B, T, ws.n, Cc, lw["w2"], 2 * I, Cc, ws.g, I, bias=lw["b2"])
            nat.check(L.ds_dwconv(ws.dw.data_ptr(), ws.g.data_ptr(), lw["dw_w"].data_ptr(), lw["dw_b"].data_ptr(),
                                  lw["slope"].data_ptr() if lw["slope"] is not None else None, B, T, I, self.k, self.k // 2,
                                  self.act, _st()))
            self._gemm(2, self.bn_c, B, T, ws.dw, I, lw["w6"], Cc, I, ws.h, Cc, vec=lw["b6"], vec_bstride=0, res=res)
            n += 5
        nat.check(L.ds_layernorm(ws.n.data_ptr(), ws.h.data_ptr(), w["norm_g"].data_ptr(), w["norm_b"].data_ptr(), rows, Cc, 1e-5,
                                 None, 0, _st()))
        self._gemm(3, self.bn_o, B, T, ws.n, Cc, w["w_out"], self.FMr, Cc, out_f32, self.FM, bias=w["b_out"], n_real=self.FM)
        return n + 2

    # ------------------------------------------------------------------ sampler programs
    @torch.no_grad()
    def run_program(self, module, prog: Program, x_tm: torch.Tensor, cond_bht: torch.Tensor, extra: Sequence[torch.Tensor] = (),
                    use_graph: bool = True) -> torch.Tensor:
        """Runs a sampler Program (core/programs.py) with this backbone: EVAL = self.eval, LINCOMB = ds_lincomb, the whole op
        list captured in one CUDA graph per (program, B, T) and replayed on later calls."""
        B, T, FM = x_tm.shape
        ws = self.workspace(B, T)
        bufs = ws.bufs(prog.n_bufs)
        bufs[0].copy_(x_tm)
        for k, e in enumerate(extra):
            bufs[2 + k].copy_(e)
        ws.set_cond(cond_bht.transpose(1, 2))
        vkey = (tuple(prog.times), self.weights_key)
        vecs = self._vec_cache.get(vkey)
        if vecs is None:
            if len(self._vec_cache) > 16:
                self._vec_cache.clear()
            vecs = self._vec_cache[vkey] = self.step_vectors(module, torch.tensor(prog.times, dtype=torch.float32), B, False)
        gkey = (id(prog), vkey)
        if not use_graph or len(prog.ops) > 1024:
            ws.last_launches = self._run_ops(ws, prog, bufs, vecs)
            return bufs[0].clone()
        graph = ws.graphs.get(gkey)
        if graph is None:
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph):
                ws.graph_launches = self._run_ops(ws, prog, bufs, vecs)
            ws.graphs[gkey] = graph
            ws._graph_refs = getattr(ws, "_graph_refs", []) + [prog, vecs]  # keep ids / device pointers valid
        graph.replay()
        ws.last_launches = ws.graph_launches
        return bufs[0].clone()

    def _run_ops(self, ws: LynxWorkspace, prog: Program, bufs: List[torch.Tensor], vecs: torch.Tensor) -> int:
        L = nat.lib()
        numel = bufs[0].numel()
        launches = 0
        for kind, dst, time, srcs, coefs in prog.ops:
            if kind == nat.DS_OP_EVAL:
                launches += self.eval(ws, bufs[srcs[0]], vecs[time], bufs[dst])
            else:
                ptrs = (C.c_void_p * len(srcs))(*[bufs[s].data_ptr() for s in srcs])
                cf = (C.c_float * len(coefs))(*coefs)
                nat.check(L.ds_lincomb(bufs[dst].data_ptr(), ptrs, cf, len(srcs), numel, _st()))
                launches += 1
        return launches
