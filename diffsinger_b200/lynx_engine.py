"""Host-side composition of the LYNXNet denoiser from the primitive ops of libdsb200.so (include/dsb200.h).

One evaluation (reference ``LYNXNet.forward``, modules/backbones/lynxnet.py:133-172) on time-major activations
``[B*T, channels]`` (bf16 between ops, fp32 at the sampler boundary):

    cast x -> bf16 | GEMM input_projection (+GELU)                                             lynxnet.py:145-147
    per layer:       GEMM conditioner_projection, epilogue  + x + (b_c + W_d d + b_d)[b]       :82-89
                     (sampler programs, strong_cond = False: W_c cond is step-invariant, so it is computed once per
                     run for all layers and added -- with the step vector -- on the way into the LayerNorm)
                     LayerNorm | GEMM C -> 2I, SwiGLU epilogue | depthwise k conv + act        :52-60
                     GEMM I -> C, epilogue + residual + bias                                   :61,92
    LayerNorm | GEMM output_projection -> fp32                                                 :157-160

PyTorch supplies device memory and the current stream; the per-step embedding vectors ``W_d * mlp(emb(t)) + b_d`` are the
only torch arithmetic (O(C^2) per step value, computed once per sampler configuration).
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Dict, List, Sequence, Tuple

import torch
import torch.nn.functional as F

from . import _native as nat
from . import ops
from .engine import Program


_st, _bn, _pad64 = ops.stream, ops.tile_n, ops.pad64


class LynxWorkspace:
    """Activation buffers of one (B, T) on the stream that created them."""

    def __init__(self, eng: "LynxEngine", B: int, T: int):
        self.eng, self.B, self.T = eng, B, T
        rows, dev, bf = B * T, eng.device, torch.bfloat16
        mk = lambda ch: torch.empty(rows, ch, dtype=bf, device=dev)
        self.xb, self.cond = mk(eng.FMp), mk(eng.H)
        self.h, self.u, self.n, self.g, self.dw = mk(eng.C), mk(eng.C), mk(eng.C), mk(eng.I), mk(eng.I)
        # partial sums of the folded LayerNorm (strong_cond): [rows, 2 * C / bn_c, 4] fp32
        self.stats = torch.empty(rows, 2 * (eng.C // eng.bn_c), 4, dtype=torch.float32, device=dev) if eng.fold_ln else None
        self.cproj = None  # [L, rows, C] bf16: W_c cond of every layer, hoisted out of a sampler program's loop
        # ragged batch: frame count of every utterance (device array read by the SwiGLU GEMM's epilogue: the depthwise conv then sees
        # zeros beyond an utterance's end); None = B full-length utterances.  The tensor persists so that captured graphs follow it.
        self.lengths = None
        self._len_dev = torch.full((B,), T, dtype=torch.int32, device=dev)
        self._bufs: List[torch.Tensor] = []
        self.graphs: Dict[Tuple, torch.cuda.CUDAGraph] = {}
        self.last_launches = 0

    def set_cond(self, cond_bth: torch.Tensor):
        c = cond_bth.detach().to(torch.float32).contiguous()
        assert c.shape == (self.B, self.T, self.eng.H), c.shape
        ops.cast_bf16(self.cond, c.view(self.B * self.T, self.eng.H))

    def set_lengths(self, lengths):
        if lengths is None:
            if self.lengths is not None:
                self._len_dev.fill_(self.T)
                self.lengths = None
            return
        lengths = tuple(int(v) for v in lengths)
        if len(lengths) != self.B or min(lengths) < 1 or max(lengths) > self.T:
            raise nat.NativeError(f"ragged batch: need {self.B} lengths in [1, {self.T}], got {lengths}")
        if lengths != self.lengths:
            self._len_dev.copy_(torch.tensor(lengths, dtype=torch.int32), non_blocking=False)
            self.lengths = lengths

    def hoist_cond(self):
        """cproj[l] = W_c_l cond (bias excluded: it travels with the step vector).  L launches, once per sampler run."""
        eng = self.eng
        if self.cproj is None:
            self.cproj = torch.empty(eng.L, self.B * self.T, eng.C, dtype=torch.bfloat16, device=eng.device)
        for i, lw in enumerate(eng.w["layers"]):
            ops.gemm(ops.EPI_LINEAR, eng.bn_c, self.B, self.T, self.cond, lw["wc"], self.cproj[i], bias=eng.zero_c)
        return eng.L

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
        # strong_cond: the per-layer LayerNorm is folded into the GEMMs around it (pack()); DSB200_LYNX_LN=separate keeps the kernel
        self.fold_ln = self.strong and os.environ.get("DSB200_LYNX_LN", "fold") != "separate"
        for name, v in (("num_channels", self.C), ("hidden_size", self.H), ("inner dim", self.I)):
            if v % 64:
                raise nat.NativeError(f"LYNXNet {name}={v} must be a multiple of 64")
        if self.k not in (3, 7, 31):
            raise nat.NativeError(f"depthwise kernel_size={self.k} not instantiated (3, 7, 31)")
        self.bn_c, self.bn_g = _bn(self.C), (256 if (2 * self.I) % 256 == 0 else 128)
        self.bn_o = 256 if self.FM > 128 else (128 if self.FM > 64 else 64)
        self.FMr = (self.FM + self.bn_o - 1) // self.bn_o * self.bn_o
        self.workspaces: Dict[Tuple, LynxWorkspace] = {}
        self.zero_c = torch.zeros(self.C, dtype=torch.float32, device=device)
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
        if self.strong:
            # strong_cond (the shipped default): the stream a layer keeps is u = x + W_c cond + b_c, and the NEXT layer's injection is
            # linear in cond, so it rides in the producing GEMM as extra K blocks:  u_{l+1} = u_l + [W6_l | Wc_{l+1}] [dw_l ; cond] +
            # (b6_l + bc_{l+1}); the head likewise (no GELU in this mode).  One launch and one [rows, C] round trip less per layer.
            self.w["w_in_c"] = torch.cat([w_in, layers[0]["wc"]], dim=1).contiguous()
            self.w["b_in_c"] = self.w["b_in"] + layers[0]["bc"]
            for i, lw in enumerate(layers):
                nxt = layers[i + 1] if i + 1 < self.L else None
                lw["w6c"] = torch.cat([lw["w6"], nxt["wc"]], dim=1).contiguous() if nxt else lw["w6"]
                lw["b6c"] = lw["b6"] + nxt["bc"] if nxt else lw["b6"]
        if self.fold_ln:
            # LayerNorm(u + d) -> Conv1d(C, 2I, 1) as ONE GEMM over the raw stream u:  y = rstd (W~ u + W~ d - mean W~ 1) + (W beta + b),
            # W~ = W diag(gamma).  mean / rstd come from partial sums the producing GEMM's epilogue writes (ops.gemm: stats_out);
            # s = W~ 1 uses the bf16-rounded W~ the tensor cores multiply with, so that the mean cancels exactly.
            for i, lw in enumerate(layers):
                w2f = f32(sd[f"residual_layers.{i}.convmodule.net.2.weight"])[:, :, 0][perm]
                lw["w2t"] = (w2f * lw["ln_g"][None, :]).to(bf).contiguous()
                lw["ln_s"] = lw["w2t"].float().sum(dim=1).contiguous()
                lw["b2t"] = (lw["b2"] + w2f @ lw["ln_b"]).contiguous()
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

    @torch.no_grad()
    def ln_vectors(self, vecs: torch.Tensor) -> torch.Tensor:
        """Folded LayerNorm: fp32 [..., L, Bt, 2I + 4] = (W~_l d, sum d, sum d^2, 0, 0) for step vectors d = vecs[..., l, b, :]."""
        out = torch.zeros(*vecs.shape[:-1], 2 * self.I + 4, dtype=torch.float32, device=vecs.device)
        for l, lw in enumerate(self.w["layers"]):
            d = vecs[..., l, :, :]
            out[..., l, :, :2 * self.I] = d @ lw["w2t"].float().t()
            out[..., l, :, 2 * self.I] = d.sum(-1)
            out[..., l, :, 2 * self.I + 1] = (d * d).sum(-1)
        return out.contiguous()

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

    def max_chunk_batch(self, B: int, T: int, n_bufs: int) -> int:
        """Largest per-run batch whose workspace (activations, hoisted conditioner projections, sampler buffers) fits the
        device; DSB200_MAX_FRAMES (frames per chunk) overrides the memory probe (core/runtime.py:max_chunk_batch)."""
        cap = os.environ.get("DSB200_MAX_FRAMES")
        if cap:
            return max(1, min(B, int(cap) // T))
        if (B, T, _st()) in self.workspaces:
            return B
        per_utt = T * (2 * (self.FMp + self.H + 3 * self.C + 2 * self.I + self.L * self.C) + 4 * n_bufs * self.FM)
        free, _total = torch.cuda.mem_get_info(self.device)
        free += torch.cuda.memory_reserved(self.device) - torch.cuda.memory_allocated(self.device)
        fit = int(0.8 * free) // per_utt
        if fit < 1:
            raise RuntimeError(f"not enough device memory for one utterance of {T} frames ({per_utt} bytes needed)")
        if fit >= B:
            return B
        n_chunks = -(-B // fit)
        return -(-B // n_chunks)

    # ------------------------------------------------------------------ one evaluation
    def eval(self, ws: LynxWorkspace, x_f32: torch.Tensor, vecs: torch.Tensor, out_f32: torch.Tensor, hoisted: bool = False,
             evecs: torch.Tensor | None = None) -> int:
        """x / out: fp32 [B, T, F*M] time-major; vecs: fp32 [L, Bt, C] step vectors of this evaluation.  hoisted: ws.cproj
        holds W_c cond of every layer (strong_cond = False only).  evecs: ln_vectors(vecs) (folded LayerNorm; computed here when
        the caller has none).  Returns #launches."""
        hoisted = hoisted and not self.strong
        w, B, T, Cc = self.w, ws.B, ws.T, self.C
        rows = B * T
        Bt = vecs.shape[1]
        lens = ws._len_dev if ws.lengths is not None else None
        ops.cast_bf16(ws.xb, x_f32.view(rows, self.FM))
        if self.fold_ln:
            if evecs is None:
                evecs = self.ln_vectors(vecs)
            st = ws.stats
            ops.gemm(ops.EPI_LINEAR, self.bn_c, B, T, ws.xb, w["w_in_c"], ws.u, a1=ws.cond, bias=w["b_in_c"], act=ops.ACT_NONE,
                     stats_out=st, stats_d=vecs[0])
            for i, lw in enumerate(w["layers"]):
                last = i + 1 == self.L
                ops.gemm(ops.EPI_SWIGLU, self.bn_g, B, T, ws.u, lw["w2t"], ws.g, bias=lw["b2t"], lengths=lens,
                         ln=dict(stats=st, channels=Cc, eps=1e-5, s=lw["ln_s"], e=evecs[i]))
                ops.dwconv(ws.dw, ws.g, lw["dw_w"], lw["dw_b"], B, T, self.k, self.act, lw["slope"])
                ops.gemm(ops.EPI_RES, self.bn_c, B, T, ws.dw, lw["w6c"], ws.u, a1=None if last else ws.cond, vec=lw["b6c"], vec_bstride=0,
                         res=ws.u, stats_out=None if last else st, stats_d=None if last else vecs[i + 1])
            ops.layernorm(ws.n, ws.u, w["norm_g"], w["norm_b"], 1e-5)
            ops.gemm(ops.EPI_OUTF32, self.bn_o, B, T, ws.n, w["w_out"], out_f32.view(rows, self.FM), bias=w["b_out"], n_real=self.FM)
            return 3 * self.L + 4
        if self.strong:
            ops.gemm(ops.EPI_LINEAR, self.bn_c, B, T, ws.xb, w["w_in_c"], ws.u, a1=ws.cond, bias=w["b_in_c"], act=ops.ACT_NONE)
            for i, lw in enumerate(w["layers"]):
                ops.layernorm(ws.n, ws.u, lw["ln_g"], lw["ln_b"], 1e-5, vecs[i], T if Bt > 1 else rows)
                ops.gemm(ops.EPI_SWIGLU, self.bn_g, B, T, ws.n, lw["w2"], ws.g, bias=lw["b2"], lengths=lens)
                ops.dwconv(ws.dw, ws.g, lw["dw_w"], lw["dw_b"], B, T, self.k, self.act, lw["slope"])
                ops.gemm(ops.EPI_RES, self.bn_c, B, T, ws.dw, lw["w6c"], ws.u, a1=ws.cond if i + 1 < self.L else None,
                         vec=lw["b6c"], vec_bstride=0, res=ws.u)
            ops.layernorm(ws.n, ws.u, w["norm_g"], w["norm_b"], 1e-5)
            ops.gemm(ops.EPI_OUTF32, self.bn_o, B, T, ws.n, w["w_out"], out_f32.view(rows, self.FM), bias=w["b_out"], n_real=self.FM)
            return 4 * self.L + 4
        ops.gemm(ops.EPI_LINEAR, self.bn_c, B, T, ws.xb, w["w_in"], ws.h, bias=w["b_in"], act=ops.ACT_GELU)
        for i, lw in enumerate(w["layers"]):
            dvec = vecs[i]  # [Bt, C]
            if hoisted:      # res_x = x; LayerNorm(x + cproj + (bias + step vector))
                ops.layernorm(ws.n, ws.h, lw["ln_g"], lw["ln_b"], 1e-5, dvec, T if Bt > 1 else rows, add=ws.cproj[i])
                res = ws.h
            else:            # res_x = x; u = x + cond projection + (bias + step vector)
                ops.gemm(ops.EPI_RES, self.bn_c, B, T, ws.cond, lw["wc"], ws.u, vec=dvec, vec_bstride=Cc if Bt > 1 else 0, res=ws.h)
                res, ln_vec = ws.h, None
            if not hoisted:
                ops.layernorm(ws.n, ws.u, lw["ln_g"], lw["ln_b"], 1e-5, ln_vec, T if Bt > 1 else rows)
            ops.gemm(ops.EPI_SWIGLU, self.bn_g, B, T, ws.n, lw["w2"], ws.g, bias=lw["b2"], lengths=lens)
            ops.dwconv(ws.dw, ws.g, lw["dw_w"], lw["dw_b"], B, T, self.k, self.act, lw["slope"])
            ops.gemm(ops.EPI_RES, self.bn_c, B, T, ws.dw, lw["w6"], ws.h, vec=lw["b6"], vec_bstride=0, res=res)
        ops.layernorm(ws.n, ws.h, w["norm_g"], w["norm_b"], 1e-5)
        ops.gemm(ops.EPI_OUTF32, self.bn_o, B, T, ws.n, w["w_out"], out_f32.view(rows, self.FM), bias=w["b_out"], n_real=self.FM)
        return (4 if hoisted else 5) * self.L + 4

    # ------------------------------------------------------------------ sampler programs
    @torch.no_grad()
    def run_program(self, module, prog: Program, x_tm: torch.Tensor, cond_bht: torch.Tensor, extra: Sequence[torch.Tensor] = (),
                    use_graph: bool = True, lengths=None) -> torch.Tensor:
        """Runs a sampler Program (core/programs.py) with this backbone: EVAL = self.eval, LINCOMB = ds_lincomb, the whole op
        list captured in one CUDA graph per (program, B, T) and replayed on later calls."""
        B, T, FM = x_tm.shape
        bc = self.max_chunk_batch(B, T, prog.n_bufs)
        if bc < B and lengths is not None:
            raise nat.NativeError("a ragged batch must fit the device in one piece")
        if bc < B:  # utterances are independent: a batch that does not fit runs as consecutive chunks (same graph)
            out = torch.empty_like(x_tm)
            launches = 0
            for lo in range(0, B, bc):
                hi = min(B, lo + bc)
                out[lo:hi] = self.run_program(module, prog, x_tm[lo:hi], cond_bht[lo:hi], [e[lo:hi] for e in extra], use_graph)
                launches += self.workspace(hi - lo, T).last_launches
            self.last_launches = launches
            return out
        ws = self.workspace(B, T)
        ws.set_lengths(lengths)  # None: back to B full-length utterances
        bufs = ws.bufs(prog.n_bufs)
        bufs[0].copy_(x_tm)
        for k, e in enumerate(extra):
            bufs[2 + k].copy_(e)
        ws.set_cond(cond_bht.transpose(1, 2))
        hoist = (not self.strong) and prog.nfe > 1
        pre = ws.hoist_cond() if hoist else 0
        vkey = (tuple(prog.times), self.weights_key)
        vecs = self._vec_cache.get(vkey)
        if vecs is None:
            if len(self._vec_cache) > 16:
                self._vec_cache.clear()
            vecs = self.step_vectors(module, torch.tensor(prog.times, dtype=torch.float32), B, False)
            vecs = self._vec_cache[vkey] = (vecs, self.ln_vectors(vecs) if self.fold_ln else None)
        gkey = (id(prog), vkey, ws.lengths is not None)  # ragged launches carry the lengths pointer
        if not use_graph or len(prog.ops) > 1024:
            ws.last_launches = pre + self._run_ops(ws, prog, bufs, vecs, hoist)
            return bufs[0].clone()
        entry = ws.graphs.get(gkey)
        if entry is None:
            if len(ws.graphs) >= 8:  # bounded: every captured graph owns a private activation pool
                ws.graphs.pop(next(iter(ws.graphs)))
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph):
                n_launch = self._run_ops(ws, prog, bufs, vecs, hoist)
            # the entry keeps the Program (its id is the key) and the step vectors (device pointers baked into the graph) alive
            # exactly as long as the graph: clearing ws.graphs frees all three
            entry = ws.graphs[gkey] = (graph, n_launch, prog, vecs)
        entry[0].replay()
        ws.last_launches = pre + entry[1]
        return bufs[0].clone()

    def _run_ops(self, ws: LynxWorkspace, prog: Program, bufs: List[torch.Tensor], vecs: torch.Tensor, hoisted: bool = False) -> int:
        launches = 0
        for kind, dst, time, srcs, coefs in prog.ops:
            if kind == nat.DS_OP_EVAL:
                launches += self.eval(ws, bufs[srcs[0]], vecs[0][time], bufs[dst], hoisted, None if vecs[1] is None else vecs[1][time])
            else:
                ops.lincomb(bufs[dst], [bufs[s] for s in srcs], coefs)
                launches += 1
        return launches
