"""Host-side owner of the native plan / sessions for one WaveNet module.

PyTorch supplies device memory (packed weights, workspaces, sampler buffers) and the current CUDA
stream; every FLOP of the no-grad path runs in libdsb200.so.
"""
from __future__ import annotations

import ctypes as C
from typing import Dict, List, Sequence, Tuple

import torch

from . import _native as nat


def _ptr(t: torch.Tensor) -> int:
    return t.data_ptr()


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _f32c(t: torch.Tensor) -> torch.Tensor:
    return t.detach().to(dtype=torch.float32).contiguous()


class Program:
    """A sampler flattened into EVAL / LINCOMB ops over numbered fp32 buffers [B, T, F*M].

    times: model-input diffusion-step values, one per distinct EVAL time index (fp32).
    """

    def __init__(self, n_bufs: int):
        self.n_bufs = n_bufs
        self.ops: List[Tuple] = []
        self.times: List[float] = []

    def add_time(self, t: float) -> int:
        self.times.append(float(t))
        return len(self.times) - 1

    def eval(self, dst: int, src: int, time: int):
        self.ops.append((nat.DS_OP_EVAL, dst, time, [src], []))

    def lincomb(self, dst: int, terms: Sequence[Tuple[float, int]]):
        terms = [(float(c), int(s)) for c, s in terms]
        assert 1 <= len(terms) <= 8
        self.ops.append((nat.DS_OP_LINCOMB, dst, 0, [s for _, s in terms], [c for c, _ in terms]))

    @property
    def nfe(self) -> int:
        return sum(1 for o in self.ops if o[0] == nat.DS_OP_EVAL)

    def to_native(self):
        arr = (nat.Op * len(self.ops))()
        for i, (kind, dst, time, srcs, coefs) in enumerate(self.ops):
            arr[i].kind, arr[i].dst, arr[i].time, arr[i].n = kind, dst, time, len(srcs)
            for k, s in enumerate(srcs):
                arr[i].src[k] = s
            for k, c in enumerate(coefs):
                arr[i].coef[k] = c
        return arr


class Session:
    def __init__(self, engine: "WaveNetEngine", B: int, T: int, max_times: int, per_sample: bool):
        L = nat.lib()
        self.engine, self.B, self.T, self.max_times, self.per_sample = engine, B, T, max_times, per_sample
        nbytes = L.ds_session_workspace_bytes(engine.plan, B, T, max_times, int(per_sample))
        if nbytes == 0:
            raise nat.NativeError("ds_session_workspace_bytes returned 0 (bad shape)")
        self.workspace = torch.empty(nbytes + 1024, dtype=torch.uint8, device=engine.device)
        base = (self.workspace.data_ptr() + 1023) // 1024 * 1024
        self.handle = C.c_void_p()
        nat.check(L.ds_session_create(engine.plan, B, T, max_times, int(per_sample), base, nbytes, C.byref(self.handle)))
        self.numel = B * T * engine.FM
        self._cond_key = None
        self._times_key = None
        self._t_dev = None
        self._lengths = None

    def set_cond(self, cond_bth: torch.Tensor):
        """cond [B, T, H] fp32 contiguous."""
        assert cond_bth.shape == (self.B, self.T, self.engine.H), (cond_bth.shape, (self.B, self.T, self.engine.H))
        cond_bth = _f32c(cond_bth)
        self._cond_key = None  # callers that cache by tensor identity (WaveNet._forward_native) set it after this call
        nat.check(nat.lib().ds_session_set_cond(self.handle, _ptr(cond_bth), _stream()))

    def set_lengths(self, lengths: Sequence[int] | None):
        """Ragged batch (include/dsb200.h: ds_session_set_lengths): utterance b has lengths[b] <= T frames; None = all T."""
        if lengths is None:
            if self._lengths is not None:
                nat.check(nat.lib().ds_session_set_lengths(self.handle, None, 0, _stream()))
                self._lengths = None
            return
        lengths = tuple(int(v) for v in lengths)
        if lengths == self._lengths:
            return
        arr = (C.c_int32 * len(lengths))(*lengths)
        nat.check(nat.lib().ds_session_set_lengths(self.handle, arr, len(lengths), _stream()))
        self._lengths = lengths

    def set_times(self, t: torch.Tensor):
        """t: fp32 device tensor [n] (shared over the batch) or [n, B] (per-sample session)."""
        if not t.is_cuda:  # pinned staging: the copy is then legal inside a CUDA-graph capture too (it becomes a memcpy node)
            self._t_host = t.to(dtype=torch.float32).contiguous().pin_memory()
            t = self._t_host.to(device=self.engine.device, non_blocking=True)
        t = t.to(device=self.engine.device, dtype=torch.float32).contiguous()
        n = t.shape[0]
        assert t.numel() == n * (self.B if self.per_sample else 1)
        self._t_dev = t  # keep alive until the kernels have run
        self._times_key = None  # set_program_times re-validates
        nat.check(nat.lib().ds_session_set_times(self.handle, _ptr(t), n, _stream()))

    def set_program_times(self, times: Sequence[float]):
        """Step-embedding tables for a sampler program's time values.  They depend only on (times, weights), so a session
        that runs the same program again (every call of a sampler configuration) skips the upload and the table kernels."""
        key = (tuple(times), self.engine.weights_key)
        if self._times_key == key:
            return
        self.set_times(torch.tensor(list(times), dtype=torch.float32))
        self._times_key = key

    def forward(self, x_btc: torch.Tensor, time_index: int, out: torch.Tensor | None = None) -> torch.Tensor:
        assert x_btc.is_contiguous() and x_btc.dtype == torch.float32 and x_btc.numel() == self.numel
        if out is None:
            out = torch.empty_like(x_btc)
        nat.check(nat.lib().ds_session_forward(self.handle, _ptr(x_btc), time_index, _ptr(out), _stream()))
        return out

    def run(self, program: Program, bufs: Sequence[torch.Tensor], use_graph: bool = True):
        assert len(bufs) == program.n_bufs
        for b in bufs:
            assert b.is_contiguous() and b.dtype == torch.float32 and b.numel() == self.numel and b.is_cuda
        ops = program.to_native()
        ptrs = (C.c_void_p * len(bufs))(*[_ptr(b) for b in bufs])
        nat.check(nat.lib().ds_session_run(self.handle, ops, len(ops), ptrs, len(bufs), int(use_graph), _stream()))

    @property
    def last_launches(self) -> int:
        return int(nat.lib().ds_session_last_launches(self.handle))

    @property
    def path(self) -> str:
        """'fused' | 'small' | 'medium' | 'stack': the kernels this session's residual layers run on (include/dsb200.h:
        ds_session_path)."""
        return ("fused", "small", "medium", "stack")[int(nat.lib().ds_session_path(self.handle))]

    def close(self):
        if self.handle:
            nat.lib().ds_session_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class WaveNetEngine:
    """Plan + packed weights of one WaveNet module on one device; sessions cached per (B, T)."""

    SESSION_MEM_FRACTION = 0.5  # cached session workspaces never hold more than this share of the device memory
    MAX_SESSIONS = 32  # config #1 walks ~20 distinct segment lengths; a session of a few hundred frames is a few MB

    def __init__(self, module, device: torch.device):
        nat.require_device()
        L = nat.lib()
        self.device = device
        self.M, self.F = module.in_dims, module.n_feats
        self.FM = self.M * self.F
        self.C, self.Lyr, self.H = module.num_channels, module.num_layers, module.hidden_size
        self.precision = getattr(module, 'precision', 'bf16')
        if self.precision not in ('bf16', 'fp32'):
            raise ValueError(f"precision must be 'bf16' or 'fp32', got {self.precision!r}")
        cfg = nat.WaveNetConfig(self.M, self.F, self.Lyr, self.C, module.dilation_cycle_length, self.H,
                                1 if self.precision == 'fp32' else 0)
        self.plan = C.c_void_p()
        nat.check(L.ds_plan_create(C.byref(cfg), C.byref(self.plan)))
        self.packed = torch.empty(L.ds_plan_packed_bytes(self.plan) + 256, dtype=torch.uint8, device=device)
        self._packed_base = (self.packed.data_ptr() + 255) // 256 * 256
        self._packed_bytes = L.ds_plan_packed_bytes(self.plan)
        self.sessions: Dict[Tuple, Session] = {}
        self.weights_key = None

    @staticmethod
    def weights_fingerprint(module) -> Tuple:
        return tuple((p.data_ptr(), p._version) for p in module.parameters())

    def pack(self, module):
        sd = {k: _f32c(v) for k, v in module.state_dict().items()}
        L = self.Lyr
        w = nat.WaveNetWeights()
        keep = [sd]

        def arr(fmt):
            a = (C.c_void_p * L)(*[_ptr(sd[fmt.format(i)]) for i in range(L)])
            keep.append(a)
            return C.cast(a, C.POINTER(C.c_void_p))

        w.input_projection_weight = _ptr(sd["input_projection.weight"])
        w.input_projection_bias = _ptr(sd["input_projection.bias"])
        w.mlp0_weight, w.mlp0_bias = _ptr(sd["mlp.0.weight"]), _ptr(sd["mlp.0.bias"])
        w.mlp2_weight, w.mlp2_bias = _ptr(sd["mlp.2.weight"]), _ptr(sd["mlp.2.bias"])
        w.skip_projection_weight = _ptr(sd["skip_projection.weight"])
        w.skip_projection_bias = _ptr(sd["skip_projection.bias"])
        w.output_projection_weight = _ptr(sd["output_projection.weight"])
        w.output_projection_bias = _ptr(sd["output_projection.bias"])
        w.dilated_conv_weight = arr("residual_layers.{}.dilated_conv.weight")
        w.dilated_conv_bias = arr("residual_layers.{}.dilated_conv.bias")
        w.diffusion_projection_weight = arr("residual_layers.{}.diffusion_projection.weight")
        w.diffusion_projection_bias = arr("residual_layers.{}.diffusion_projection.bias")
        w.conditioner_projection_weight = arr("residual_layers.{}.conditioner_projection.weight")
        w.conditioner_projection_bias = arr("residual_layers.{}.conditioner_projection.bias")
        w.layer_output_projection_weight = arr("residual_layers.{}.output_projection.weight")
        w.layer_output_projection_bias = arr("residual_layers.{}.output_projection.bias")
        nat.check(nat.lib().ds_plan_pack_weights(self.plan, C.byref(w), self._packed_base, self._packed_bytes, _stream()))
        torch.cuda.current_stream().synchronize()  # the fp32 staging copies in `keep` may be freed after this
        self.weights_key = self.weights_fingerprint(module)

    def invalidate(self):
        """Forget the packed weights: the next use repacks from the module's current parameters.  The fingerprint
        (data_ptr, _version) does not see writes through `.data` / storage aliases (EMA helpers, weight swaps): call this -- or
        module.invalidate_native() -- after such an update.  load_state_dict does it by itself (post hook)."""
        self.weights_key = None
        for s in self.sessions.values():
            s._times_key = None
            s._cond_key = None

    def session(self, B: int, T: int, max_times: int = 64, per_sample: bool = False) -> Session:
        key = (B, T, per_sample, _stream())  # a session's buffers belong to the stream its kernels run on
        s = self.sessions.get(key)
        if s is not None and s.max_times >= max_times:
            self.sessions[key] = self.sessions.pop(key)  # most recently used last: eviction below is LRU
            return s
        if s is not None:
            s.close()
            del self.sessions[key]
        need = nat.lib().ds_session_workspace_bytes(self.plan, B, T, max(max_times, 64), int(per_sample))
        budget = self.SESSION_MEM_FRACTION * torch.cuda.get_device_properties(self.device).total_memory
        while self.sessions and (len(self.sessions) >= self.MAX_SESSIONS
                                 or need + sum(x.workspace.numel() for x in self.sessions.values()) > budget):
            old = next(iter(self.sessions))  # least recently used
            self.sessions.pop(old).close()
        s = Session(self, B, T, max(max_times, 64), per_sample)
        self.sessions[key] = s
        return s

    def close(self):
        for s in self.sessions.values():
            s.close()
        self.sessions.clear()
        if self.plan:
            nat.lib().ds_plan_destroy(self.plan)
            self.plan = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
