"""ctypes binding of libdsb200.so (include/dsb200.h).  No compute happens in Python.

The library is built in-tree by ``build_native()`` (nvcc, sm_100a only).  There is NO fallback:
if the shared object is missing or no B200 is visible, every entry point raises ``NativeError``.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
import sys

_HERE = os.path.dirname(os.path.abspath(__file__))
_ROOT = os.path.dirname(_HERE)
LIB_PATH = os.path.join(_HERE, "libdsb200.so")
SRC = os.path.join(_HERE, "csrc", "api.cu")
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
              "-Xcompiler", "-fPIC", "-shared"]

DS_OP_EVAL, DS_OP_LINCOMB = 0, 1
EXPORTS = [
    "ds_last_error", "ds_abi_version", "ds_device_check", "ds_plan_create", "ds_plan_destroy",
    "ds_plan_packed_bytes", "ds_plan_pack_weights", "ds_session_workspace_bytes", "ds_session_create",
    "ds_session_destroy", "ds_session_set_cond", "ds_session_set_times", "ds_session_forward",
    "ds_session_run", "ds_session_last_launches", "ds_session_path", "ds_plan_stack_capacity", "ds_session_set_lengths", "ds_lincomb", "ds_transpose", "ds_shallow_start", "ds_debug_gemm", "ds_debug_set_trace", "ds_debug_mma_rate",
    "ds_debug_layer", "ds_static_scratch_bytes", "ds_forward_static", "ds_plan_layer_mode", "ds_plan_precision", "ds_gemm", "ds_layernorm", "ds_dwconv", "ds_cast_bf16", "ds_gate_fwd", "ds_gate_bwd", "ds_wgrad", "ds_source_conv", "ds_mean3_lrelu", "ds_conv_small", "ds_attention", "ds_mask_rows", "ds_length_regulate",
]


class NativeError(RuntimeError):
    pass


class WaveNetConfig(C.Structure):
    _fields_ = [(n, C.c_int32) for n in ("in_dims", "n_feats", "num_layers", "num_channels",
                                          "dilation_cycle_length", "hidden_size", "precision")]


_FP = C.c_void_p  # const float*
_FPP = C.POINTER(C.c_void_p)  # const float* const*


class WaveNetWeights(C.Structure):
    _fields_ = [(n, _FP) for n in (
        "input_projection_weight", "input_projection_bias", "mlp0_weight", "mlp0_bias", "mlp2_weight",
        "mlp2_bias", "skip_projection_weight", "skip_projection_bias", "output_projection_weight",
        "output_projection_bias")] + [(n, _FPP) for n in (
        "dilated_conv_weight", "dilated_conv_bias", "diffusion_projection_weight", "diffusion_projection_bias",
        "conditioner_projection_weight", "conditioner_projection_bias", "layer_output_projection_weight",
        "layer_output_projection_bias")]


class Op(C.Structure):
    _fields_ = [("kind", C.c_int32), ("dst", C.c_int32), ("time", C.c_int32), ("n", C.c_int32),
                ("src", C.c_int32 * 8), ("coef", C.c_float * 8)]


class WgradOperands(C.Structure):
    """ds_wgrad_operands (include/dsb200.h)."""
    _fields_ = [("g_bf16", C.c_void_p), ("x_bf16", C.c_void_p), ("g_channels", C.c_int32), ("g_batch_rows", C.c_int32),
                ("x_channels", C.c_int32), ("x_batch_rows", C.c_int32)]


class CurveEmbed(C.Structure):
    """ds_curve_embed (include/dsb200.h)."""
    _fields_ = [("value", C.c_void_p), ("stride_b", C.c_int64), ("stride_t", C.c_int64), ("weight", C.c_void_p), ("bias", C.c_void_p)]


class GemmDesc(C.Structure):
    _fields_ = [
        ("epi", C.c_int32), ("BN", C.c_int32), ("B", C.c_int32), ("T", C.c_int32),
        ("a0", C.c_void_p), ("a0_channels", C.c_int32), ("a0_c0", C.c_int32), ("kb_a0", C.c_int32),
        ("n_taps", C.c_int32), ("dil", C.c_int32),
        ("a1", C.c_void_p), ("a1_channels", C.c_int32), ("kb_a1", C.c_int32),
        ("w", C.c_void_p), ("w_rows", C.c_int32), ("w_k", C.c_int32), ("b_row0", C.c_int32), ("n_tiles", C.c_int32),
        ("bias", C.c_void_p), ("vec_sub", C.c_void_p), ("vec_add", C.c_void_p), ("vec_bstride", C.c_int32),
        ("res_in", C.c_void_p), ("ld_res", C.c_int32),
        ("out", C.c_void_p), ("ld_out", C.c_int32), ("out_c0", C.c_int32), ("n_real", C.c_int32),
        ("relu", C.c_int32), ("scale", C.c_float),
        ("out2", C.c_void_p), ("slope2", C.c_float),
        ("stats_out", C.c_void_p), ("stats_d", C.c_void_p), ("stats_d_bstride", C.c_int32),
        ("ln_stats", C.c_void_p), ("ln_parts", C.c_int32), ("ln_channels", C.c_int32), ("ln_eps", C.c_float),
        ("ln_s", C.c_void_p), ("ln_e", C.c_void_p), ("ln_e_bstride", C.c_int32),
        ("lengths", C.c_void_p),
    ]


def build_native(force: bool = False, verbose: bool = False) -> str:
    """Compile csrc/api.cu into diffsinger_b200/libdsb200.so (nvcc cross-compiles without a GPU)."""
    srcs = [os.path.join(_HERE, "csrc", f) for f in os.listdir(os.path.join(_HERE, "csrc"))]
    srcs.append(os.path.join(_ROOT, "include", "dsb200.h"))
    if not force and os.path.exists(LIB_PATH) and all(os.path.getmtime(LIB_PATH) >= os.path.getmtime(s) for s in srcs):
        return LIB_PATH
    nvcc = os.environ.get("NVCC", "nvcc")
    # several ranks (torchrun) may find the library stale at once: one builds under an exclusive lock, into a temporary file that
    # replaces the library atomically, so nobody ever dlopens a half-written .so; the others re-check after the lock
    import fcntl
    with open(LIB_PATH + ".lock", "w") as lock:
        fcntl.flock(lock, fcntl.LOCK_EX)
        try:
            if not force and os.path.exists(LIB_PATH) and all(os.path.getmtime(LIB_PATH) >= os.path.getmtime(s) for s in srcs):
                return LIB_PATH
            tmp = f"{LIB_PATH}.{os.getpid()}.tmp"
            cmd = [nvcc] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-o", tmp, SRC]
            res = subprocess.run(cmd, capture_output=True, text=True)
            if res.returncode != 0:
                if os.path.exists(tmp):
                    os.remove(tmp)
                raise NativeError("nvcc failed:\n" + res.stdout + res.stderr)
            os.replace(tmp, LIB_PATH)
        finally:
            fcntl.flock(lock, fcntl.LOCK_UN)
    if verbose:
        sys.stderr.write(res.stderr)
    return LIB_PATH


_lib = None


def lib() -> C.CDLL:
    """Loads the library (building it if the sources are newer and nvcc is present)."""
    global _lib
    if _lib is not None:
        return _lib
    try:
        build_native()  # no-op when the in-tree .so is newer than every source
    except (OSError, NativeError) as e:
        if not os.path.exists(LIB_PATH):
            raise NativeError(f"libdsb200.so is missing and could not be built: {e}") from e
    L = C.CDLL(LIB_PATH)
    L.ds_last_error.restype = C.c_char_p
    L.ds_plan_packed_bytes.restype = C.c_size_t
    L.ds_plan_packed_bytes.argtypes = [C.c_void_p]
    L.ds_session_workspace_bytes.restype = C.c_size_t
    L.ds_session_workspace_bytes.argtypes = [C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.c_int32]
    L.ds_session_last_launches.restype = C.c_int64
    L.ds_session_last_launches.argtypes = [C.c_void_p]
    L.ds_shallow_start.argtypes = [C.c_void_p] * 5 + [C.c_float, C.c_float, C.c_int32, C.c_int32, C.c_int32, C.c_void_p]
    L.ds_session_set_lengths.argtypes = [C.c_void_p, C.POINTER(C.c_int32), C.c_int32, C.c_void_p]
    L.ds_plan_stack_capacity.restype = C.c_int
    L.ds_plan_stack_capacity.argtypes = [C.c_void_p]
    L.ds_session_path.restype = C.c_int
    L.ds_session_path.argtypes = [C.c_void_p]
    L.ds_plan_create.argtypes = [C.POINTER(WaveNetConfig), C.POINTER(C.c_void_p)]
    L.ds_plan_destroy.argtypes = [C.c_void_p]
    L.ds_plan_pack_weights.argtypes = [C.c_void_p, C.POINTER(WaveNetWeights), C.c_void_p, C.c_size_t, C.c_void_p]
    L.ds_session_create.argtypes = [C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_void_p, C.c_size_t,
                                    C.POINTER(C.c_void_p)]
    L.ds_session_destroy.argtypes = [C.c_void_p]
    L.ds_session_set_cond.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
    L.ds_session_set_times.argtypes = [C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p]
    L.ds_session_forward.argtypes = [C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p]
    L.ds_session_run.argtypes = [C.c_void_p, C.POINTER(Op), C.c_int32, C.POINTER(C.c_void_p), C.c_int32, C.c_int32,
                                 C.c_void_p]
    L.ds_lincomb.argtypes = [C.c_void_p, C.POINTER(C.c_void_p), C.POINTER(C.c_float), C.c_int32, C.c_size_t, C.c_void_p]
    L.ds_transpose.argtypes = [C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.c_void_p]
    L.ds_debug_gemm.argtypes = [C.POINTER(GemmDesc), C.c_void_p]
    L.ds_debug_set_trace.argtypes = [C.c_void_p, C.c_int32]
    L.ds_debug_mma_rate.argtypes = [C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_void_p, C.c_void_p]
    L.ds_debug_layer.argtypes = [C.c_void_p, C.c_int32, C.c_int32, C.c_void_p]
    L.ds_plan_layer_mode.argtypes = [C.c_void_p]
    L.ds_plan_precision.argtypes = [C.c_void_p]
    L.ds_gemm.argtypes = [C.POINTER(GemmDesc), C.c_void_p]
    L.ds_gate_fwd.argtypes = [C.c_void_p, C.c_int64, C.c_void_p, C.c_int64, C.c_int64, C.c_int32, C.c_void_p]
    L.ds_gate_bwd.argtypes = [C.c_void_p, C.c_int64, C.c_void_p, C.c_int64, C.c_void_p, C.c_int64, C.c_int64, C.c_int32, C.c_void_p]
    L.ds_wgrad.argtypes = [C.POINTER(WgradOperands), C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_void_p, C.c_int32, C.c_void_p,
                           C.c_void_p]
    L.ds_source_conv.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.c_int64, C.c_int64, C.c_int32, C.c_int32,
                                 C.c_int32, C.c_int32, C.c_void_p]
    L.ds_conv_small.argtypes = [C.c_void_p, C.c_int32, C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32,
                                C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_float, C.c_void_p]
    L.ds_mean3_lrelu.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_float, C.c_void_p]
    L.ds_layernorm.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int32, C.c_float, C.c_void_p,
                               C.c_int64, C.c_void_p, C.c_void_p]
    L.ds_dwconv.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.c_int32,
                            C.c_int32, C.c_int32, C.c_void_p]
    L.ds_cast_bf16.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_int32, C.c_int32, C.c_void_p]
    L.ds_attention.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_float,
                               C.c_void_p]
    L.ds_mask_rows.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_int32, C.c_void_p]
    L.ds_length_regulate.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.POINTER(CurveEmbed),
                                     C.c_int32, C.c_void_p, C.c_int64, C.c_int64, C.c_void_p]
    L.ds_static_scratch_bytes.restype = C.c_size_t
    L.ds_static_scratch_bytes.argtypes = []
    L.ds_forward_static.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p, C.c_int32, C.c_int32,
                                    C.c_void_p]
    _lib = L
    return L


def check(status: int) -> None:
    if status != 0:
        raise NativeError(f"dsb200 error {status}: {lib().ds_last_error().decode()}")


def require_device() -> None:
    """Fails loudly when the CUDA path cannot run (no silent CPU fallback anywhere)."""
    check(lib().ds_device_check())
