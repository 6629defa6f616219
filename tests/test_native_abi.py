"""CPU-side checks of the C ABI: the library builds/loads without a GPU, exports every symbol that
include/dsb200.h declares, validates arguments, and refuses to compute without a device."""
import ctypes as C
import os
import re

import pytest

from diffsinger_b200 import _native as nat

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    hdr = open(os.path.join(ROOT, "include", "dsb200.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    return sorted(set(re.findall(r"\b(ds_[a-z_0-9]+)\s*\(", hdr)))


def test_library_exports_every_declared_symbol():
    L = nat.lib()
    names = declared_symbols()
    assert len(names) >= 18
    for n in names:
        assert hasattr(L, n), f"{n} declared in dsb200.h but not exported"
    assert sorted(nat.EXPORTS) == names
    assert L.ds_abi_version() == 1


def test_plan_validation_without_device():
    L = nat.lib()
    plan = C.c_void_p()
    bad = nat.WaveNetConfig(128, 1, 20, 100, 4, 256)  # channels not a multiple of 64
    assert L.ds_plan_create(C.byref(bad), C.byref(plan)) == -1
    assert b"multiple of 64" in L.ds_last_error()
    good = nat.WaveNetConfig(128, 1, 20, 512, 4, 256)
    assert L.ds_plan_create(C.byref(good), C.byref(plan)) == 0
    nbytes = L.ds_plan_packed_bytes(plan)
    # bf16 matrices of the 20x512 net are ~95 MB, the second gate pack for the small-problem path
    # 73 MB, plus fp32 step-embedding weights
    assert 160e6 < nbytes < 220e6
    assert L.ds_plan_precision(plan) == 0 and L.ds_plan_layer_mode(plan) == 5
    # fp32-accurate plan: hi/lo halves of every bf16 matrix, unfused path, no second gate pack
    p32 = C.c_void_p()
    assert L.ds_plan_create(C.byref(nat.WaveNetConfig(128, 1, 20, 512, 4, 256, 1)), C.byref(p32)) == 0
    assert L.ds_plan_precision(p32) == 1 and L.ds_plan_layer_mode(p32) == 5  # the fused kernel has a 3 x bf16 split instantiation
    assert 200e6 < L.ds_plan_packed_bytes(p32) < 260e6
    L.ds_plan_destroy(p32)
    assert L.ds_plan_create(C.byref(nat.WaveNetConfig(128, 1, 20, 512, 4, 256, 7)), C.byref(p32)) == -1
    assert b"precision" in L.ds_last_error()
    assert L.ds_session_workspace_bytes(plan, 2, 300, 20, 0) > 2 * 300 * 20 * 512 * 2
    assert L.ds_plan_destroy(plan) == 0


def test_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    import diffsinger_b200 as ds
    ds.hparams.update(hidden_size=64)
    net = ds.WaveNet(16, 1, num_layers=2, num_channels=64, dilation_cycle_length=2).eval()
    with torch.no_grad(), pytest.raises(RuntimeError, match="no CPU inference path"):
        net(torch.zeros(1, 1, 16, 8), torch.zeros(1), torch.zeros(1, 64, 8))
    with pytest.raises(nat.NativeError):
        nat.require_device()
