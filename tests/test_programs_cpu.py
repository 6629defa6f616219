"""Host logic (no GPU): the sampler Programs built by diffsinger_b200/core/programs.py, interpreted on
CPU with the oracle backbone, must reproduce the reference's golden outputs.  fp32 tolerance 2e-5 rel-L2
(coefficients are float64-on-host instead of fp32-on-device, so agreement is to round-off, not bitwise)."""
import pytest
import torch

from conftest import load_golden, rel_l2
from diffsinger_b200.core import programs
from oracle.program_interp import run_program_cpu
from oracle.wavenet_oracle import OracleWaveNet
from oracle import samplers_oracle as so

TOL = 2e-5
SPEC_MIN, SPEC_MAX = torch.tensor([-12.0]), torch.tensor([0.0])


def tm(x):
    B, F, M, T = x.shape
    return x.reshape(B, F * M, T).transpose(1, 2).contiguous()


def denorm(x_tm, lo=SPEC_MIN, hi=SPEC_MAX):
    return (x_tm + 1) / 2 * (hi - lo) + lo


@pytest.mark.parametrize("name", ["ddim", "pndm", "dpm_solver_s50", "dpm_solver_s125", "unipc", "ddpm", "ddim_shallow",
                                  "ddim_cosine", "dpm_solver_cosine", "unipc_cosine", "dpm_solver_shallow", "unipc_shallow"])
def test_gaussian_programs(name):
    d, sd, cfg = load_golden("ddpm_" + name)
    M, F, timesteps, k_step, speedup, shallow = [int(v) for v in d["meta"]]
    sch = so.Schedule(so.cosine_betas(timesteps) if name.endswith("cosine") else so.linear_betas(timesteps))
    model = OracleWaveNet(sd, cfg["n_feats"], cfg["dilation_cycle_length"])
    t_max = k_step
    x = d["noise"]
    if "src" in d:
        src = (d["src"] - SPEC_MIN) / (SPEC_MAX - SPEC_MIN) * 2 - 1
        x = so.q_sample(sch, src.transpose(-2, -1)[:, None], t_max - 1, x)
    extra = []
    if name.startswith("ddim"):
        prog = programs.ddim_program(sch.alphas_cumprod, t_max, speedup)
    elif name == "pndm":
        prog = programs.pndm_program(sch.alphas_cumprod, t_max, speedup)
    elif name == "ddpm":
        prog = programs.ddpm_program(sch, t_max)
        extra = [tm(n) for n in d["step_noise"]]
    elif name.startswith("dpm"):
        prog = programs.dpm_solver_pp_program(sch.betas[:t_max], t_max // speedup)
    else:
        prog = programs.unipc_program(sch.betas[:t_max], t_max // speedup)
    y = denorm(run_program_cpu(model, prog, tm(x), d["cond"].transpose(1, 2), F, extra))
    assert rel_l2(y, d["y"]) < TOL
    expect = {"pndm": t_max // speedup + 1}.get(name, t_max // max(speedup, 1))
    assert prog.nfe == expect == model.calls


@pytest.mark.parametrize("name", ["euler", "rk2", "rk4", "rk5", "euler_shallow"])
def test_reflow_programs(name):
    d, sd, cfg = load_golden("reflow_" + name)
    M, F, steps, shallow = [int(v) for v in d["meta"]]
    t_start = float(d["t_start"])
    model = OracleWaveNet(sd, cfg["n_feats"], cfg["dilation_cycle_length"])
    x = d["noise"]
    if shallow:
        src = (d["src"] - SPEC_MIN) / (SPEC_MAX - SPEC_MIN) * 2 - 1
        x = t_start * src.transpose(-2, -1)[:, None] + (1 - t_start) * x
    prog = programs.reflow_program(name.split("_")[0], steps, t_start, 1000.0)
    y = denorm(run_program_cpu(model, prog, tm(x), d["cond"].transpose(1, 2), F))
    assert rel_l2(y, d["y"]) < TOL
    assert prog.nfe == {"euler": 1, "rk2": 2, "rk4": 4, "rk5": 6}[name.split("_")[0]] * steps


def test_lower_order_final_small_k():
    """DPM-Solver++ uses a first-order last step only when K < 10 (dpm_solver_pytorch.py:1198-1201)."""
    from oracle import solvers_oracle as sv
    d, sd, cfg = load_golden("ddpm_dpm_solver_s125")
    sch = so.Schedule(so.linear_betas(1000))
    for K in (2, 5):
        model = OracleWaveNet(sd, cfg["n_feats"], cfg["dilation_cycle_length"])
        cond = d["cond"].transpose(1, 2)
        ref = sv.sample_dpm_solver_pp(model, sch.betas, d["noise"], cond, K)
        prog = programs.dpm_solver_pp_program(sch.betas, K)
        y = run_program_cpu(model, prog, tm(d["noise"]), cond, 1)
        assert rel_l2(y, tm(ref)) < TOL


def test_ancestral_chain_in_blocks():
    """The ancestral sampler runs block by block (GaussianDiffusion.ANCESTRAL_BLOCK steps per program, so that only one block of
    per-step noise is alive at a time): chaining the block programs reproduces the reference's 12-step golden run."""
    d, sd, cfg = load_golden("ddpm_ddpm")
    M, F, timesteps, k_step, speedup, shallow = [int(v) for v in d["meta"]]
    sch = so.Schedule(so.linear_betas(timesteps))
    model = OracleWaveNet(sd, cfg["n_feats"], cfg["dilation_cycle_length"])
    src = (d["src"] - SPEC_MIN) / (SPEC_MAX - SPEC_MIN) * 2 - 1
    x = tm(so.q_sample(sch, src.transpose(-2, -1)[:, None], k_step - 1, d["noise"]))
    noise = [tm(n) for n in d["step_noise"]]
    for hi in range(k_step, 0, -5):
        lo = max(hi - 5, 0)
        prog = programs.ddpm_program(sch, hi, lo)
        assert prog.n_bufs == 2 + hi - lo and prog.nfe == hi - lo
        x = run_program_cpu(model, prog, x, d["cond"].transpose(1, 2), F, noise[k_step - hi:k_step - lo])
    assert rel_l2(denorm(x), d["y"]) < TOL
    assert model.calls == k_step
