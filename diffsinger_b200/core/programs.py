"""Host-side sampler "compilers": each reference sampling loop becomes a flat Program of
EVAL (one backbone evaluation) and LINCOMB (dst = sum coef_i * buf_i) ops.

Every coefficient in the reference's loops is a scalar that depends only on (schedule, t_max, K,
step index) -- the reference recomputes them on the device with hundreds of tiny ops per step
(SURVEY.md section 2.2, appendix C); here they are computed once on the host in float64 from the same
fp32 schedule buffers and baked into the launch arguments.  Model-input time values are produced
with the reference's own fp32 arithmetic (torch CPU ops) because the step embedding sees them.

Buffer 0 is always the sampler state x (normalised domain, time-major [B, T, F*M]).
"""
from __future__ import annotations

import math
from typing import List

import numpy as np
import torch

from ..engine import Program

X, E = 0, 1  # state, model output


# ---------------------------------------------------------------------------------- DDPM family
def ddim_program(alphas_cumprod: torch.Tensor, t_max: int, interval: int) -> Program:
    """p_sample_ddim, modules/core/ddpm.py:159-167, loop :334-343."""
    ac = alphas_cumprod.detach().cpu().double().numpy()
    p = Program(2)
    for i in reversed(range(0, t_max, interval)):
        a_t, a_p = ac[i], ac[max(i - interval, 0)]
        ti = p.add_time(i)
        p.eval(E, X, ti)
        c_x = math.sqrt(a_p) / math.sqrt(a_t)
        c_e = math.sqrt(a_p) * (math.sqrt((1 - a_p) / a_p) - math.sqrt((1 - a_t) / a_t))
        p.lincomb(X, [(c_x, X), (c_e, E)])
    return p


def pndm_program(alphas_cumprod: torch.Tensor, t_max: int, interval: int) -> Program:
    """p_sample_plms, ddpm.py:170-204, loop :323-333.  Buffers: x, 4 rotating eps slots, x_pred, eps'."""
    ac = alphas_cumprod.detach().cpu().double().numpy()
    H0, XP, E2 = 1, 5, 6  # eps history slots are buffers 1..4
    p = Program(7)
    hist: List[int] = []  # buffer ids, most recent last

    def ab(i):  # x_pred = A*x + Bc*e  (get_x_pred, :176-185)
        a_t, a_p = ac[i], ac[max(i - interval, 0)]
        st, sp = math.sqrt(a_t), math.sqrt(a_p)
        A = 1.0 + (a_p - a_t) / (st * (st + sp))
        Bc = -(a_p - a_t) / (st * (math.sqrt((1 - a_p) * a_t) + math.sqrt((1 - a_t) * a_p)))
        return A, Bc

    for i in reversed(range(0, t_max, interval)):
        free = [b for b in range(H0, H0 + 4) if b not in hist[-3:]]
        cur = free[0]
        ti = p.add_time(i)
        p.eval(cur, X, ti)
        A, Bc = ab(i)
        if len(hist) == 0:
            p.lincomb(XP, [(A, X), (Bc, cur)])
            tj = p.add_time(max(i - interval, 0))
            p.eval(E2, XP, tj)
            p.lincomb(X, [(A, X), (Bc * 0.5, cur), (Bc * 0.5, E2)])
        elif len(hist) == 1:
            p.lincomb(X, [(A, X), (Bc * 1.5, cur), (Bc * -0.5, hist[-1])])
        elif len(hist) == 2:
            p.lincomb(X, [(A, X), (Bc * 23 / 12, cur), (Bc * -16 / 12, hist[-1]), (Bc * 5 / 12, hist[-2])])
        else:
            p.lincomb(X, [(A, X), (Bc * 55 / 24, cur), (Bc * -59 / 24, hist[-1]), (Bc * 37 / 24, hist[-2]),
                          (Bc * -9 / 24, hist[-3])])
        hist.append(cur)
        hist = hist[-4:]
    return p


def ddpm_program(sched, t_hi: int, t_lo: int = 0) -> Program:
    """Ancestral p_sample, ddpm.py:123-156, loop :346-349, for the timesteps t_hi-1 ... t_lo (a block of the full loop: the
    sampler runs long ancestral chains block by block so that only one block of per-step noise is alive at a time).  Buffers 2..
    hold the per-step noise (buffer 2+k is the randn drawn at the block's iteration k; t = 0 uses none)."""
    f = lambda name: getattr(sched, name).detach().cpu().double().numpy()
    c1, c2 = f('sqrt_recip_alphas_cumprod'), f('sqrt_recipm1_alphas_cumprod')
    pc1, pc2, plv = f('posterior_mean_coef1'), f('posterior_mean_coef2'), f('posterior_log_variance_clipped')
    p = Program(2 + (t_hi - t_lo))
    for k, i in enumerate(reversed(range(t_lo, t_hi))):
        ti = p.add_time(i)
        p.eval(E, X, ti)
        terms = [(pc1[i] * c1[i] + pc2[i], X), (-pc1[i] * c2[i], E)]
        if i != 0:
            terms.append((math.exp(0.5 * plv[i]), 2 + k))
        p.lincomb(X, terms)
    return p


# ---------------------------------------------------------------------------------- VP schedule helpers
class VPTable:
    """NoiseScheduleVP('discrete', betas=...) -- inference/dpm_solver_pytorch.py:98-167 (clip=True,
    lambda-clip at -5.1 :114-125) and inference/uni_pc.py:77-152 (clip=False)."""

    def __init__(self, betas: torch.Tensor, clip: bool):
        b = betas.detach().cpu().float()
        log_alphas = 0.5 * torch.log(1 - b).cumsum(dim=0)  # fp32 like the reference
        if clip:
            log_sigmas = 0.5 * torch.log(1.0 - torch.exp(2.0 * log_alphas))
            lambs = log_alphas - log_sigmas
            idx = int(torch.searchsorted(torch.flip(lambs, [0]), torch.tensor(-5.1)))
            if idx > 0:
                log_alphas = log_alphas[:-idx]
        self.N = int(log_alphas.shape[0])
        self.log_alpha = log_alphas.double().numpy()
        self.t_array = torch.linspace(0.0, 1.0, self.N + 1)[1:].double().numpy()

    def log_mean(self, t: float) -> float:
        """interpolate_fn (dpm_solver_pytorch.py:1253-1292): piecewise linear, outermost segments extrapolate."""
        xp, yp = self.t_array, self.log_alpha
        lo = int(np.searchsorted(xp, t, side='left')) - 1
        lo = min(max(lo, 0), self.N - 2)
        return float(yp[lo] + (t - xp[lo]) * (yp[lo + 1] - yp[lo]) / (xp[lo + 1] - xp[lo]))

    def alpha(self, t):
        return math.exp(self.log_mean(t))

    def sigma(self, t):
        return math.sqrt(1.0 - math.exp(2.0 * self.log_mean(t)))

    def lam(self, t):
        lm = self.log_mean(t)
        return lm - 0.5 * math.log(1.0 - math.exp(2.0 * lm))

    def timesteps(self, steps: int) -> torch.Tensor:
        """get_time_steps('time_uniform'): linspace(T=1, 1/N, steps+1), fp32 (:474)."""
        return torch.linspace(1.0, 1.0 / self.N, steps + 1)

    def model_time(self, t: torch.Tensor) -> float:
        """model_wrapper.get_model_input_time (:271-280): (t - 1/N) * N in fp32."""
        return float((t - 1.0 / self.N) * self.N)


def dpm_solver_pp_program(betas: torch.Tensor, steps: int) -> Program:
    """DPM-Solver++(2M): DPM_Solver.sample(method='multistep', order=2, skip_type='time_uniform')
    (dpm_solver_pytorch.py:1171-1213) with data prediction (:433-442), first update (:569-580) and
    second update (:812-837).  Buffers: x, eps, and two rotating x0 slots."""
    ns = VPTable(betas, clip=True)
    assert steps >= 2
    ts32 = ns.timesteps(steps)
    ts = [float(v) for v in ts32]
    p = Program(4)
    M = [2, 3]  # M[-1] most recent x0 prediction

    def x0_into(slot, k):  # x0 = (x - sigma_t * eps) / alpha_t
        ti = p.add_time(ns.model_time(ts32[k]))
        p.eval(E, X, ti)
        a, s = ns.alpha(ts[k]), ns.sigma(ts[k])
        p.lincomb(slot, [(1.0 / a, X), (-s / a, E)])

    def first(s_, t_, m0):
        h = ns.lam(t_) - ns.lam(s_)
        p.lincomb(X, [(ns.sigma(t_) / ns.sigma(s_), X), (-ns.alpha(t_) * math.expm1(-h), m0)])

    def second(t1, t0, t_, m1, m0):
        l1, l0, lt = ns.lam(t1), ns.lam(t0), ns.lam(t_)
        h0, h = l0 - l1, lt - l0
        r0 = h0 / h
        ap = ns.alpha(t_) * math.expm1(-h)
        p.lincomb(X, [(ns.sigma(t_) / ns.sigma(t0), X), (-ap * (1.0 + 0.5 / r0), m0), (0.5 * ap / r0, m1)])

    x0_into(M[0], 0)
    first(ts[0], ts[1], M[0])
    x0_into(M[1], 1)
    t_prev = [ts[0], ts[1]]
    for step in range(2, steps + 1):
        order = min(2, steps + 1 - step) if steps < 10 else 2  # lower_order_final only when steps < 10
        if order == 2:
            second(t_prev[0], t_prev[1], ts[step], M[0], M[1])
        else:
            first(t_prev[1], ts[step], M[1])
        t_prev = [t_prev[1], ts[step]]
        M = [M[1], M[0]]  # oldest slot becomes the target of the next x0
        if step < steps:
            x0_into(M[1], step)
        else:
            M = [M[0], M[0]]
    return p


def unipc_program(betas: torch.Tensor, steps: int) -> Program:
    """UniPC bh2 / data prediction / multistep order 2: UniPC.sample (uni_pc.py:606-658) +
    multistep_uni_pc_bh_update (:471-588).  Buffers: x, eps, x_pred, three rotating x0 slots."""
    ns = VPTable(betas, clip=False)
    assert steps >= 2
    ts32 = ns.timesteps(steps)
    ts = [float(v) for v in ts32]
    XP = 2
    slots = [3, 4, 5]
    p = Program(6)

    def x0_into(slot, src, k):
        ti = p.add_time(ns.model_time(ts32[k]))
        p.eval(E, src, ti)
        a, s = ns.alpha(ts[k]), ns.sigma(ts[k])
        p.lincomb(slot, [(1.0 / a, src), (-s / a, E)])

    m_prev = [slots[0]]
    t_prev = [ts[0]]
    x0_into(m_prev[0], X, 0)
    for step in range(1, steps + 1):
        t = ts[step]
        order = 1 if step == 1 else min(2, steps + 1 - step)
        use_corrector = step != steps
        m0, t0 = m_prev[-1], t_prev[-1]
        l0, lt = ns.lam(t0), ns.lam(t)
        h = lt - l0
        hh = -h
        h_phi_1 = math.expm1(hh)
        B_h = math.expm1(hh)
        a_t = ns.alpha(t)
        cx, cm0 = ns.sigma(t) / ns.sigma(t0), -a_t * h_phi_1  # x_t_ = cx*x + cm0*m0
        if order == 2:
            m1 = m_prev[-2]
            rk = (ns.lam(t_prev[-2]) - l0) / h
            # predictor: x_t = x_t_ - a_t*B_h*0.5*(m1 - m0)/rk
            cp = -a_t * B_h * 0.5 / rk
            pred_terms = [(cx, X), (cm0 - cp, m0), (cp, m1)]
        else:
            pred_terms = [(cx, X), (cm0, m0)]
        if not use_corrector:
            p.lincomb(X, pred_terms)
            break
        p.lincomb(XP, pred_terms)
        mt = [s for s in slots if s not in m_prev[-2:]][0]
        x0_into(mt, XP, step)
        if order == 1:
            rho_t = 0.5
            terms = [(cx, X), (cm0 + a_t * B_h * rho_t, m0), (-a_t * B_h * rho_t, mt)]
        else:
            h_phi_k = h_phi_1 / hh - 1.0
            b1 = h_phi_k * 1.0 / B_h
            h_phi_k = h_phi_k / hh - 1.0 / 2.0
            b2 = h_phi_k * 2.0 / B_h
            R = np.array([[1.0, 1.0], [rk, 1.0]])
            rho = np.linalg.solve(R, np.array([b1, b2]))
            c1 = -a_t * B_h * rho[0] / rk  # * (m1 - m0)
            c2 = -a_t * B_h * rho[1]       # * (mt - m0)
            terms = [(cx, X), (cm0 - c1 - c2, m0), (c1, m_prev[-2]), (c2, mt)]
        p.lincomb(X, terms)
        m_prev = (m_prev + [mt])[-2:]
        t_prev = (t_prev + [t])[-2:]
    return p


# ---------------------------------------------------------------------------------- rectified flow
def reflow_program(algorithm: str, steps: int, t_start: float, time_scale_factor: float) -> Program:
    """sample_euler / rk2 / rk4 / rk5 (modules/core/reflow.py:66-102) inside the loop of :123-136."""
    dt = (1.0 - t_start) / max(1, steps)
    dts = torch.tensor([dt], dtype=torch.float32)
    tsf = time_scale_factor

    def mt(i, frac):  # time_scale_factor * (t + frac*dt) with t = t_start + i*dts (fp32 tensor), as the reference
        t = t_start + i * dts
        if frac != 0.0:
            t = t + frac * dt
        return float((tsf * t)[0])

    if algorithm == 'euler':
        p = Program(2)
        for i in range(steps):
            p.eval(E, X, p.add_time(mt(i, 0.0)))
            p.lincomb(X, [(1.0, X), (dt, E)])
    elif algorithm == 'rk2':
        K1, K2, XS = 1, 2, 3
        p = Program(4)
        for i in range(steps):
            p.eval(K1, X, p.add_time(mt(i, 0.0)))
            p.lincomb(XS, [(1.0, X), (0.5 * dt, K1)])
            p.eval(K2, XS, p.add_time(mt(i, 0.5)))
            p.lincomb(X, [(1.0, X), (dt, K2)])
    elif algorithm == 'rk4':
        K1, K2, K3, K4, XS = 1, 2, 3, 4, 5
        p = Program(6)
        for i in range(steps):
            p.eval(K1, X, p.add_time(mt(i, 0.0)))
            p.lincomb(XS, [(1.0, X), (0.5 * dt, K1)])
            p.eval(K2, XS, p.add_time(mt(i, 0.5)))
            p.lincomb(XS, [(1.0, X), (0.5 * dt, K2)])
            p.eval(K3, XS, p.add_time(mt(i, 0.5)))
            p.lincomb(XS, [(1.0, X), (dt, K3)])
            p.eval(K4, XS, p.add_time(mt(i, 1.0)))
            p.lincomb(X, [(1.0, X), (dt / 6, K1), (dt / 3, K2), (dt / 3, K3), (dt / 6, K4)])
    elif algorithm == 'rk5':
        K1, K2, K3, K4, K5, K6, XS = 1, 2, 3, 4, 5, 6, 7
        p = Program(8)
        for i in range(steps):
            p.eval(K1, X, p.add_time(mt(i, 0.0)))
            p.lincomb(XS, [(1.0, X), (0.25 * dt, K1)])
            p.eval(K2, XS, p.add_time(mt(i, 0.25)))
            p.lincomb(XS, [(1.0, X), (0.125 * dt, K2), (0.125 * dt, K1)])
            p.eval(K3, XS, p.add_time(mt(i, 0.25)))
            p.lincomb(XS, [(1.0, X), (-0.5 * dt, K2), (1.0 * dt, K3)])
            p.eval(K4, XS, p.add_time(mt(i, 0.5)))
            p.lincomb(XS, [(1.0, X), (0.1875 * dt, K1), (0.5625 * dt, K4)])
            p.eval(K5, XS, p.add_time(mt(i, 0.75)))
            p.lincomb(XS, [(1.0, X), (-3 * dt / 7, K1), (2 * dt / 7, K2), (12 * dt / 7, K3), (-12 * dt / 7, K4),
                           (8 * dt / 7, K5)])
            p.eval(K6, XS, p.add_time(mt(i, 1.0)))
            p.lincomb(X, [(1.0, X), (7 * dt / 90, K1), (32 * dt / 90, K3), (12 * dt / 90, K4), (32 * dt / 90, K5),
                          (7 * dt / 90, K6)])
    else:
        raise ValueError(f'Unsupported algorithm for Rectified Flow: {algorithm}.')
    return p
