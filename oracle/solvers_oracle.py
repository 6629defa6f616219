"""TEST INFRASTRUCTURE ONLY -- CPU restatement (torch fp32) of the reference's higher-order paths.

  * DPM-Solver++ (2M, multistep, time_uniform, discrete schedule): the only branch
    ``ddpm.py:246-284`` reaches in ``inference/dpm_solver_pytorch.py``
    (NoiseScheduleVP :98-167, model_wrapper :271-289, data_prediction_fn :433-442,
    first_update :569-580, second_update :812-837, sample :1171-1213, interpolate_fn :1253-1292).
  * UniPC (bh2, data prediction, order 2): ``ddpm.py:285-322`` ->
    ``inference/uni_pc.py`` (NoiseScheduleVP :77-86 -- no lambda clip, bh_update :471-588,
    sample :606-658).
  * Rectified-flow ODE samplers: ``modules/core/reflow.py:66-138``.
"""
from __future__ import annotations

import torch


def interp(x: torch.Tensor, xp: torch.Tensor, yp: torch.Tensor) -> torch.Tensor:
    """Piecewise-linear f(x) through (xp, yp), linearly extrapolated from the outermost segment
    (interpolate_fn, dpm_solver_pytorch.py:1253-1292).  x scalar tensor; xp, yp 1-D ascending xp."""
    K = xp.shape[0]
    idx = int(torch.searchsorted(xp, x.reshape(1), right=False).item())  # #keypoints < x
    # the reference sorts [x, xp...] (stable) so ties put x first: same as 'left' search
    lo = min(max(idx - 1, 0), K - 2)
    x0, x1, y0, y1 = xp[lo], xp[lo + 1], yp[lo], yp[lo + 1]
    return y0 + (x - x0) * (y1 - y0) / (x1 - x0)


class VPSchedule:
    def __init__(self, betas: torch.Tensor, clip: bool):
        log_alphas = 0.5 * torch.log(1 - betas).cumsum(dim=0)
        if clip:  # dpm_solver_pytorch.py:114-125 (trap T9)
            log_sigmas = 0.5 * torch.log(1.0 - torch.exp(2.0 * log_alphas))
            lambs = log_alphas - log_sigmas
            idx = int(torch.searchsorted(torch.flip(lambs, [0]), torch.tensor(-5.1)))
            if idx > 0:
                log_alphas = log_alphas[:-idx]
        self.log_alpha = log_alphas.float()
        self.N = self.log_alpha.shape[0]
        self.t_array = torch.linspace(0.0, 1.0, self.N + 1)[1:].float()

    def log_mean(self, t):
        return interp(t, self.t_array, self.log_alpha)

    def alpha(self, t):
        return torch.exp(self.log_mean(t))

    def std(self, t):
        return torch.sqrt(1.0 - torch.exp(2.0 * self.log_mean(t)))

    def lam(self, t):
        lm = self.log_mean(t)
        return lm - 0.5 * torch.log(1.0 - torch.exp(2.0 * lm))


def _x0_fn(model, ns: VPSchedule, cond):
    def f(x, t):
        t_in = ((t - 1.0 / ns.N) * ns.N).expand(x.shape[0])
        eps = model(x, t_in, cond)
        return (x - ns.std(t) * eps) / ns.alpha(t)
    return f


@torch.no_grad()
def sample_dpm_solver_pp(model, betas: torch.Tensor, x, cond, steps: int):
    ns = VPSchedule(betas, clip=True)
    x0_fn = _x0_fn(model, ns, cond)
    ts = torch.linspace(1.0, 1.0 / ns.N, steps + 1)
    assert steps >= 2
    t_prev = [ts[0]]
    m_prev = [x0_fn(x, ts[0])]

    def first(x, s, t, m_s):
        h = ns.lam(t) - ns.lam(s)
        return ns.std(t) / ns.std(s) * x - ns.alpha(t) * torch.expm1(-h) * m_s

    def second(x, t):
        m1, m0 = m_prev[-2], m_prev[-1]
        l1, l0, lt = ns.lam(t_prev[-2]), ns.lam(t_prev[-1]), ns.lam(t)
        h0, h = l0 - l1, lt - l0
        r0 = h0 / h
        D1 = (1.0 / r0) * (m0 - m1)
        phi = torch.expm1(-h)
        a_t = ns.alpha(t)
        return (ns.std(t) / ns.std(t_prev[-1])) * x - (a_t * phi) * m0 - 0.5 * (a_t * phi) * D1

    x = first(x, t_prev[-1], ts[1], m_prev[-1])
    t_prev.append(ts[1])
    m_prev.append(x0_fn(x, ts[1]))
    for step in range(2, steps + 1):
        t = ts[step]
        order = min(2, steps + 1 - step) if steps < 10 else 2
        x = second(x, t) if order == 2 else first(x, t_prev[-1], t, m_prev[-1])
        t_prev = [t_prev[1], t]
        m_prev = [m_prev[1], m_prev[1]]
        if step < steps:
            m_prev[-1] = x0_fn(x, t)
    return x


@torch.no_grad()
def sample_unipc(model, betas: torch.Tensor, x, cond, steps: int):
    ns = VPSchedule(betas, clip=False)
    x0_fn = _x0_fn(model, ns, cond)
    ts = torch.linspace(1.0, 1.0 / ns.N, steps + 1)
    assert steps >= 2

    def update(x, m_list, t_list, t, order, use_corrector):
        t0, m0 = t_list[-1], m_list[-1]
        l0, lt = ns.lam(t0), ns.lam(t)
        h = lt - l0
        rks, D1s = [], []
        for i in range(1, order):
            rk = (ns.lam(t_list[-(i + 1)]) - l0) / h
            rks.append(rk)
            D1s.append((m_list[-(i + 1)] - m0) / rk)
        rks.append(torch.tensor(1.0))
        rks = torch.stack([r.reshape(()) for r in rks])
        hh = -h
        h_phi_1 = torch.expm1(hh)
        h_phi_k = h_phi_1 / hh - 1
        B_h = torch.expm1(hh)
        R, b, fact = [], [], 1
        for i in range(1, order + 1):
            R.append(torch.pow(rks, i - 1))
            b.append((h_phi_k * fact / B_h).reshape(1))
            fact *= (i + 1)
            h_phi_k = h_phi_k / hh - 1 / fact
        R, b = torch.stack(R), torch.cat(b)
        a_t = ns.alpha(t)
        x_t_ = ns.std(t) / ns.std(t0) * x - a_t * h_phi_1 * m0
        pred = 0.5 * D1s[0] if len(D1s) > 0 else 0  # order 2 shortcut rhos_p = [0.5]
        x_t = x_t_ - a_t * B_h * pred
        m_t = None
        if use_corrector:
            rhos_c = torch.tensor([0.5]) if order == 1 else torch.linalg.solve(R, b)
            m_t = x0_fn(x_t, t)
            corr = rhos_c[0] * D1s[0] if len(D1s) > 0 else 0
            x_t = x_t_ - a_t * B_h * (corr + rhos_c[-1] * (m_t - m0))
        return x_t, m_t

    t_list, m_list = [ts[0]], [x0_fn(x, ts[0])]
    x, m = update(x, m_list, t_list, ts[1], 1, True)
    t_list.append(ts[1])
    m_list.append(m)
    for step in range(2, steps + 1):
        t = ts[step]
        order = min(2, steps + 1 - step)
        x, m = update(x, m_list, t_list, t, order, step != steps)
        t_list = [t_list[1], t]
        m_list = [m_list[1], m_list[1]]
        if step < steps:
            m_list[-1] = m if m is not None else x0_fn(x, t)
    return x


@torch.no_grad()
def sample_reflow(model, x, cond, steps: int, algorithm: str, t_start: float = 0.0, tsf: float = 1000.0):
    """reflow.py:66-102 + loop :123-136.  x is updated out of place here (same values)."""
    dt = (1.0 - t_start) / max(1, steps)
    dts = torch.tensor([dt]).to(x)
    v = lambda x_, t_: model(x_, tsf * t_, cond)
    for i in range(steps):
        t = t_start + i * dts
        if algorithm == "euler":
            x = x + v(x, t) * dt
        elif algorithm == "rk2":
            k1 = v(x, t)
            k2 = v(x + 0.5 * k1 * dt, t + 0.5 * dt)
            x = x + k2 * dt
        elif algorithm == "rk4":
            k1 = v(x, t)
            k2 = v(x + 0.5 * k1 * dt, t + 0.5 * dt)
            k3 = v(x + 0.5 * k2 * dt, t + 0.5 * dt)
            k4 = v(x + k3 * dt, t + dt)
            x = x + (k1 + 2 * k2 + 2 * k3 + k4) * dt / 6
        elif algorithm == "rk5":
            k1 = v(x, t)
            k2 = v(x + 0.25 * k1 * dt, t + 0.25 * dt)
            k3 = v(x + 0.125 * (k2 + k1) * dt, t + 0.25 * dt)
            k4 = v(x + 0.5 * (-k2 + 2 * k3) * dt, t + 0.5 * dt)
            k5 = v(x + 0.0625 * (3 * k1 + 9 * k4) * dt, t + 0.75 * dt)
            k6 = v(x + (-3 * k1 + 2 * k2 + 12 * k3 - 12 * k4 + 8 * k5) * dt / 7, t + dt)
            x = x + (7 * k1 + 32 * k3 + 12 * k4 + 32 * k5 + 7 * k6) * dt / 90
        else:
            raise ValueError(f"Unsupported algorithm for Rectified Flow: {algorithm}.")
    return x
