"""Fit the rational approximations used by the direct-space tile kernel.

  F(t) = (erf(z)/z - 2/sqrt(pi) exp(-t)) / t ,  V(t) = erf(z)/z ,   t = z^2 = (alpha r)^2

so that   erfc(alpha r)/r + 2 alpha/sqrt(pi) exp(-alpha^2 r^2)  =  1/r - alpha^3 r^2 F(t)   (force)
          erfc(alpha r)/r                                       =  1/r - alpha V(t)          (energy)
(the quantities of coulombLennardJones.cc:9-23 in the reference, reorganised so no exp/erfc is needed).
Run:  python tools/fit_ewald_kernels.py   -> prints C arrays pasted into csrc/b200_ewald_coeffs.h
"""
import numpy as np
from scipy.special import erf
import mpmath as mp

def F_exact(t):
    t = np.asarray(t, dtype=np.float64)
    out = np.empty_like(t)
    small = t < 0.5
    # series for small t:  F = 2/sqrt(pi) * sum_{k>=0} (-1)^k t^k (1/(k!(2k+1)) - ... ) -> use mpmath for safety
    for i, tv in enumerate(t):
        z = mp.sqrt(mp.mpf(tv))
        if tv == 0:
            out[i] = float(mp.mpf(4) / (3 * mp.sqrt(mp.pi)))
        else:
            out[i] = float((mp.erf(z) / z - 2 / mp.sqrt(mp.pi) * mp.exp(-tv)) / tv)
    return out

def V_exact(t):
    t = np.asarray(t, dtype=np.float64)
    out = np.empty_like(t)
    for i, tv in enumerate(t):
        z = mp.sqrt(mp.mpf(tv))
        out[i] = float(2 / mp.sqrt(mp.pi)) if tv == 0 else float(mp.erf(z) / z)
    return out

def ratfit(f, T, n, m, iters=30, npts=2000):
    """SK-iteration weighted least squares rational fit p_n/q_m on [0,T], q(0)=1, relative error."""
    k = np.arange(npts)
    x = 0.5 * T * (1 - np.cos(np.pi * (k + 0.5) / npts))
    y = f(x)
    s = x / T  # scaled variable for conditioning
    w = 1.0 / np.abs(y)
    qprev = np.ones_like(x)
    for it in range(iters):
        A = np.zeros((npts, n + 1 + m))
        for i in range(n + 1):
            A[:, i] = s ** i
        for j in range(1, m + 1):
            A[:, n + j] = -y * s ** j
        W = (w / qprev)[:, None]
        coef, *_ = np.linalg.lstsq(A * W, y * W[:, 0], rcond=None)
        p = coef[: n + 1]
        q = np.concatenate([[1.0], coef[n + 1:]])
        qv = sum(q[j] * s ** j for j in range(m + 1))
        pv = sum(p[i] * s ** i for i in range(n + 1))
        err = np.abs(pv / qv - y) / np.abs(y)
        # minimax-ish reweighting
        w = w * (1 + 0.5 * err / err.max())
        qprev = qv
    # unscale
    p = p / T ** np.arange(n + 1)
    q = q / T ** np.arange(m + 1)
    return p, q, err.max()

def eval32(p, q, t):
    t = t.astype(np.float32)
    pn = np.float32(p[-1]) * np.ones_like(t)
    for c in p[-2::-1]:
        pn = pn * t + np.float32(c)
    qn = np.float32(q[-1]) * np.ones_like(t)
    for c in q[-2::-1]:
        qn = qn * t + np.float32(c)
    return pn / qn

if __name__ == "__main__":
    T = 16.0
    tt = np.linspace(0.0, T, 4001)
    for name, f, degs in (("F", F_exact, [(5, 5), (6, 5), (6, 6)]), ("V", V_exact, [(5, 5), (6, 6)])):
        ex = f(tt)
        for (n, m) in degs:
            p, q, e = ratfit(f, T, n, m)
            e32 = np.max(np.abs(eval32(p, q, tt).astype(np.float64) - ex) / np.abs(ex))
            print(f"// {name}: ({n},{m}) max rel err (f64 fit) {e:.3e}, (f32 eval) {e32:.3e}")
            print(f"static const float {name}_NUM_{n}{m}[{n+1}] = {{" + ", ".join(f"{c:.9e}f" for c in p) + "};")
            print(f"static const float {name}_DEN_{n}{m}[{m+1}] = {{" + ", ".join(f"{c:.9e}f" for c in q) + "};")
