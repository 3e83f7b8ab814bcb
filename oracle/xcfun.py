"""Closed-form exchange-correlation energy densities -- TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).

The reference's `xc(rho[, grad], id)` arithmetic function (/root/reference/src/arithmetic@proc.F90:1645-1680) calls
libxc (un-vendored third-party dependency, absent from /root/reference; version not pinned by the snapshot: `configure.ac`
only asks for a libxc with the f90 interface) with XC_UNPOLARIZED, rho = max(rho, 1e-14), sigma = |grad rho|^2, and returns
zk * rho, zk = energy per particle.  The four functionals used by tests_libxc/h2o.cri:4-5 are restated here from their
published closed forms with libxc's parameter values:

  id   1  XC_LDA_X          Slater/Dirac exchange            e_x = -3/4 (3/pi)^(1/3) rho^(1/3)
  id   8  XC_LDA_C_VWN_RPA  Vosko, Wilk, Nusair, Can. J. Phys. 58 (1980) 1200: their formula with the RPA fit
                            (paramagnetic: A = 0.0310907, b = 13.0720, c = 42.7198, x0 = -0.409286)
  id 101  XC_GGA_X_PBE      Perdew, Burke, Ernzerhof, PRL 77 (1996) 3865: kappa = 0.8040, mu = beta pi^2 / 3
  id 130  XC_GGA_C_PBE      same paper; beta = 0.06672455060314922, gamma = (1 - ln 2)/pi^2, on the Perdew-Wang 92 LDA
                            with libxc's "modified" (full-digit) PW92 parameters.
"""
from __future__ import annotations

import numpy as np

RHO_FLOOR = 1e-14   # arithmetic@proc.F90:1664,1668


def _rs(rho):
    return (3.0 / (4.0 * np.pi * rho)) ** (1.0 / 3.0)


def lda_x(rho):
    return -0.75 * (3.0 / np.pi) ** (1.0 / 3.0) * rho ** (1.0 / 3.0)


def lda_c_vwn_rpa(rho):
    A, b, c, x0 = 0.0310907, 13.0720, 42.7198, -0.409286
    x = np.sqrt(_rs(rho))
    X = x * x + b * x + c
    X0 = x0 * x0 + b * x0 + c
    Q = np.sqrt(4.0 * c - b * b)
    at = np.arctan(Q / (2.0 * x + b))
    return A * (np.log(x * x / X) + 2.0 * b / Q * at
                - b * x0 / X0 * (np.log((x - x0) ** 2 / X) + 2.0 * (b + 2.0 * x0) / Q * at))


def lda_c_pw_mod(rho):
    """Perdew-Wang 92, unpolarised, libxc's XC_LDA_C_PW_MOD parameters (the LDA inside its PBE correlation)"""
    a, alpha1 = 0.0310906908696548950, 0.21370
    b1, b2, b3, b4 = 7.5957, 3.5876, 1.6382, 0.49294
    rs = _rs(rho)
    den = 2.0 * a * (b1 * np.sqrt(rs) + b2 * rs + b3 * rs ** 1.5 + b4 * rs * rs)
    return -2.0 * a * (1.0 + alpha1 * rs) * np.log1p(1.0 / den)


_PBE_BETA = 0.06672455060314922
_PBE_GAMMA = (1.0 - np.log(2.0)) / np.pi ** 2
_PBE_KAPPA = 0.8040
_PBE_MU = _PBE_BETA * np.pi ** 2 / 3.0       # 0.2195149727645171


def gga_x_pbe(rho, sigma):
    s2 = sigma / (4.0 * (3.0 * np.pi ** 2) ** (2.0 / 3.0) * rho ** (8.0 / 3.0))
    fx = 1.0 + _PBE_KAPPA - _PBE_KAPPA / (1.0 + _PBE_MU * s2 / _PBE_KAPPA)
    return lda_x(rho) * fx


def gga_c_pbe(rho, sigma):
    ec = lda_c_pw_mod(rho)
    kf = (3.0 * np.pi ** 2 * rho) ** (1.0 / 3.0)
    ks2 = 4.0 * kf / np.pi
    t2 = sigma / (4.0 * ks2 * rho * rho)
    A = _PBE_BETA / _PBE_GAMMA / np.expm1(-ec / _PBE_GAMMA)
    At2 = A * t2
    H = _PBE_GAMMA * np.log1p(_PBE_BETA / _PBE_GAMMA * t2 * (1.0 + At2) / (1.0 + At2 + At2 * At2))
    return ec + H


def xc(rho, gmod=None, ident: int = 1):
    """The value `xc(...)` pushes on the arithmetic stack: zk * rho with rho floored at 1e-14 (:1664-1680)."""
    rho = np.maximum(np.asarray(rho, float), RHO_FLOOR)
    if ident == 1:
        zk = lda_x(rho)
    elif ident == 8:
        zk = lda_c_vwn_rpa(rho)
    else:
        sigma = np.asarray(gmod, float) ** 2
        zk = gga_x_pbe(rho, sigma) if ident == 101 else gga_c_pbe(rho, sigma)
        if ident not in (101, 130):
            raise ValueError("functional not restated")
    return zk * rho
