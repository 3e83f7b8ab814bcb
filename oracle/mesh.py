"""Molecular integration mesh of the reference (MOLCALC's default) -- TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).

Restates, in numpy, /root/reference/src/meshmod@proc.f90:
  * genmesh_franchini :231-398  atom-centred radial x Lebedev product mesh with Franchini's partition weights
                                (J. Comput. Chem. 34 (2013) 1819): w = vp0 / max(vp0, sum_atoms vp) * wrad * wang,
                                vp(r) = fscal * exp(-2r) / max(r,1e-10)^3, fscal = 0.3 for hydrogen, 1 otherwise;
  * rmesh_franchini   :531-563  Gauss-Legendre nodes q in (-1,1) mapped by r = zeta/ln2 (1+q) ln(2/(1-q));
  * z2nr / z2nang     :569-622  node counts per quality level (the default is mesh_level_good = 3,
                                global@proc.f90:120-121: nr = ceiling(nr0 * 3.42), nang = 302);
  * gauleg            tools_math@proc.f90:1289-1350  Newton iteration on the Legendre recurrence, eps = 3e-16.
The Lebedev-Laikov rules themselves (tools_math's lebedev submodule) are NOT in the reference snapshot; they are the
published rules (Lebedev & Laikov, Doklady Mathematics 59 (1999) 477), taken here from scipy.integrate.lebedev_rule, whose
weights sum to 4 pi like the reference's (genmesh divides them by fourpi, :326).  The order of the angular nodes inside a
shell may differ from the reference's table: it changes the summation order only.

This mesh pins the oracle's rho and |grad rho| (and the GPU point-list path) against the one reference-held numeric
vector that goes through rho2: /root/reference/tests_libxc/ref.txt (MOLCALC integrals of LDA / PBE energy densities of
tests_libxc/h2o.wfx).  See tests/test_oracle.py::test_libxc_reference_energies_pin_rho2 and
tests/test_gpu_parity.py::test_libxc_reference_energies_through_the_point_list_path.
"""
from __future__ import annotations

import numpy as np

MESH_LEVEL_SMALL, MESH_LEVEL_NORMAL, MESH_LEVEL_GOOD, MESH_LEVEL_VGOOD, MESH_LEVEL_AMAZING = 1, 2, 3, 4, 5

# meshmod@proc.f90:541-552.  The reference writes the table with default-real literals inside a real*8 array constructor,
# so every entry carries single precision (0.9 -> 0.89999997615814209): reproduced with float32.
_ZETA = np.array([
    0.8, 0.9, 1.8, 1.4, 1.3, 1.1, 0.9, 0.9, 0.9, 0.9,
    1.4, 1.3, 1.3, 1.2, 1.1, 1.0, 1.0, 1.0, 1.5, 1.4,
    1.3, 1.2, 1.2, 1.2, 1.2, 1.2, 1.2, 1.1, 1.1, 1.1,
    1.1, 1.0, 0.9, 0.9, 0.9, 0.9, 1.4, 1.4, 1.1, 1.3,
    1.0, 1.2, 0.9, 0.9, 0.9, 1.0, 0.9, 1.0, 1.0, 1.3,
    1.2, 1.2, 0.9, 1.0, 1.7, 1.5, 1.5, 1.3, 1.3, 1.4,
    1.8, 1.4, 1.2, 1.3, 1.3, 1.4, 1.1, 1.1, 1.2, 1.6,
    1.4, 1.3, 1.2, 1.0, 1.0, 0.9, 1.3, 1.2, 1.2, 1.0,
    1.2, 1.2, 1.1, 1.2, 1.1, 2.1, 2.2, 1.8, 1.7, 1.3,
    1.4, 1.2, 1.2, 1.3, 1.4, 1.4, 1.7, 1.9, 1.9, 2.0,
    2.0, 1.6, 2.0], dtype=np.float32).astype(np.float64)

_LEBEDEV_DEGREE = {110: 17, 194: 23, 302: 29, 590: 41, 770: 47}


def z2nr(z: int, lvl: int) -> int:
    """meshmod@proc.f90:569-598.  `nr * 2.37` etc. are integer x default real: single-precision products."""
    nr = 15
    if z > 2:
        nr = 20
    if z > 10:
        nr = 25
    if z > 18:
        nr = 35
    if z > 36:
        nr = 60
    if z > 54:
        nr = 85
    if z > 86:
        nr = 110
    fac = {MESH_LEVEL_SMALL: 2.37, MESH_LEVEL_NORMAL: 3.08, MESH_LEVEL_GOOD: 3.42, MESH_LEVEL_VGOOD: 4.27,
           MESH_LEVEL_AMAZING: 6.72}[lvl]
    return int(np.ceil(np.float32(nr) * np.float32(fac)))


def z2nang(z: int, lvl: int) -> int:
    """meshmod@proc.f90:603-622"""
    return {MESH_LEVEL_SMALL: 110, MESH_LEVEL_NORMAL: 194, MESH_LEVEL_GOOD: 302, MESH_LEVEL_VGOOD: 590,
            MESH_LEVEL_AMAZING: 770}[lvl]


def gauleg(x1: float, x2: float, n: int):
    """tools_math@proc.f90:1289-1350: Gauss-Legendre nodes/weights by Newton iteration, same start values and tolerance."""
    x, w = np.zeros(n), np.zeros(n)
    m = (n + 1) // 2
    xm, xl = 0.5 * (x2 + x1), 0.5 * (x2 - x1)
    for i in range(1, m + 1):
        z = np.cos(np.pi * (i - 0.25) / (n + 0.5))
        while True:
            p1, p2 = 1.0, 0.0
            for j in range(1, n + 1):
                p3, p2 = p2, p1
                p1 = ((2.0 * j - 1.0) * z * p2 - (j - 1.0) * p3) / j
            pp = n * (z * p1 - p2) / (z * z - 1.0)
            z1 = z
            z = z1 - p1 / pp
            if abs(z - z1) <= 3.0e-16:
                break
        x[i - 1] = xm - xl * z
        x[n - i] = xm + xl * z
        w[i - 1] = 2.0 * xl / ((1.0 - z * z) * pp * pp)
        w[n - i] = w[i - 1]
    return x, w


def rmesh_franchini(n: int, iz: int):
    """meshmod@proc.f90:531-563"""
    q, wq = gauleg(-1.0, 1.0, n)
    log2 = np.log(2.0)
    zeta = _ZETA[iz - 1]
    r = zeta / log2 * (1.0 + q) * np.log(2.0 / (1.0 - q))
    w = 4.0 * np.pi * r ** 2 * wq * zeta / log2 * (np.log(2.0 / (1.0 - q)) + (1.0 + q) / (1.0 - q))
    return r, w


def lebedev(nang: int):
    from scipy.integrate import lebedev_rule
    x, w = lebedev_rule(_LEBEDEV_DEGREE[nang])
    assert x.shape[1] == nang
    return x.T.copy(), w / (4.0 * np.pi)          # wang = wang / fourpi, :326


def genmesh_franchini(xyz, z, lvl: int = MESH_LEVEL_GOOD):
    """Molecule version of meshmod@proc.f90:231-398.  xyz: (nat, 3) nuclear positions (bohr, the frame the field is
    evaluated in); z: atomic numbers.  Returns x (n, 3), w (n): nodes ordered atom / radial shell / angular node (:379-393).
    For a molecule list_near_atoms(up2d = rmax) returns every atom within rmax = max(largest radial node, 12 bohr) of the
    node (:289-290,:340): plain distances here."""
    xyz = np.asarray(xyz, float).reshape(-1, 3)
    z = np.asarray(z).reshape(-1)
    ok = (z >= 1) & (z <= 118)
    # rmax from the atom with most radial nodes (largest Z on ties), :272-290
    mr, izmr = -1, -1
    for iz in z[ok]:
        nr = z2nr(int(iz), lvl)
        if nr > mr or (nr == mr and iz > izmr):
            mr, izmr = nr, int(iz)
    rmax = max(rmesh_franchini(mr, izmr)[0][-1], 12.0)
    fsc = np.where(z == 1, 0.3, 1.0)
    xs, ws = [], []
    for i in range(len(z)):
        if not ok[i]:
            continue
        iz = int(z[i])
        fscal = 0.3 if iz == 1 else 1.0
        rads, wrads = rmesh_franchini(z2nr(iz, lvl), iz)
        ang, wang = lebedev(z2nang(iz, lvl))
        x = xyz[i][None, None, :] + rads[:, None, None] * ang[None, :, :]          # (nr, nang, 3)
        vp0 = fscal * np.exp(-2.0 * rads) / np.maximum(rads, 1e-10) ** 3             # :330
        d = np.linalg.norm(x[:, :, None, :] - xyz[None, None, :, :], axis=-1)       # (nr, nang, nat)
        vp = fsc[None, None, :] * np.exp(-2.0 * d) / np.maximum(d, 1e-10) ** 3
        vp = np.where((d <= rmax) & ok[None, None, :], vp, 0.0)
        vpsum = np.maximum(vp0[:, None], vp.sum(-1))                                 # :362
        w = vp0[:, None] / np.maximum(vpsum, 1e-40) * wrads[:, None] * wang[None, :]  # :365
        xs.append(x.reshape(-1, 3))
        ws.append(w.reshape(-1))
    return np.concatenate(xs), np.concatenate(ws)

