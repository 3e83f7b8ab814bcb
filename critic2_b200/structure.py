"""Host-side mirror of the geometry set-up the hot path consumes.

* ``molecular_cell`` -- crystal%struct_new for MOLECULE input without a cell
  (/root/reference/src/crystalmod@proc.f90:217-233, 250-256, 285-304, 334-339).
* ``crystal_cell`` -- CRYSTAL input from cell parameters (m_x2c_from_cellpar,
  /root/reference/src/tools_math@proc.f90:69-97); the reduced cell is taken equal to the cell
  (m_x2xr = identity), which is what Delaunay reduction gives for the cells used here.
* ``Structure.environment_atoms`` -- which periodic images the environment holds
  (environ_build, /root/reference/src/environmod@proc.f90:139-213).

Only bookkeeping: no field is evaluated here.
"""
from __future__ import annotations

import math
from dataclasses import dataclass

import numpy as np

BOHRTOA = 0.52917720859  # param.F90:42
RBORDER_DEF = 10.0 / BOHRTOA  # global.f90:113

# cutrad, global.f90:51-93 (bohr), index Z-1
CUTRAD = np.array([
    2.149886192475e+01, 1.169139170668e+01, 3.430831385801e+01, 2.502075396007e+01, 2.801001395722e+01,
    2.167675592180e+01, 1.749805708313e+01, 1.465173060207e+01, 1.263885024136e+01, 1.110521599057e+01,
    3.523728402162e+01, 2.763528367271e+01, 3.395549316507e+01, 2.847261278601e+01, 2.487715217494e+01,
    2.222930087269e+01, 2.022231415676e+01, 1.857150175607e+01, 3.884428729523e+01, 3.144587224767e+01,
    2.970981796151e+01, 2.864438811442e+01, 2.784088946336e+01, 2.925194799711e+01, 2.660566532177e+01,
    2.609916866690e+01, 2.558901439004e+01, 2.517359152887e+01, 2.691554955610e+01, 2.435659411320e+01,
    3.467478603212e+01, 2.914443825602e+01, 2.572575006996e+01, 2.323452863278e+01, 2.134146595122e+01,
    1.981582897591e+01, 3.976877622180e+01, 3.266858263171e+01, 3.027851405458e+01, 2.899491720657e+01,
    2.967865003580e+01, 2.914637014504e+01, 2.697201600611e+01, 2.844039136970e+01, 2.814409350112e+01,
    1.659926809140e+01, 2.771163603049e+01, 2.519886588880e+01, 3.538116802480e+01, 3.026454800251e+01,
    2.695514982633e+01, 2.460024202780e+01, 2.277601677390e+01, 2.130658017554e+01, 4.137458546886e+01,
    3.442036204804e+01, 3.242561614450e+01, 3.212250868201e+01, 3.306457792690e+01, 3.284026775197e+01,
    3.262654222620e+01, 3.242292112974e+01, 3.222895504883e+01, 3.128431696800e+01, 3.180465714311e+01,
    3.157435555252e+01, 3.135291924508e+01, 3.120231512615e+01, 3.099709515455e+01, 3.079969409503e+01,
    3.160515129459e+01, 2.709458469010e+01, 2.614193052742e+01, 2.548104664032e+01, 2.489113924347e+01,
    2.441668377017e+01, 2.405143298004e+01, 2.466268008529e+01, 2.439924398342e+01, 2.305709117567e+01,
    3.643576493190e+01, 3.110226831614e+01, 2.780342993946e+01, 2.541102668192e+01, 2.360240806573e+01,
    2.210165966674e+01, 4.053200388132e+01, 3.407838067822e+01, 3.585071927373e+01, 3.175945034367e+01,
    3.478340806986e+01, 3.489038964505e+01, 3.514212336660e+01, 3.120895952111e+01] + [37.0] * 24)


@dataclass
class Structure:
    """The subset of ``type crystal`` + ``type environ`` the path reads. Matrices are numpy (row, col) = Fortran (i,j)."""
    ismolecule: bool
    m_x2c: np.ndarray      # crystallographic -> Cartesian
    m_c2x: np.ndarray
    m_xr2c: np.ndarray     # reduced crystallographic -> Cartesian
    m_c2xr: np.ndarray
    m_x2xr: np.ndarray
    x: np.ndarray          # (ncell,3) crystallographic coordinates in [0,1)
    r: np.ndarray          # (ncell,3) Cartesian, cell frame (bohr)
    z: np.ndarray          # (ncell,) atomic numbers
    molx0: np.ndarray      # origin shift: real-space position = r + molx0 (molecules)
    molborder: np.ndarray  # fractional border of the molecular cell
    omega: float

    @property
    def ncell(self) -> int:
        return int(self.r.shape[0])

    def species(self):
        """(zspc, is): species table in order of first appearance; is is 1-based."""
        zspc, ismap, iss = [], {}, []
        for zz in self.z:
            if int(zz) not in ismap:
                ismap[int(zz)] = len(zspc) + 1
                zspc.append(int(zz))
            iss.append(ismap[int(zz)])
        return np.array(zspc, np.int32), np.array(iss, np.int32)

    def sphmax(self) -> float:
        """Half-diagonal of the reduced cell, environmod@proc.f90:116-121."""
        c = np.array([[-.5, -.5, -.5], [.5, -.5, -.5], [-.5, .5, -.5], [-.5, -.5, .5]])
        return float(np.max(np.linalg.norm(c @ self.m_xr2c.T, axis=1)))

    def dmax0(self) -> float:
        """environmod@proc.f90:123-133 (no dmax0 argument)."""
        return max(float(np.max(CUTRAD[self.z - 1])), self.sphmax()) + 1e-2

    def environment_atoms(self, dmax0: float | None = None):
        """Atoms of the environment: the cell atoms first (folded to the main reduced cell for crystals), then all
        periodic images with |r - centre| < sphmax + dmax0 (environmod@proc.f90:143-213, addcell :301-330).
        Returns (r_env (n,3), cidx (n,) 1-based)."""
        if self.ismolecule:
            return self.r.copy(), np.arange(1, self.ncell + 1, dtype=np.int32)
        if dmax0 is None:
            dmax0 = self.dmax0()
        sph = self.sphmax()
        xr = self.x @ self.m_x2xr.T
        xr = xr - np.floor(xr)
        r0 = xr @ self.m_xr2c.T
        centre = self.m_xr2c @ np.array([.5, .5, .5])
        rs, cs = [r0], [np.arange(1, self.ncell + 1, dtype=np.int32)]
        imax = 0
        while True:
            imax += 1
            found = False
            rng = range(-imax, imax + 1)
            for i1 in rng:
                for i2 in rng:
                    for i3 in rng:
                        if max(abs(i1), abs(i2), abs(i3)) != imax:
                            continue
                        xc = (xr + np.array([i1, i2, i3], float)) @ self.m_xr2c.T
                        m = np.linalg.norm(xc - centre, axis=1) < sph + dmax0
                        if m.any():
                            found = True
                            rs.append(xc[m])
                            cs.append(np.nonzero(m)[0].astype(np.int32) + 1)
            if not found:
                break
        return np.concatenate(rs), np.concatenate(cs)


def m_x2c_from_cellpar(aa, bb) -> np.ndarray:
    """tools_math@proc.f90:69-97."""
    deg2rad = math.atan(1.0) / 45.0
    ca, cb, cg = (math.cos(b * deg2rad) for b in bb)
    h1 = math.sqrt(max(1 - cg * cg, 0.0))
    h2 = (ca - cb * cg) / h1
    h3 = math.sqrt(max(1 - cb * cb - h2 * h2, 0.0))
    m = np.zeros((3, 3))
    m[0, 0] = aa[0]
    m[0, 1] = aa[1] * cg
    m[0, 2] = aa[2] * cb
    m[1, 1] = aa[1] * h1
    m[1, 2] = aa[2] * h2
    m[2, 2] = aa[2] * h3
    return m


def molecular_cell(xyz: np.ndarray, z: np.ndarray, border: float = RBORDER_DEF, cubic: bool = False) -> Structure:
    """MOLECULE without an explicit cell: crystalmod@proc.f90:217-233, 250-256, 285-304, 334-339."""
    xyz = np.asarray(xyz, float)
    xmax = (xyz + border).max(axis=0)
    xmin = (xyz - border).min(axis=0)
    xcm = xyz.sum(axis=0) / xyz.shape[0]
    if cubic:
        xmin[:] = xmin.min()
        xmax[:] = xmax.max()
    aa = xmax - xmin
    m_x2c = m_x2c_from_cellpar(aa, (90.0, 90.0, 90.0))
    m_c2x = np.linalg.inv(m_x2c)
    x = (xyz - xcm + 0.5 * aa) / aa
    molx0 = -0.5 * aa + xcm
    molborder = np.maximum(border - max(2.0, 0.8 * border), 0.0) / (xmax - xmin)
    x = x - np.floor(x)
    r = x @ m_x2c.T
    eye = np.eye(3)
    return Structure(True, m_x2c, m_c2x, m_x2c.copy(), m_c2x.copy(), eye, x, r, np.asarray(z, np.int32), molx0,
                     molborder, float(abs(np.linalg.det(m_x2c))))


def crystal_cell(aa, bb, xfrac: np.ndarray, z: np.ndarray) -> Structure:
    """CRYSTAL from cell lengths (bohr) / angles (deg) and fractional coordinates."""
    m_x2c = m_x2c_from_cellpar(aa, bb)
    m_c2x = np.linalg.inv(m_x2c)
    x = np.asarray(xfrac, float)
    x = x - np.floor(x)
    r = x @ m_x2c.T
    return Structure(False, m_x2c, m_c2x, m_x2c.copy(), m_c2x.copy(), np.eye(3), x, r, np.asarray(z, np.int32),
                     np.zeros(3), np.zeros(3), float(abs(np.linalg.det(m_x2c))))


def quartz_supercell(n: int = 4) -> Structure:
    """alpha-quartz (P3_2 21 setting, literature values: a=4.913 A, c=5.405 A, Si 3a (0.4697,0,0),
    O 6c (0.4135,0.2669,0.1191)) repeated n x n x n: BASELINE config 3 (576 atoms for n=4)."""
    a, c = 4.913 / BOHRTOA, 5.405 / BOHRTOA
    u = 0.4697
    ox, oy, oz = 0.4135, 0.2669, 0.1191
    si = [(u, 0, 0), (0, u, 2 / 3), (-u, -u, 1 / 3)]
    # general position orbit of P3_2 21 (No. 154)
    o = [(ox, oy, oz), (-oy, ox - oy, oz + 2 / 3), (-ox + oy, -ox, oz + 1 / 3),
         (oy, ox, -oz), (ox - oy, -oy, 2 / 3 - oz), (-ox, -ox + oy, 1 / 3 - oz)]
    base = np.array(si + o, float)
    zs = [14] * 3 + [8] * 6
    xs, zz = [], []
    for i in range(n):
        for j in range(n):
            for k in range(n):
                xs.append((base + np.array([i, j, k])) / n)
                zz += zs
    return crystal_cell((a * n, a * n, c * n), (90.0, 90.0, 120.0), np.concatenate(xs), np.array(zz, np.int32))
