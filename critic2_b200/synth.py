"""Synthetic wavefunctions of the shapes BASELINE.json names.

A def2-TZVP-shaped primitive basis (heavy atoms 11s6p2d1f = 51 Cartesian primitives, H 5s1p = 8) on a jittered
simple-cubic lattice, with i.i.d. N(0,1)/sqrt(npri) primitive-level MO coefficients and occupation 2.  Everything is
seeded, so the same arrays are produced on the CPU container and on the GPU box.
"""
from __future__ import annotations

import numpy as np

from .wfn import Molwfn, WFN_RHF, complete_struct

# public def2-TZVP exponents (carbon; N and O are scaled by (Z/6)^2) and hydrogen
_C_S = [13575.349682, 2035.2333680, 463.22562359, 131.20019598, 42.853015891, 15.584185766,
        6.2067138508, 2.5764896527, 0.57696339419, 0.22972831358, 0.095164440028]
_C_P = [34.697232244, 7.9582622826, 2.3780826883, 0.81433208183, 0.28887547253, 0.10056823671]
_C_D = [1.097, 0.318]
_C_F = [0.761]
_H_S = [34.0613410, 5.1235746, 1.1646626, 0.32723041, 0.10307241]
_H_P = [0.8]


def _atom_prims(z: int, gh: bool = False):
    """[(itype, exponent)] for one atom.  gh: heavy atoms also get one g shell (types 21-35) and one h shell (36-56, the
    largest angular momentum of calculate_mo_gto's type table, wfn_private@proc.F90:2695-2705)."""
    out = []
    if z == 1:
        out += [(1, a) for a in _H_S]
        for a in _H_P:
            out += [(t, a) for t in (2, 3, 4)]
        return out
    sc = (z / 6.0) ** 2
    out += [(1, a * sc) for a in _C_S]
    for a in _C_P:
        out += [(t, a * sc) for t in (2, 3, 4)]
    for a in _C_D:
        out += [(t, a * sc) for t in range(5, 11)]
    for a in _C_F:
        out += [(t, a * sc) for t in range(11, 21)]
    if gh:
        out += [(t, 1.011 * sc) for t in range(21, 36)]
        out += [(t, 0.873 * sc) for t in range(36, 57)]
    return out


def synthetic_molecule(lattice=(8, 8, 8), ndrop=12, composition=((6, 160), (7, 20), (8, 20), (1, 300)),
                       seed=4, spacing=4.0, jitter=0.5, nmo=None, uhf=False, with_edf=False, gh=False) -> Molwfn:
    """Config 4 by default (500 atoms, 12 600 primitives, 780 MOs).  Config 5: lattice=(13,13,12), ndrop=28,
    composition=((6,640),(7,80),(8,80),(1,1200)), seed=5."""
    rng = np.random.default_rng(seed)
    nx, ny, nz = lattice
    sites = np.array([(i, j, k) for i in range(nx) for j in range(ny) for k in range(nz)], float)
    keep = np.ones(len(sites), bool)
    if ndrop > 0:
        keep[rng.choice(len(sites), size=ndrop, replace=False)] = False
    sites = sites[keep]
    natom = len(sites)
    xyz = sites * spacing + rng.uniform(-jitter, jitter, size=sites.shape)
    zs = np.concatenate([np.full(n, z, np.int32) for z, n in composition])
    if len(zs) != natom:
        raise ValueError(f"composition has {len(zs)} atoms, lattice has {natom} sites")
    rng.shuffle(zs)
    icenter, itype, e = [], [], []
    for ia, z in enumerate(zs):
        for t, a in _atom_prims(int(z), gh):
            icenter.append(ia + 1)
            itype.append(t)
            e.append(a)
    npri = len(e)
    nelec = int(zs.sum())
    if nmo is None:
        nmo = nelec if uhf else nelec // 2
    cmo = rng.standard_normal((npri, nmo)) / np.sqrt(npri)
    if uhf:
        occ, wfntyp, nalpha = np.ones(nmo), 1, (nmo + 1) // 2
    else:
        occ, wfntyp, nalpha = np.full(nmo, 2.0), WFN_RHF, nmo
    f = Molwfn(natom=natom, xyz=xyz, z=zs, icenter=np.array(icenter, np.int32), itype=np.array(itype, np.int32),
               e=np.array(e), occ=occ, cmo=np.ascontiguousarray(cmo), nmoocc=nmo, wfntyp=wfntyp, nalpha=nalpha)
    if with_edf:
        # a small ECP-like core on every heavy atom: two s-type EDF Gaussians (read_wfx :864-873)
        ic, it, ee, cc = [], [], [], []
        for ia, z in enumerate(zs):
            if z > 1:
                for a, c in ((12.0 * z, 3.0 * z), (3.0 * z, 0.7 * z)):
                    ic.append(ia + 1); it.append(1); ee.append(a); cc.append(c)
        f.icenter_edf = np.array(ic, np.int32)
        f.itype_edf = np.array(it, np.int32)
        f.e_edf = np.array(ee)
        f.c_edf = np.array(cc)
    complete_struct(f)
    return f


def config4(seed=4) -> Molwfn:
    return synthetic_molecule(seed=seed)


def config5(seed=5) -> Molwfn:
    return synthetic_molecule(lattice=(13, 13, 12), ndrop=28,
                              composition=((6, 640), (7, 80), (8, 80), (1, 1200)), seed=seed)


def small_synthetic(seed=11, uhf=False, with_edf=False, nmo=None, gh=False) -> Molwfn:
    """A 24-atom cousin of config 4 (3x3x3 lattice, 3 sites dropped) for parity tests the oracle finishes in seconds."""
    return synthetic_molecule(lattice=(3, 3, 3), ndrop=3, composition=((6, 7), (7, 2), (8, 2), (1, 13)), seed=seed,
                              uhf=uhf, with_edf=with_edf, nmo=nmo, gh=gh)
