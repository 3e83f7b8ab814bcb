"""Shared builders for the parity tests: golden wavefunctions, structures, grids and the tolerance checks.

Tolerances are BASELINE.json's: rho relative error <= 1e-10, derivatives / RDG <= 1e-8 (relative to the magnitude of
the whole vector / tensor at the point, with an absolute floor for points where the reference value is ~0), and
identical NCI classification."""
import os

import numpy as np

from critic2_b200 import structure, synth
from critic2_b200.wfn import Molwfn, complete_struct

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
TOL_RHO, TOL_DER = 1e-10, 1e-8


def golden_wfn(name: str) -> Molwfn:
    d = np.load(os.path.join(GOLDEN, f"wfn_{name}.npz"))
    f = Molwfn(natom=int(d["xyz"].shape[0]), xyz=d["xyz"], z=d["z"], icenter=d["icenter"], itype=d["itype"], e=d["e"],
               occ=d["occ"], cmo=np.ascontiguousarray(d["cmo"]), nmoocc=int(d["nmoocc"]), wfntyp=int(d["wfntyp"]),
               nalpha=int(d["nalpha"]))
    complete_struct(f)
    return f


def tables():
    d = np.load(os.path.join(GOLDEN, "promol_tables.npz"))
    out = {}
    for z in (1, 6, 7, 8, 14):
        out[z] = {k: (d[f"{z}_{k}"] if d[f"{z}_{k}"].ndim else float(d[f"{z}_{k}"])) for k in ("a", "b", "r", "f", "fp", "fpp")}
    return out


def core_tables():
    """{(Z, zpsp): table} for C 4, N 5, O 6, Si 4 (tools/make_golden.py)"""
    d = np.load(os.path.join(GOLDEN, "core_tables.npz"))
    out = {}
    for z, q in ((6, 4), (7, 5), (8, 6), (14, 4)):
        out[(z, q)] = {k: (d[f"{z}_{q}_{k}"] if d[f"{z}_{q}_{k}"].ndim else float(d[f"{z}_{q}_{k}"]))
                       for k in ("a", "b", "r", "f", "fp", "fpp")}
        out[(z, q)]["q"] = q
    return out


def mol_structure(f: Molwfn, border=None):
    return structure.molecular_cell(f.xyz, f.z) if border is None else structure.molecular_cell(f.xyz, f.z, border)


def bbox_grid(st, n, margin=4.0):
    """n^3 lattice over the atoms' bounding box +- margin (cell frame); returns x0, xd, nn."""
    lo, hi = st.r.min(0) - margin, st.r.max(0) + margin
    nn = np.array([n, n, n] if np.isscalar(n) else n, np.int32)
    xd = np.diag((hi - lo) / nn)
    return lo, xd, nn


def check_rho(got, ref, what="rho", tol=TOL_RHO, floor=1e-30):
    got, ref = np.asarray(got), np.asarray(ref)
    err = np.abs(got - ref) / np.maximum(np.abs(ref), floor)
    # exact-zero mask (outside every cutoff the reference returns exactly 0)
    assert np.array_equal(ref == 0.0, got == 0.0), f"{what}: zero masks differ"
    assert err.max() <= tol, f"{what}: max rel err {err.max():.3e} at {np.unravel_index(err.argmax(), err.shape)}"
    return float(err.max())


def check_vec(got, ref, what, tol=TOL_DER, floor=1e-20, rho=None):
    """got/ref: (..., k) components of one vector/tensor per point; error relative to its Frobenius norm.
    Where the vector (nearly) vanishes by symmetry -- e.g. grad rho at a ring critical point that is a grid node --
    the norm is floored at 1e-4*rho of the point (rho given): the absolute tolerance there is tol*1e-4*rho."""
    got, ref = np.asarray(got), np.asarray(ref)
    nrm = np.sqrt((ref ** 2).sum(-1, keepdims=True))
    if rho is not None:
        floor = np.maximum(1e-4 * np.abs(np.asarray(rho))[..., None], floor)
    err = np.abs(got - ref) / np.maximum(nrm, floor)
    assert err.max() <= tol, f"{what}: max rel err {err.max():.3e}"
    return float(err.max())


def quartz(n=1):
    return structure.quartz_supercell(n)
