"""ctypes wrapper around oracle/_build/liboracle.so -- the CPU restatement of the reference path.

TEST INFRASTRUCTURE ONLY.  Import this only from tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs.
The product package (critic2_b200) must never import it.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = os.path.join(_HERE, "_build", "liboracle.so")
_SRC = ["oracle_env.c", "oracle_wfn.c", "oracle_promol.c", "oracle_field.c", "oracle_text.c", "oracle.h", "oracle_util.h"]


def build(force: bool = False) -> str:
    """Compile the oracle with gcc (a few seconds)."""
    newest = max(os.path.getmtime(os.path.join(_HERE, s)) for s in _SRC)
    if force or not os.path.exists(_LIB) or os.path.getmtime(_LIB) < newest:
        cc = "/usr/bin/gcc" if os.path.exists("/usr/bin/gcc") else "gcc"
        subprocess.check_call(["make", "-C", _HERE, f"CC={cc}"], stdout=subprocess.DEVNULL)
    return _LIB


_lib = None
dp = C.POINTER(C.c_double)
ip = C.POINTER(C.c_int)


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_LIB):
            build()
        _lib = C.CDLL(_LIB)
        L = _lib
        L.orc_environ_build.restype = C.c_void_p
        L.orc_environ_build.argtypes = [C.c_int, C.c_int, ip, C.c_int, dp, ip, dp, dp, dp, C.c_double]
        L.orc_environ_free.argtypes = [C.c_void_p]
        L.orc_wfn_create.restype = C.c_void_p
        L.orc_wfn_create.argtypes = [C.c_int, C.c_int, ip, ip, dp, C.c_int, C.c_int, dp, dp, C.c_int, C.c_int, C.c_int,
                                     ip, ip, dp, dp]
        L.orc_wfn_free.argtypes = [C.c_void_p]
        L.orc_wfn_tables.argtypes = [C.c_void_p, dp, dp, ip, ip, ip, ip]
        L.orc_rho2.argtypes = [C.c_void_p, C.c_void_p, dp, C.c_int, dp]
        L.orc_grid1_from_table.restype = C.c_void_p
        L.orc_grid1_from_table.argtypes = [C.c_int, C.c_int, C.c_int, C.c_double, C.c_double, dp, dp, dp, dp]
        L.orc_grid1_read_critic.restype = C.c_void_p
        L.orc_grid1_read_critic.argtypes = [C.c_char_p, C.c_int, C.c_int]
        L.orc_grid1_free.argtypes = [C.c_void_p]
        L.orc_grid1_interp.argtypes = [C.c_void_p, C.c_double, dp, dp, dp]
        L.orc_promolecular.argtypes = [C.c_void_p, C.POINTER(C.c_void_p), dp, C.c_int, C.c_char_p, dp, dp, dp]
        L.orc_grd.argtypes = [C.c_void_p, dp, C.c_int, C.c_int, dp]
        L.orc_grd0.restype = C.c_double
        L.orc_grd0.argtypes = [C.c_void_p, dp, C.c_int]
        L.orc_eig.restype = C.c_int
        L.orc_eig.argtypes = [dp, dp, C.c_int]
        L.orc_cube.argtypes = [C.c_void_p, dp, dp, ip, C.c_int, C.c_uint, C.POINTER(dp), C.c_int]
        L.orc_nci.argtypes = [C.c_void_p, C.c_void_p, dp, dp, ip, dp, dp, dp, C.c_int]
        L.orc_nci_cutoffs.restype = C.c_long
        L.orc_nci_cutoffs.argtypes = [ip, dp, dp, dp, C.c_double, C.c_double, C.c_double, dp, C.c_int, C.c_double,
                                      C.c_double, dp, dp, C.c_long]
        L.orc_nci_fragments.argtypes = [C.c_int, ip, dp, ip, C.POINTER(C.c_void_p), dp, dp, ip, dp, C.c_int]
        L.orc_nci_cutoffs_frag.restype = C.c_long
        L.orc_nci_cutoffs_frag.argtypes = [ip, dp, dp, dp, C.c_int, dp, C.c_double, C.c_double, C.c_double, C.c_double,
                                           C.c_double, dp, C.c_int, C.c_double, C.c_double, dp, dp, C.c_long]
        L.orc_text_record_bytes.restype = C.c_long
        L.orc_text_record_bytes.argtypes = [C.c_int, C.c_int]
        L.orc_cube_text.argtypes = [dp, ip, C.c_int, C.c_int, C.c_char_p, C.c_int]
        L.orc_promolecular_grid.argtypes = [C.c_void_p, C.POINTER(C.c_void_p), ip, C.c_char_p, dp, C.c_int]
        L.orc_promol_nat_points.argtypes = [C.c_void_p, C.POINTER(C.c_void_p), C.c_long, dp, C.c_char_p, C.POINTER(C.c_long), C.c_int]
        L.orc_grd_points.argtypes = [C.c_void_p, C.c_long, dp, C.c_int, C.c_int, dp, C.c_int]
        L.orc_mo_points.argtypes = [C.c_void_p, C.c_long, dp, C.c_int, C.c_int, dp, C.c_int]
        L.orc_nact_points.argtypes = [C.c_void_p, C.c_void_p, C.c_long, dp, C.POINTER(C.c_long), C.c_int]
        L.orc_cutrad.restype = C.c_double
        L.orc_cutrad.argtypes = [C.c_int]
        L.orc_identify_atom.restype = C.c_int
        L.orc_identify_atom.argtypes = [C.c_void_p, dp, C.c_double]
    return _lib


def _d(a):
    return a.ctypes.data_as(dp)


def _i(a):
    return a.ctypes.data_as(ip)


def _f64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


def _i32(a):
    return np.ascontiguousarray(a, dtype=np.int32)


def _colmajor(m):
    """numpy (i,j) matrix -> flat Fortran order"""
    return np.ascontiguousarray(np.asarray(m, float).T).reshape(-1)


class _OrcField(C.Structure):
    _fields_ = [("type", C.c_int), ("env", C.c_void_p), ("wfn", C.c_void_p), ("agrid", C.POINTER(C.c_void_p)),
                ("fragmask", C.c_char_p), ("cgrid", C.POINTER(C.c_void_p))]


RES_LEN = 30
PROPS = ("f", "gx", "gy", "gz", "gmod", "hxx", "hxy", "hxz", "hyy", "hyz", "hzz", "lap")


class Environ:
    """orc_environ built from a critic2_b200.structure.Structure"""

    def __init__(self, st, dmax0: float = -1.0):
        zspc, iss = st.species()
        self.st = st
        self.ptr = lib().orc_environ_build(int(st.ismolecule), len(zspc), _i(_i32(zspc)), st.ncell, _d(_f64(st.x)),
                                           _i(_i32(iss)), _d(_colmajor(st.m_x2c)), _d(_colmajor(st.m_x2xr)),
                                           _d(_colmajor(st.m_xr2c)), float(dmax0))

    def identify_atom(self, x, distmax=1e-5):
        return lib().orc_identify_atom(self.ptr, _d(_f64(x)), distmax)

    def __del__(self):
        if getattr(self, "ptr", None):
            lib().orc_environ_free(self.ptr)
            self.ptr = None


class Grid1Set:
    """agrid(:) -- radial tables indexed by atomic number. tables: {Z: dict(a,b,r,f,fp,fpp[,q])}; with q = zpsp(Z) > 0
    the set is the core selection cgrid(Z, zpsp(Z)) that grd's core augmentation uses."""

    def __init__(self, tables):
        self.arr = (C.c_void_p * 119)()
        self.ptrs = []
        for z, t in tables.items():
            r, f, fp, fpp = (_f64(t[k]) for k in ("r", "f", "fp", "fpp"))
            p = lib().orc_grid1_from_table(int(z), int(t.get("q", 0)), len(r), float(t["a"]), float(t["b"]), _d(r), _d(f), _d(fp), _d(fpp))
            self.arr[int(z)] = p
            self.ptrs.append(p)

    def interp(self, z, r0):
        f, fp, fpp = C.c_double(), C.c_double(), C.c_double()
        lib().orc_grid1_interp(self.arr[z], r0, C.byref(f), C.byref(fp), C.byref(fpp))
        return f.value, fp.value, fpp.value

    def __del__(self):
        for p in getattr(self, "ptrs", []):
            lib().orc_grid1_free(p)
        self.ptrs = []


def read_critic(path: str, z: int, q: int = 0):
    """grid1%read_db/read_critic on a dat/wfc file; returns dict(a,b,r,f,fp,fpp) (numpy copies)."""

    class G(C.Structure):
        _fields_ = [("isinit", C.c_int), ("ngrid", C.c_int), ("z", C.c_int), ("qat", C.c_int), ("a", C.c_double),
                    ("b", C.c_double), ("rmax", C.c_double), ("rmax2", C.c_double), ("r", dp), ("f", dp), ("fp", dp),
                    ("fpp", dp)]

    p = lib().orc_grid1_read_critic(path.encode(), z, q)
    if not p:
        raise FileNotFoundError(path)
    g = C.cast(p, C.POINTER(G)).contents
    n = g.ngrid
    out = dict(a=g.a, b=g.b, r=np.array(g.r[:n]), f=np.array(g.f[:n]), fp=np.array(g.fp[:n]), fpp=np.array(g.fpp[:n]))
    lib().orc_grid1_free(p)
    return out


class Wfn:
    """orc_wfn from a critic2_b200.wfn.Molwfn (complete_struct is recomputed independently in C)."""

    def __init__(self, f):
        self.f = f
        ne = f.nedf
        self.ptr = lib().orc_wfn_create(
            f.natom, f.npri, _i(_i32(f.icenter)), _i(_i32(f.itype)), _d(_f64(f.e)), f.nmoocc, f.nmoall, _d(_f64(f.cmo)),
            _d(_f64(f.occ)), f.wfntyp, f.nalpha, ne, _i(_i32(f.icenter_edf)) if ne else None,
            _i(_i32(f.itype_edf)) if ne else None, _d(_f64(f.e_edf)) if ne else None, _d(_f64(f.c_edf)) if ne else None)

    def rho2(self, env: "Environ", x, nder=2):
        """molwfn%rho2 at one Cartesian point: dict of rho(3), rhoval(3), grad(3,3), gradval(3,3), h(3,3,3), hval(3,3,3),
        gkin(3), vir, stress(3,3) -- last index = total / alpha / beta (arrays indexed [spin][...])"""
        out = np.zeros(91)
        lib().orc_rho2(self.ptr, env.ptr, _d(_f64(x)), int(nder), _d(out))
        F = lambda a, shp: np.asarray(a).reshape(shp[::-1]).transpose(*range(len(shp) - 1, -1, -1))   # Fortran order
        return dict(rho=out[0:3].copy(), rhoval=out[3:6].copy(), grad=F(out[6:15], (3, 3)).T.copy(),
                    gradval=F(out[15:24], (3, 3)).T.copy(), h=np.moveaxis(F(out[24:51], (3, 3, 3)), 2, 0).copy(),
                    hval=np.moveaxis(F(out[51:78], (3, 3, 3)), 2, 0).copy(), gkin=out[78:81].copy(), vir=out[81],
                    stress=F(out[82:91], (3, 3)).copy())

    def tables(self):
        f = self.f
        dran, spc = np.zeros(f.npri), np.zeros(f.natom)
        lmax, icord = np.zeros(f.natom, np.int32), np.zeros(f.npri, np.int32)
        lo, hi = np.zeros(f.natom, np.int32), np.zeros(f.natom, np.int32)
        lib().orc_wfn_tables(self.ptr, _d(dran), _d(spc), _i(lmax), _i(icord), _i(lo), _i(hi))
        return dran, spc, lmax, icord, lo, hi

    def __del__(self):
        if getattr(self, "ptr", None):
            lib().orc_wfn_free(self.ptr)
            self.ptr = None


class Field:
    """orc_field: type_wfn (wfn given) or type_promol (grids given)"""

    def __init__(self, env: Environ, wfn: Wfn | None = None, grids: Grid1Set | None = None, fragmask=None,
                 core: Grid1Set | None = None):
        self.env, self.wfn, self.grids, self.core = env, wfn, grids, core
        self._frag = None if fragmask is None else np.ascontiguousarray(fragmask, np.uint8).tobytes()
        self.c = _OrcField(0 if wfn is not None else 1, env.ptr, wfn.ptr if wfn is not None else None,
                           grids.arr if grids is not None else None, self._frag,
                           core.arr if core is not None else None)

    @property
    def ref(self):
        return C.byref(self.c)

    def grd(self, v, nder=2, periodic=True):
        res = np.zeros(RES_LEN)
        lib().orc_grd(self.ref, _d(_f64(v)), nder, int(periodic), _d(res))
        return res

    def grd_points(self, pts, nder=2, periodic=True, nthreads=0):
        pts = _f64(pts)
        res = np.zeros((pts.shape[0], RES_LEN))
        lib().orc_grd_points(self.ref, pts.shape[0], _d(pts), nder, int(periodic), _d(res), nthreads)
        return res

    def mo_points(self, pts, imo: int, periodic=True, nthreads=0):
        """grd(nder=-1, fder=imo): MO values (imo 1-based)"""
        pts = _f64(pts)
        out = np.zeros(pts.shape[0])
        lib().orc_mo_points(self.ref, pts.shape[0], _d(pts), int(imo), int(periodic), _d(out), nthreads)
        return out

    def nact_points(self, pts, nthreads=0):
        pts = _f64(pts)
        out = np.zeros(pts.shape[0], np.int64)
        lib().orc_nact_points(self.wfn.ptr, self.env.ptr, pts.shape[0], _d(pts), out.ctypes.data_as(C.POINTER(C.c_long)),
                              nthreads)
        return out

    def cube(self, x0, xd, nn, props=PROPS, periodic=False, nthreads=0):
        """xd: numpy (3,3) with xd[:,i] the i-th step vector. Returns {prop: array (nn3,nn2,nn1) C-order = x fastest}."""
        nn = _i32(nn)
        npt = int(nn[0]) * int(nn[1]) * int(nn[2])
        mask = 0
        outs = []
        for nti, p in enumerate(PROPS):
            if p in props:
                mask |= 1 << nti
                outs.append(np.zeros(npt))
        arr = (dp * len(outs))(*[_d(o) for o in outs])
        lib().orc_cube(self.ref, _d(_f64(x0)), _d(_colmajor(xd)), _i(nn), int(periodic), mask, arr, nthreads)
        names = [p for p in PROPS if p in props]
        return {p: o.reshape(nn[2], nn[1], nn[0]) for p, o in zip(names, outs)}

    def nci(self, x0, xmat, nstep, promol: "Field | None" = None, nthreads=0):
        """Returns crho, cgrad[, rhoat] as arrays (n1,n2,n3) C-order = z fastest, like the reference's (k,j,i)."""
        nstep = _i32(nstep)
        npt = int(nstep[0]) * int(nstep[1]) * int(nstep[2])
        crho, cgrad = np.zeros(npt), np.zeros(npt)
        rhoat = np.zeros(npt) if promol is not None else None
        lib().orc_nci(self.ref, promol.ref if promol is not None else None, _d(_f64(x0)), _d(_colmajor(xmat)), _i(nstep),
                      _d(crho), _d(cgrad), _d(rhoat) if rhoat is not None else None, nthreads)
        shp = (nstep[0], nstep[1], nstep[2])
        return crho.reshape(shp), cgrad.reshape(shp), (rhoat.reshape(shp) if rhoat is not None else None)


def nci_fragments(frags, grids: Grid1Set, x0, xmat, nstep, nthreads=0):
    """rhofrag of nciplot (nci@proc.f90:534-550).  frags: list of (xyz (nat,3) Cartesian bohr, z (nat,)).
    Returns (nfrag, n1, n2, n3), z fastest."""
    nstep = _i32(nstep)
    npt = int(nstep[0]) * int(nstep[1]) * int(nstep[2])
    nat = _i32([len(f[1]) for f in frags])
    xyz = _f64(np.concatenate([np.asarray(f[0], float).reshape(-1, 3) for f in frags]))
    z = _i32(np.concatenate([np.asarray(f[1]).reshape(-1) for f in frags]))
    out = np.zeros(len(frags) * npt)
    lib().orc_nci_fragments(len(frags), _i(nat), _d(xyz), _i(z), grids.arr, _d(_f64(x0)), _d(_colmajor(xmat)), _i(nstep),
                            _d(out), nthreads)
    return out.reshape(len(frags), nstep[0], nstep[1], nstep[2])


def nci_cutoffs(crho, cgrad, rhoat=None, rhocut=0.2, dimcut=2.0, rhoplot=0.1, srhorange=(-1e30, 1e30), onlyneg=False,
                rho_void=-1.0, omega=1.0, rhofrag=None, rhoparam=0.95, rhoparam2=0.75):
    """nciplot cutoff pass; modifies cgrad in place. Returns (sums[3], dat (ndat,2)).
    rhofrag: None or (nfrag, n1, n2, n3) fragment densities (then rhoat is required)."""
    shp = crho.shape
    nstep = _i32(shp)
    sums = np.zeros(3)
    maxdat = crho.size
    dat = np.zeros((maxdat, 2))
    sr = _f64(srhorange)
    nfrag = 0 if rhofrag is None else int(rhofrag.shape[0])
    rf = None if rhofrag is None else _f64(rhofrag)
    n = lib().orc_nci_cutoffs_frag(_i(nstep), _d(crho), _d(cgrad), _d(rhoat) if rhoat is not None else None, nfrag,
                                   _d(rf) if rf is not None else None, rhoparam, rhoparam2, rhocut, dimcut, rhoplot,
                                   _d(sr), int(onlyneg), rho_void, omega, _d(sums), _d(dat), maxdat)
    return sums, dat[:n].copy()


def eig(h):
    h = np.array(h, float, order="F").copy(order="F")
    n = h.shape[0]
    ev = np.zeros(n)
    flat = np.ascontiguousarray(h.T).reshape(-1)  # column-major flat
    ierr = lib().orc_eig(_d(flat), _d(ev), n)
    if ierr:
        raise RuntimeError("eig: error in diagonalization")
    return ev


def promolecular_grid(env: Environ, grids: Grid1Set, n, nthreads=0, fragmask=None):
    """crystal%promolecular_grid(f, n, zpsp, fr): grids = agrid, or the core selection cgrid(iz, zpsp(iz)) (tables with q);
    fragmask: per cell atom, the `fr` argument."""
    n = _i32(n)
    out = np.zeros(int(n[0]) * int(n[1]) * int(n[2]))
    fm = None if fragmask is None else np.ascontiguousarray(fragmask, np.uint8).tobytes()
    lib().orc_promolecular_grid(env.ptr, grids.arr, _i(n), fm, _d(out), nthreads)
    return out.reshape(n[2], n[1], n[0])


def promol_nat_points(env: Environ, grids: Grid1Set, pts, nthreads=0, fragmask=None):
    """atoms entering promolecular's radial sum at each point (Cartesian bohr)"""
    pts = _f64(pts)
    out = np.zeros(pts.shape[0], np.int64)
    fm = None if fragmask is None else np.ascontiguousarray(fragmask, np.uint8).tobytes()
    lib().orc_promol_nat_points(env.ptr, grids.arr, pts.shape[0], _d(pts), fm, out.ctypes.data_as(C.POINTER(C.c_long)), nthreads)
    return out


from . import cube_text  # noqa: E402,F401  (formatted-output restatement, pure Python)


def cube_text_c(g, fmt: int, order: int, nthreads: int = 0) -> bytes:
    """C version of cube_text.cube_body (same bytes) for large arrays and the timed CPU baseline."""
    g = _f64(g)
    n = _i32(g.shape[::-1] if order == 0 else g.shape)
    nb = int(n[0]) * int(n[1]) * int(lib().orc_text_record_bytes(int(n[2]), fmt))
    buf = C.create_string_buffer(nb)
    lib().orc_cube_text(_d(g), _i(n), fmt, order, buf, nthreads)
    return buf.raw
