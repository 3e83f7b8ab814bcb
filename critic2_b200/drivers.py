"""Host-side mirror of the two task drivers whose evaluation loops the CUDA module replaces.

* ``cube_grid`` / ``rhoplot_cube`` -- CUBE: grid definition and property selection of
  /root/reference/src/rhoplot@proc.f90:381-447, 536-544 (end point excluded), evaluation loop :582-631.
* ``nci_grid`` / ``nciplot``       -- NCIPLOT: box, step and defaults of /root/reference/src/nci@proc.f90:119-162,
  350-410; evaluation loop :499-561; cutoff pass :569-604.

Only bookkeeping lives here; all field arithmetic is in csrc/ behind the C ABI (critic2_b200.gpu).
"""
from __future__ import annotations

import math
from dataclasses import dataclass

import numpy as np

from . import gpu
from .structure import Structure


def cube_grid(st: Structure, nn, x0=None, x1=None, cell: bool = True):
    """CUBE CELL nx ny nz (cell=True) or CUBE x0 x1 nx ny nz.  Returns (x0 (3,), xd (3,3) with xd[:,i] the i-th step
    vector, nn): rhoplot@proc.f90:412-426, 542-544.  For molecules x0/x1 are real-space Cartesian (bohr) and are
    shifted by molx0 like :401-404."""
    nn = np.maximum(np.asarray(nn, np.int32), 2)
    if cell:
        xd = np.eye(3)
        o = np.zeros(3)
    else:
        x0 = np.asarray(x0, float)
        x1 = np.asarray(x1, float)
        if st.ismolecule:
            x0 = st.m_c2x @ (x0 - st.molx0)
            x1 = st.m_c2x @ (x1 - st.molx0)
        xd = np.diag(x1 - x0)
        o = x0
    o = st.m_x2c @ o
    xd = st.m_x2c @ xd
    xd = xd / nn.astype(float)[None, :]
    return o, xd, nn


def rhoplot_cube(field: "gpu._Field", st: Structure, nn, props=("f",), x0=None, x1=None, cell: bool = True):
    """Evaluate the CUBE properties on the grid; one nder=2 evaluation serves every requested property
    (the reference re-runs the whole loop per CUBE command, :595)."""
    o, xd, nn = cube_grid(st, nn, x0, x1, cell)
    return field.cube(o, xd, nn, props=props, nder=2, periodic=not st.ismolecule), (o, xd, nn)


@dataclass
class NciOptions:
    """Defaults of nci@proc.f90:137-162."""
    rhocut: float = 0.2
    dimcut: float = 2.0
    rhoplot: float = 0.1
    dimplot: float = 0.5
    isden: bool = True            # reference field is a density (not promolecular): :119,137-153
    srhorange: tuple = (-1e30, 1e30)
    onlyneg: bool = False
    rho_void: float = -1.0
    xinc: float = 0.1
    rhoparam: float = 0.95        # fragments: :120-121, used by the `inter` test :579-582
    rhoparam2: float = 0.75

    @staticmethod
    def promolecular():
        return NciOptions(rhocut=0.2, dimcut=1.0, rhoplot=0.12, dimplot=0.3, isden=False)


def nci_grid(st: Structure, nstep=None, xinc: float = 0.1, rthres: float | None = None, frag_xyz=None):
    """NCIPLOT box and step: nci@proc.f90:122-134, 350-410.  Returns (x0, xmat, nstep).
    frag_xyz: list of (nat,3) Cartesian fragment coordinates -> the box is the fragments' bounding box (findlimits = 1,
    :350-364, non-periodic) and rthres defaults to 2 bohr (:159, ithres = 1 :321)."""
    if frag_xyz is not None:
        allr = np.concatenate([np.asarray(r, float).reshape(-1, 3) for r in frag_xyz])
        x0, x1 = allr.min(0), allr.max(0)
        periodic = False
        if rthres is None:
            rthres = 2.0
    elif st.ismolecule:
        x0 = st.m_x2c @ st.molborder
        x1 = st.m_x2c @ (1.0 - st.molborder)
        periodic = False
    else:
        x0 = st.m_x2c @ np.zeros(3)
        x1 = st.m_x2c @ np.ones(3)
        periodic = True
    if rthres is not None:
        x0 = x0 - rthres
        x1 = x1 + rthres
    x0 = st.m_c2x @ x0
    x1 = st.m_c2x @ x1
    xmat = np.zeros((3, 3))
    if periodic:
        ns = np.zeros(3, np.int32)
        for i in range(3):
            e = np.zeros(3)
            e[i] = 1.0
            v = st.m_x2c @ e
            ns[i] = int(nstep[i]) if nstep is not None else math.ceil(np.linalg.norm(v) / xinc)
            xmat[:, i] = v / float(ns[i])
        x0c = st.m_x2c @ x0
    else:
        a = st.m_x2c @ x0
        b = st.m_x2c @ x1
        x0c, x1c = np.minimum(a, b), np.maximum(a, b)
        if nstep is None:
            ns = np.ceil(np.abs(x1c - x0c) / xinc).astype(np.int32)
            inc = np.full(3, xinc)
        else:
            ns = np.asarray(nstep, np.int32)
            inc = (x1c - x0c) / ns.astype(float)
        for i in range(3):
            xmat[i, i] = inc[i]
    return x0c, xmat, ns


def nciplot(field: "gpu._Field", st: Structure, nstep=None, opts: NciOptions | None = None,
            promol: "gpu._Field | None" = None, rthres: float | None = None, fragments=None, frag_xyz=None):
    """NCIPLOT evaluation + cutoff pass.  Returns dict(crho, cgrad, rhoat, rhofrag, sums, dat, ndat, grid).
    fragments: list of gpu.FragmentField with frag_xyz their Cartesian coordinates (FRAGMENT keyword, :320-338): the box
    becomes the fragments' bounding box + rthres, the promolecular density is evaluated (dopromol, :413) and only the
    points between fragments survive the cutoff pass (:579-582)."""
    opts = opts or NciOptions()
    nfrag = 0 if not fragments else len(fragments)
    x0, xmat, ns = nci_grid(st, nstep, opts.xinc, rthres, frag_xyz if nfrag else None)
    use_promol = promol if (opts.rho_void > 0 or nfrag > 0) else None
    if nfrag > 0 and use_promol is None:
        raise ValueError("NCIPLOT with fragments needs the promolecular field (dopromol, nci@proc.f90:413)")
    crho, cgrad, rhoat = field.nci(x0, xmat, ns, promol=use_promol)
    rhofrag = gpu.nci_fragments(fragments, x0, xmat, ns) if nfrag > 0 else None
    cg = cgrad.reshape(-1).copy()
    sums, dat, ndat = gpu.nci_cutoffs(crho.reshape(-1), cg, None if rhoat is None else rhoat.reshape(-1),
                                      rhocut=opts.rhocut, dimcut=opts.dimcut, rhoplot=opts.rhoplot,
                                      srhorange=opts.srhorange, onlyneg=opts.onlyneg, rho_void=opts.rho_void,
                                      omega=st.omega, rhofrag=rhofrag, rhoparam=opts.rhoparam, rhoparam2=opts.rhoparam2)
    return dict(crho=crho, cgrad=cgrad, cgrad_masked=cg.reshape(cgrad.shape), rhoat=rhoat, rhofrag=rhofrag, sums=sums,
                dat=dat, ndat=ndat, grid=(x0, xmat, ns))


# ---- writers: crystal%writegrid_cube (crystalmod@proc.f90:4965-5041) and the NCIPLOT cubes (nci@proc.f90:769-883) ------
def _fe(x: float, w: int, d: int) -> str:
    """Fortran Ew.d (header lines only; the bodies are formatted on the device)."""
    if x == 0.0:
        digits, e10 = "0" * d, -1
    else:
        mant, ex = ("%.*e" % (d - 1, abs(x))).split("e")
        digits, e10 = mant.replace(".", ""), int(ex)
    ex = e10 + 1
    es = ("E%+03d" % ex) if abs(ex) <= 99 else ("%+04d" % ex)
    return (("-" if math.copysign(1.0, x) < 0 else "") + "0." + digits + es).rjust(w)


def cube_header(st: Structure, n, x0=None, xd=None, precise: bool = True) -> bytes:
    """Header of writegrid_cube (:5014-5031): two title lines, natoms + origin, three (n, step vector) lines, atoms."""
    n = [int(v) for v in n]
    if xd is None:
        xd = st.m_x2c / np.asarray(n, float)[None, :]
    x0 = st.molx0 if x0 is None else np.asarray(x0, float)
    out = ["critic2-cube", "critic2-cube"]
    if precise:
        out.append("%5d " % st.ncell + "".join(_fe(float(v), 22, 14) + " " for v in x0))
        for i in range(3):
            out.append("%5d " % n[i] + "".join(_fe(float(v), 22, 14) + " " for v in xd[:, i]))
        for z, r in zip(st.z, st.r):
            out.append("%4d%5.1f" % (int(z), 0.0) + "".join(_fe(float(v), 22, 14) + " " for v in r + st.molx0))
    else:
        out.append("%5d" % st.ncell + "".join("%12.6f" % v for v in x0))
        for i in range(3):
            out.append("%5d" % n[i] + "".join("%12.6f" % v for v in xd[:, i]))
        for z, r in zip(st.z, st.r):
            out.append("%4d%5.1f" % (int(z), 0.0) + "".join("%11.6f" % v for v in r + st.molx0))
    return ("\n".join(out) + "\n").encode()


def writegrid_cube(path: str, st: Structure, g, x0=None, xd=None, precise: bool = True) -> int:
    """Write the cube file of a rhoplot_cube property array g[iz][iy][ix] (lf(ix,iy,iz)).  The body -- one record per
    (ix, iy), iz fastest, 6 values per line (:5026-5036) -- is formatted by the CUDA module; returns the bytes written."""
    g = np.ascontiguousarray(g, float)
    n = g.shape[::-1]
    body = gpu.cube_text(g, gpu.FMT_PRECISE if precise else gpu.FMT_CUBE, order=0)
    head = cube_header(st, n, x0, xd, precise)
    with open(path, "wb") as fh:
        fh.write(head)
        fh.write(body)
    return len(head) + len(body)


def write_nci_cube_body(fh, c) -> int:
    """write_cube_body (nci@proc.f90:867-883): c[i][j][k] = crho/cgrad(k,j,i), '(6(1x,1p,e13.5e3))'; the header written
    by write_cube_header stays with the caller."""
    body = gpu.cube_text(np.ascontiguousarray(c, float), gpu.FMT_NCI, order=1)
    fh.write(body)
    return len(body)


def write_nci_dat(path: str, dat) -> int:
    """The .dat file of NCIPLOT: one '(1p,E15.7,E15.7)' line per (rho, s) pair (nci@proc.f90:593-595)."""
    text = gpu.dat_text(dat)
    with open(path, "wb") as fh:
        fh.write(text)
    return len(text)


# ---- NCIPLOT checkpoint (.chk_nci): writechk / readchk, nci@proc.f90:962-1059 -------------------------------------
# gfortran unformatted sequential records: <int32 nbytes> payload <int32 nbytes>; a record longer than 2^31-9 bytes is
# split into subrecords whose head marker is negative when another subrecord follows and whose tail marker is negative
# when one precedes (libgfortran io/unix.c); default LOGICAL and INTEGER are 4 bytes.
_SUBREC = 2147483639


def _f_record(fh, payload) -> None:
    mv = memoryview(payload).cast("B")
    n, pos, first = len(mv), 0, True
    while True:
        m = min(_SUBREC, n - pos)
        more = pos + m < n
        fh.write(np.int32(-m if more else m).tobytes())
        fh.write(mv[pos:pos + m])
        fh.write(np.int32(m if first else -m).tobytes())
        pos, first = pos + m, False
        if not more:
            return


def _f_read_record(fh) -> bytes:
    out = []
    while True:
        head = int(np.frombuffer(fh.read(4), np.int32)[0])
        out.append(fh.read(abs(head)))
        tail = int(np.frombuffer(fh.read(4), np.int32)[0])
        if abs(tail) != abs(head):
            raise ValueError("corrupt unformatted record")
        if head >= 0:
            return b"".join(out)


def write_chk_nci(path: str, x0, xmat, nstep, crho, cgrad, rhoat=None, rhofrag=None) -> None:
    """writechk (nci@proc.f90:1027-1059).  crho/cgrad/rhoat indexed [i][j][k] (the reference's (k,j,i) memory order),
    rhofrag [ifr][i][j][k]; xmat[:, i] is the i-th step vector."""
    nfrag = 0 if rhofrag is None else int(np.asarray(rhofrag).shape[0])
    xmat = np.asarray(xmat, float)
    with open(path, "wb") as fh:
        _f_record(fh, np.array([rhoat is not None, rhofrag is not None, nfrag], np.int32))
        _f_record(fh, np.ascontiguousarray(x0, float))
        for i in range(3):
            _f_record(fh, np.ascontiguousarray(xmat[:, i]))
        _f_record(fh, np.ascontiguousarray(nstep, np.int32))
        _f_record(fh, np.concatenate([np.ascontiguousarray(crho, float).reshape(-1),
                                      np.ascontiguousarray(cgrad, float).reshape(-1)]))
        if rhoat is not None:
            _f_record(fh, np.ascontiguousarray(rhoat, float))
        if rhofrag is not None:
            _f_record(fh, np.ascontiguousarray(rhofrag, float))


def read_chk_nci(path: str, x0, xmat, nstep, want_rhoat: bool = False, nfrag: int = 0):
    """readchk (nci@proc.f90:962-1024): returns (crho, cgrad, rhoat, rhofrag) or None when the file does not match the
    current lattice / options (the reference then restarts the evaluation)."""
    xmat = np.asarray(xmat, float)
    ns = [int(v) for v in nstep]
    with open(path, "rb") as fh:
        llat, llfrag, nfrag0 = (int(v) for v in np.frombuffer(_f_read_record(fh), np.int32))
        if (not llat and want_rhoat) or (not llfrag and nfrag > 0) or nfrag0 != nfrag:
            return None
        chk = [np.asarray(x0, float)] + [xmat[:, i] for i in range(3)]
        for ref in chk:
            if np.any(np.abs(np.frombuffer(_f_read_record(fh), np.float64) - ref) > 1e-13):
                return None
        if list(np.frombuffer(_f_read_record(fh), np.int32)) != ns:
            return None
        npt = ns[0] * ns[1] * ns[2]
        both = np.frombuffer(_f_read_record(fh), np.float64)
        crho, cgrad = both[:npt].reshape(ns).copy(), both[npt:].reshape(ns).copy()
        rhoat = rhofrag = None
        if llat and want_rhoat:
            rhoat = np.frombuffer(_f_read_record(fh), np.float64).reshape(ns).copy()
            if llfrag and nfrag > 0:
                rhofrag = np.frombuffer(_f_read_record(fh), np.float64).reshape([nfrag] + ns).copy()
    return crho, cgrad, rhoat, rhofrag


# ---- point-list callers of field%grd: LINE, PLANE, mesh fill -----------------------------------
_RES_OF_NTI = {0: 0, 1: 1, 2: 2, 3: 3, 4: 13, 5: 4, 6: 7, 7: 10, 8: 8, 9: 11, 10: 12, 11: 14}   # res column of nti


def _nder_of(nti: int) -> int:
    return 0 if nti == 0 else (1 if nti <= 4 else 2)          # rhoplot@proc.f90:279-285, 891-897


def rhoplot_line(field: "gpu._Field", st: Structure, x0, x1, npts: int, prop: str = "f"):
    """LINE x0 x1 npts [prop] (rhoplot@proc.f90:276-325): x0, x1 Cartesian (cell frame, bohr).  Returns
    (xp (npts,3), rhoout, lapout): rhoout = f, lapout = the selected property, like the reference's two columns."""
    nti = gpu.PROPS.index(prop)
    x0, x1 = np.asarray(x0, float), np.asarray(x1, float)
    t = np.arange(npts, dtype=float) / float(npts - 1)
    xp = x0[None, :] + (x1 - x0)[None, :] * t[:, None]          # :278
    res = field.points(xp, nder=_nder_of(nti), periodic=not st.ismolecule)
    return xp, res[:, 0].copy(), res[:, _RES_OF_NTI[nti]].copy()


def rhoplot_plane(field: "gpu._Field", st: Structure, x0, x1, x2, nx: int, ny: int, prop: str = "f"):
    """PLANE x0 x1 x2 nx ny [prop] (rhoplot@proc.f90:880-940, without SCALE/EXTENDXY): ff(nx,ny) with
    xp = x0 + (ix-1) uu + (iy-1) vv, uu = (x1-x0)/(nx-1), vv = (x2-x0)/(ny-1).  Returns ff (nx, ny)."""
    nti = gpu.PROPS.index(prop)
    x0, x1, x2 = (np.asarray(v, float) for v in (x0, x1, x2))
    uu, vv = (x1 - x0) / float(nx - 1), (x2 - x0) / float(ny - 1)
    ix, iy = np.meshgrid(np.arange(nx, dtype=float), np.arange(ny, dtype=float), indexing="ij")
    xp = x0[None, None, :] + ix[..., None] * uu + iy[..., None] * vv
    res = field.points(xp.reshape(-1, 3), nder=_nder_of(nti), periodic=not st.ismolecule)
    return res[:, _RES_OF_NTI[nti]].reshape(nx, ny)


def mesh_fill(field: "gpu._Field", pts, props=("rho",), periodic: bool = True):
    """mesh%fill (meshmod@proc.f90:420-470).  props: 'volume' (1), 'rho' (res%f), 'gradrho' (res%gfmod), 'gkin' (res%gkin,
    wavefunction fields) or an integer > 100 = MO number prop - 100 (res%fspc of grd(nder=-1), wavefunction fields).
    pts (n,3) Cartesian mesh nodes; returns f (n, len(props))."""
    nders = {"volume": -1, "rho": 0, "gradrho": 1, "gkin": -1}
    nder = max([-1] + [nders[p] for p in props if not isinstance(p, int)])
    pts = np.asarray(pts, float).reshape(-1, 3)
    res = field.points(pts, nder=nder, periodic=periodic) if nder >= 0 else None
    cols = []
    for p in props:
        if isinstance(p, int):
            if p <= 100:
                raise ValueError("MO properties are 100 + n")
            cols.append(field.mo_points(pts, p - 100, periodic))
        elif p == "volume":
            cols.append(np.ones(pts.shape[0]))
        elif p == "rho":
            cols.append(res[:, 0])
        elif p == "gradrho":
            cols.append(res[:, 13])
        else:
            cols.append(field.gkin_points(pts, periodic))
    return np.stack(cols, axis=1)
