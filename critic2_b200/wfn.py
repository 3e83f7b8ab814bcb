"""Host-side mirror of critic2's ``molwfn`` loaders (the step *before* the hot path).

Restates, in numpy, the reference's wavefunction readers so that tests and benches can
feed the same flat arrays to both the CUDA module and the oracle:

* ``read_fchk``  -- /root/reference/src/wfn_private@proc.F90:920-1434
* ``read_wfx``   -- :774-917 (+ helpers :2934-3012)
* ``prune_primitives`` -- :3017-3069
* ``complete_struct``  -- :3077-3184 (dran, spcutoff, lmax, icord, iprilo/iprihi)
* ``gnorm`` -- :2877-2931

Nothing here evaluates a field; evaluation lives in csrc/ (CUDA) only.
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field

import numpy as np

WFN_RHF, WFN_UHF, WFN_FRAC = 0, 1, 2  # wfn_private.f90:142-144
RPRIM_THRES = 1e-12  # wfn_private@proc.F90:145

# (lx,ly,lz) per primitive type 1..56, wfn_private@proc.F90:2695-2705
LI = np.array([
    (0, 0, 0),
    (1, 0, 0), (0, 1, 0), (0, 0, 1),
    (2, 0, 0), (0, 2, 0), (0, 0, 2), (1, 1, 0), (1, 0, 1), (0, 1, 1),
    (3, 0, 0), (0, 3, 0), (0, 0, 3), (2, 1, 0), (2, 0, 1), (0, 2, 1), (1, 2, 0), (1, 0, 2), (0, 1, 2), (1, 1, 1),
    (4, 0, 0), (0, 4, 0), (0, 0, 4), (3, 1, 0), (3, 0, 1), (1, 3, 0), (0, 3, 1), (1, 0, 3),
    (0, 1, 3), (2, 2, 0), (2, 0, 2), (0, 2, 2), (2, 1, 1), (1, 2, 1), (1, 1, 2),
    (0, 0, 5), (0, 1, 4), (0, 2, 3), (0, 3, 2), (0, 4, 1), (0, 5, 0), (1, 0, 4), (1, 1, 3),
    (1, 2, 2), (1, 3, 1), (1, 4, 0), (2, 0, 3), (2, 1, 2), (2, 2, 1), (2, 3, 0), (3, 0, 2),
    (3, 1, 1), (3, 2, 0), (4, 0, 1), (4, 1, 0), (5, 0, 0)], dtype=np.int32)
LIMAX = LI.max(axis=1).astype(np.int32)  # max single-axis power; equals the limax table at :3084-3086

# :31-42
DFACM1 = [1, 1, 1, 2, 3, 8, 15, 48, 105]
NSHLT = {-4: 9, -3: 7, -2: 5, -1: 3, 0: 1, 1: 3, 2: 6, 3: 10, 4: 15}
JSHL0 = {-4: 17, -3: 10, -2: 5, -1: 2, 0: 1, 1: 2, 2: 5, 3: 11, 4: 21}
JSHL1 = {-4: 25, -3: 16, -2: 9, -1: 4, 0: 1, 1: 4, 2: 10, 3: 20, 4: 35}

# typtrans(fchk) = critic2, :953-958
TYPTRANS = [1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11, 12, 13, 17, 14, 15, 18, 19, 16, 20,
            23, 29, 32, 27, 22, 28, 35, 34, 26, 31, 33, 30, 25, 24, 21]


def _sphcar_tables():
    """dsphcar, fsphcar, gsphcar_fchk (:72-142). Stored as (nsph, ncar): column c is the Cartesian function."""
    s = math.sqrt
    s3, s3_4, s3_8, s5_8, s5_16 = s(3.), s(3. / 4), s(3. / 8), s(5. / 8), s(5. / 16)
    s6, s10, s10_8, s15, s15_4 = s(6.), s(10.), s(10. / 8), s(15.), s(15. / 4)
    s35_4, s35_8, s35_64, s45, s45_4, s45_8 = s(35. / 4), s(35. / 8), s(35. / 64), s(45.), s(45. / 4), s(45. / 8)
    s315_8, s315_16 = s(315. / 8), s(315. / 16)
    d32, d34, d38 = 1.5, 0.75, 0.375
    d = np.array([
        [-0.5, 0, 0, s3_4, 0], [-0.5, 0, 0, -s3_4, 0], [1, 0, 0, 0, 0],
        [0, 0, 0, 0, s3], [0, s3, 0, 0, 0], [0, 0, s3, 0, 0]]).T
    f = np.array([
        [0, -s3_8, 0, 0, 0, s5_8, 0], [0, 0, -s3_8, 0, 0, 0, -s5_8], [1, 0, 0, 0, 0, 0, 0],
        [0, -s3_8, 0, 0, 0, -s45_8, 0], [0, 0, -s3_8, 0, 0, 0, s45_8], [-d32, 0, 0, s15_4, 0, 0, 0],
        [0, s6, 0, 0, 0, 0, 0], [0, 0, s6, 0, 0, 0, 0], [-d32, 0, 0, -s15_4, 0, 0, 0],
        [0, 0, 0, 0, s15, 0, 0]]).T
    g = np.array([
        [1, 0, 0, 0, 0, 0, 0, 0, 0],                     # zzzz
        [0, 0, s10, 0, 0, 0, 0, 0, 0],                   # yzzz
        [-3, 0, 0, -s45_4, 0, 0, 0, 0, 0],               # yyzz
        [0, 0, -s45_8, 0, 0, 0, -s35_8, 0, 0],           # yyyz
        [d38, 0, 0, s5_16, 0, 0, 0, s35_64, 0],          # yyyy
        [0, s10, 0, 0, 0, 0, 0, 0, 0],                   # xzzz
        [0, 0, 0, 0, s45, 0, 0, 0, 0],                   # xyzz
        [0, -s45_8, 0, 0, 0, -s315_8, 0, 0, 0],          # xyyz
        [0, 0, 0, 0, -s10_8, 0, 0, 0, -s35_4],           # xyyy
        [-3, 0, 0, s45_4, 0, 0, 0, 0, 0],                # xxzz
        [0, 0, -s45_8, 0, 0, 0, s315_8, 0, 0],           # xxyz
        [d34, 0, 0, 0, 0, 0, 0, -s315_16, 0],            # xxyy
        [0, -s45_8, 0, 0, 0, s35_8, 0, 0, 0],            # xxxz
        [0, 0, 0, 0, -s10_8, 0, 0, 0, s35_4],            # xxxy
        [d38, 0, 0, -s5_16, 0, 0, 0, s35_64, 0]]).T      # xxxx
    return {-2: d, -3: f, -4: g}


SPHCAR = _sphcar_tables()


def gnorm(ityp: int, a: float) -> float:
    """Primitive normalisation factor, wfn_private@proc.F90:2877-2931."""
    pi34 = math.pi ** 0.75
    if ityp == 1:
        return 2 ** 0.75 * a ** 0.75 / pi34
    if 2 <= ityp <= 4:
        return 2 ** 1.75 * a ** 1.25 / pi34
    if 5 <= ityp <= 7:
        return 2 ** 2.75 * a ** 1.75 / pi34 / math.sqrt(3.)
    if 8 <= ityp <= 10:
        return 2 ** 2.75 * a ** 1.75 / pi34
    if 11 <= ityp <= 13:
        return 2 ** 3.75 * a ** 2.25 / pi34 / math.sqrt(15.)
    if 14 <= ityp <= 19:
        return 2 ** 3.75 * a ** 2.25 / pi34 / math.sqrt(3.)
    if ityp == 20:
        return 2 ** 3.75 * a ** 2.25 / pi34
    if 21 <= ityp <= 23:
        return 2 ** 4.75 * a ** 2.75 / pi34 / math.sqrt(35.)
    if 24 <= ityp <= 29:
        return 2 ** 4.75 * a ** 2.75 / pi34 / math.sqrt(5.)
    if 30 <= ityp <= 32:
        return 2 ** 4.75 * a ** 2.75 / pi34 * math.sqrt(3.)
    if 33 <= ityp <= 35:
        return 2 ** 4.75 * a ** 2.75 / pi34
    raise ValueError("gnorm: primitive type not supported")


@dataclass
class Molwfn:
    """Flat-array image of ``type molwfn`` (wfn_private.f90:83-131). Indices are 1-based like the reference."""
    natom: int
    xyz: np.ndarray            # (natom,3) bohr, as read from the file (before the molecular-cell shift)
    z: np.ndarray              # (natom,) atomic numbers
    icenter: np.ndarray        # (npri,) int32, 1-based
    itype: np.ndarray          # (npri,) int32, 1..56
    e: np.ndarray              # (npri,)
    occ: np.ndarray            # (nmoocc,)
    cmo: np.ndarray            # (npri, nmoall) C-order == Fortran cmo(nmoall,npri) column-major
    nmoocc: int
    wfntyp: int = WFN_RHF
    nalpha: int = 0
    icenter_edf: np.ndarray = field(default_factory=lambda: np.zeros(0, np.int32))
    itype_edf: np.ndarray = field(default_factory=lambda: np.zeros(0, np.int32))
    e_edf: np.ndarray = field(default_factory=lambda: np.zeros(0))
    c_edf: np.ndarray = field(default_factory=lambda: np.zeros(0))
    # complete_struct outputs
    dran: np.ndarray | None = None
    spcutoff: np.ndarray | None = None
    lmax: np.ndarray | None = None
    icord: np.ndarray | None = None
    iprilo: np.ndarray | None = None
    iprihi: np.ndarray | None = None
    dran_edf: np.ndarray | None = None

    @property
    def npri(self) -> int:
        return int(self.e.shape[0])

    @property
    def nmoall(self) -> int:
        return int(self.cmo.shape[1])

    @property
    def nedf(self) -> int:
        return int(self.e_edf.shape[0])


def prune_primitives(f: Molwfn) -> None:
    """Merge primitives with equal exponent (1e-12), type and centre: wfn_private@proc.F90:3017-3069."""
    n = f.npri
    iord = np.argsort(f.e, kind="stable")
    keep = np.ones(n, bool)
    es = f.e[iord]
    for a in range(n):
        i = iord[a]
        if not keep[i]:
            continue
        b = a + 1
        # the reference scans all j>i and tests |e_j - e_i| <= eps; sorted order lets us stop early
        while b < n and es[b] - es[a] <= 1e-12:
            j = iord[b]
            if keep[j] and f.itype[i] == f.itype[j] and f.icenter[i] == f.icenter[j]:
                f.cmo[i, :] += f.cmo[j, :]
                keep[j] = False
            b += 1
    f.cmo = np.ascontiguousarray(f.cmo[keep])
    f.e = np.ascontiguousarray(f.e[keep])
    f.icenter = np.ascontiguousarray(f.icenter[keep])
    f.itype = np.ascontiguousarray(f.itype[keep])


def complete_struct(f: Molwfn) -> None:
    """dran/spcutoff/lmax/icord/iprilo/iprihi: wfn_private@proc.F90:3077-3184 (molecular environment)."""
    f.dran = np.sqrt(-math.log(RPRIM_THRES) / f.e)
    f.spcutoff = np.zeros(f.natom)
    f.lmax = np.zeros(f.natom, np.int32)
    np.maximum.at(f.spcutoff, f.icenter - 1, f.dran)
    np.maximum.at(f.lmax, f.icenter - 1, LIMAX[f.itype - 1])
    f.icord = (np.argsort(f.icenter, kind="stable") + 1).astype(np.int32)
    f.iprilo = np.zeros(f.natom, np.int32)
    f.iprihi = np.full(f.natom, -1, np.int32)
    cs = f.icenter[f.icord - 1]
    for i, iat in enumerate(cs):
        if f.iprilo[iat - 1] == 0:
            f.iprilo[iat - 1] = i + 1
        f.iprihi[iat - 1] = i + 1
    if f.nedf > 0:
        f.dran_edf = np.sqrt(-math.log(RPRIM_THRES) / f.e_edf)
        np.maximum.at(f.spcutoff, f.icenter_edf - 1, f.dran_edf)
    else:
        f.dran_edf = np.zeros(0)


# --------------------------------------------------------------------------------------
# readers
# --------------------------------------------------------------------------------------
def _fchk_sections(path: str):
    """Parse a Gaussian formatted checkpoint into {label: scalar | ndarray}."""
    out = {}
    with open(path) as fh:
        lines = fh.read().split("\n")
    i = 2
    nl = len(lines)
    while i < nl:
        line = lines[i]
        i += 1
        if len(line) < 44:
            continue
        label = line[:40].rstrip()
        typ = line[43:44]
        rest = line[44:]
        if "N=" in rest:
            n = int(rest.split("N=")[1])
            if typ == "C":
                per, conv = 5, str
            elif typ == "I":
                per, conv = 6, int
            elif typ == "R":
                per, conv = 5, float
            else:
                per, conv = 72, str
            nrows = (n + per - 1) // per
            vals = []
            for r in range(nrows):
                vals.extend(lines[i + r].split())
            i += nrows
            if typ in "IR":
                out[label] = np.array([conv(v) for v in vals[:n]], dtype=np.int64 if typ == "I" else np.float64)
        else:
            tok = rest.split()
            if not tok:
                continue
            if typ == "I":
                out[label] = int(tok[0])
            elif typ == "R":
                out[label] = float(tok[0])
    return out


def read_fchk(path: str) -> Molwfn:
    """LOAD file.fchk (occupied MOs only, readvirtual=.false.): wfn_private@proc.F90:920-1434."""
    s = _fchk_sections(path)
    nelec = s["Number of electrons"]
    nalpha_e = s["Number of alpha electrons"]
    nbassph = s["Number of basis functions"]
    uhf = "Beta Orbital Energies" in s
    if not uhf:
        if nelec % 2:
            raise ValueError("read_fchk: odd nelec but closed-shell wavefunction")
        nmoocc, nalpha = nelec // 2, nelec // 2
        nmoread, namoread = nmoocc, nalpha
        wfntyp, occ = WFN_RHF, np.full(nmoocc, 2.0)
    else:
        nmoocc, nalpha = nelec, nalpha_e
        nmoread, namoread = nmoocc, nalpha
        wfntyp, occ = WFN_UHF, np.full(nmoocc, 1.0)

    ishlt = s["Shell types"].astype(int)
    ishlpri = s["Number of primitives per shell"].astype(int)
    ishlat = s["Shell to atom map"].astype(int)
    exppri = s["Primitive exponents"]
    ccontr = s["Contraction coefficients"]
    pccontr = s.get("P(S=P) Contraction coefficients", np.zeros_like(ccontr))
    # MO coefficients: motemp((i-1)*nbassph + mu) -> [imo, mu]
    mot = np.zeros((nmoread, nbassph))
    a = s["Alpha MO coefficients"].reshape(-1, nbassph)
    mot[:namoread] = a[:namoread]
    if uhf:
        b = s["Beta MO coefficients"].reshape(-1, nbassph)
        mot[namoread:] = b[: nmoread - namoread]

    # unfold SP shells :1116-1174
    t2, p2, a2, e2, c2 = [], [], [], [], []
    nm = 0
    for t, npr, at in zip(ishlt, ishlpri, ishlat):
        ex = exppri[nm:nm + npr]
        if t == -1:
            t2 += [0, 1]; p2 += [npr, npr]; a2 += [at, at]
            e2 += [ex, ex]; c2 += [ccontr[nm:nm + npr], pccontr[nm:nm + npr]]
        else:
            t2.append(t); p2.append(npr); a2.append(at); e2.append(ex); c2.append(ccontr[nm:nm + npr])
        nm += npr
    if any(abs(t) > 4 for t in t2):
        raise ValueError("read_fchk: h primitives not supported in fchk")

    # spherical -> Cartesian MO coefficients :1192-1225
    cols, ns = [], 0
    for t in t2:
        nsph = NSHLT[t]
        blk = mot[:, ns:ns + nsph]
        if nsph == NSHLT[abs(t)]:
            cols.append(blk)
        else:
            cols.append(blk @ SPHCAR[t])
        ns += nsph
    mocoef = np.concatenate(cols, axis=1)

    # primitive coefficients :1238-1286 and assignment :1404-1421
    icenter, itype, e, cm = [], [], [], []
    nl = 0
    for t, npr, at, ex, cc in zip(t2, p2, a2, e2, c2):
        L = abs(t)
        for j in range(JSHL0[L], JSHL1[L] + 1):
            ityp = TYPTRANS[j - 1]
            cnorm = np.array([cc[k] * gnorm(ityp, ex[k]) for k in range(npr)])
            norm = 0.0
            for k1 in range(npr):
                for k2 in range(npr):
                    norm += cnorm[k1] * cnorm[k2] / (ex[k1] + ex[k2]) ** (L + 1.5)
            cons = math.pi ** 1.5 * DFACM1[2 * L] / 2 ** L
            norm = 1.0 / math.sqrt(norm * cons)
            if t >= 0:  # Cartesian shells only :1262-1276
                if 8 <= ityp <= 10:
                    norm *= math.sqrt(3.)
                elif 14 <= ityp <= 19:
                    norm *= math.sqrt(5.)
                elif ityp == 20:
                    norm *= math.sqrt(15.)
                elif 24 <= ityp <= 29:
                    norm *= math.sqrt(7.)
                elif 30 <= ityp <= 32:
                    norm *= math.sqrt(35. / 3.)
                elif 33 <= ityp <= 35:
                    norm *= math.sqrt(35.)
            for k in range(npr):
                icenter.append(at); itype.append(ityp); e.append(ex[k])
                cm.append(cnorm[k] * norm * mocoef[:, nl])
            nl += 1

    xyz = s["Current cartesian coordinates"].reshape(-1, 3)
    f = Molwfn(natom=xyz.shape[0], xyz=xyz.copy(), z=s["Atomic numbers"].astype(np.int32),
               icenter=np.array(icenter, np.int32), itype=np.array(itype, np.int32), e=np.array(e, float),
               occ=occ, cmo=np.ascontiguousarray(np.array(cm)), nmoocc=nmoocc, wfntyp=wfntyp, nalpha=nalpha)
    prune_primitives(f)
    complete_struct(f)
    return f


def _wfx_block(lines, tag):
    try:
        i0 = lines.index("<" + tag + ">")
    except ValueError:
        return None
    i1 = lines.index("</" + tag + ">", i0)
    return lines[i0 + 1:i1]


def read_wfx(path: str) -> Molwfn:
    """LOAD file.wfx: wfn_private@proc.F90:774-917."""
    with open(path) as fh:
        lines = [l.strip() for l in fh]

    def ints(tag):
        b = _wfx_block(lines, tag)
        return None if b is None else np.array(" ".join(b).split(), dtype=np.int64)

    def reals(tag):
        b = _wfx_block(lines, tag)
        return None if b is None else np.array(" ".join(b).replace("D", "E").split(), dtype=np.float64)

    nmoocc = int(ints("Number of Occupied Molecular Orbitals")[0])
    npri = int(ints("Number of Primitives")[0])
    nal = ints("Number of Alpha Electrons")
    nalpha = int(nal[0]) if nal is not None else 0
    ned = ints("Number of EDF Primitives")
    nedf = int(ned[0]) if ned is not None else 0
    icenter = ints("Primitive Centers").astype(np.int32)
    itype = ints("Primitive Types").astype(np.int32)
    if (itype > 56).any():
        raise ValueError("read_wfx: primitive type not supported")
    e = reals("Primitive Exponents")
    occ = reals("Molecular Orbital Occupation Numbers")
    blk = _wfx_block(lines, "Molecular Orbital Primitive Coefficients")
    vals = []
    skip = False
    for l in blk:
        if l.startswith("<MO Number>"):
            skip = True
            continue
        if l.startswith("</MO Number>"):
            skip = False
            continue
        if not skip and l:
            vals.extend(l.replace("D", "E").split())
    cmo = np.array(vals, dtype=np.float64).reshape(nmoocc, npri)
    xyz = reals("Nuclear Cartesian Coordinates").reshape(-1, 3)
    z = ints("Atomic Numbers").astype(np.int32)
    # wavefunction type :879-906
    ioc = np.rint(occ)
    if (np.abs(ioc - occ) > 1e-10).any():
        wfntyp, nalpha = WFN_FRAC, 0
    elif not (ioc == 1).any():
        wfntyp = WFN_RHF
    elif not (ioc == 2).any():
        wfntyp = WFN_UHF
    else:
        raise ValueError("read_wfx: restricted-open wfx files not supported")
    f = Molwfn(natom=xyz.shape[0], xyz=xyz, z=z, icenter=icenter, itype=itype, e=e, occ=occ,
               cmo=np.ascontiguousarray(cmo.T), nmoocc=nmoocc, wfntyp=wfntyp, nalpha=nalpha)
    if nedf > 0:
        f.icenter_edf = ints("EDF Primitive Centers").astype(np.int32)
        f.itype_edf = ints("EDF Primitive Types").astype(np.int32)
        f.e_edf = reals("EDF Primitive Exponents")
        f.c_edf = reals("EDF Primitive Coefficients")
    prune_primitives(f)
    complete_struct(f)
    return f


def write_wfx(f: Molwfn, path: str, title: str = "synthetic") -> None:
    """Write the primitive-level arrays as a WFX file the reference's read_wfx accepts (:774-917)."""
    def block(fh, tag, arr, fmt, per):
        fh.write(f"<{tag}>\n")
        for i in range(0, len(arr), per):
            fh.write(" ".join(fmt % v for v in arr[i:i + per]) + "\n")
        fh.write(f"</{tag}>\n")
    with open(path, "w") as fh:
        fh.write(f"<Title>\n {title}\n</Title>\n<Keywords>\n GTO\n</Keywords>\n")
        fh.write(f"<Number of Nuclei>\n {f.natom}\n</Number of Nuclei>\n")
        fh.write(f"<Number of Occupied Molecular Orbitals>\n {f.nmoocc}\n</Number of Occupied Molecular Orbitals>\n")
        nel = float(f.occ.sum())
        fh.write(f"<Number of Electrons>\n {int(round(nel))}\n</Number of Electrons>\n")
        na = f.nalpha if f.wfntyp == WFN_UHF else int(round(nel)) // 2
        fh.write(f"<Number of Alpha Electrons>\n {na}\n</Number of Alpha Electrons>\n")
        fh.write("<Number of Core Electrons>\n 0\n</Number of Core Electrons>\n")
        block(fh, "Atomic Numbers", f.z, "%d", 1)
        block(fh, "Nuclear Cartesian Coordinates", f.xyz.reshape(-1), "%.14E", 3)
        fh.write(f"<Number of Primitives>\n {f.npri}\n</Number of Primitives>\n")
        block(fh, "Primitive Centers", f.icenter, "%d", 20)
        block(fh, "Primitive Types", f.itype, "%d", 20)
        block(fh, "Primitive Exponents", f.e, "%.14E", 5)
        block(fh, "Molecular Orbital Occupation Numbers", f.occ, "%.14E", 5)
        fh.write("<Molecular Orbital Primitive Coefficients>\n")
        for i in range(f.nmoocc):
            fh.write(f"<MO Number>\n {i + 1}\n</MO Number>\n")
            c = f.cmo[:, i]
            for k in range(0, len(c), 5):
                fh.write(" ".join("%.14E" % v for v in c[k:k + 5]) + "\n")
        fh.write("</Molecular Orbital Primitive Coefficients>\n")
        if f.nedf > 0:
            fh.write(f"<Number of EDF Primitives>\n {f.nedf}\n</Number of EDF Primitives>\n")
            block(fh, "EDF Primitive Centers", f.icenter_edf, "%d", 20)
            block(fh, "EDF Primitive Types", f.itype_edf, "%d", 20)
            block(fh, "EDF Primitive Exponents", f.e_edf, "%.14E", 5)
            block(fh, "EDF Primitive Coefficients", f.c_edf, "%.14E", 5)


# ---- AIMPAC .wfn: read_wfn, wfn_private@proc.F90:648-771 (fixed formats 101-106 at :764-769) ---------------------------
_ELEM = ("H He Li Be B C N O F Ne Na Mg Al Si P S Cl Ar K Ca Sc Ti V Cr Mn Fe Co Ni Cu Zn Ga Ge As Se Br Kr Rb Sr Y Zr Nb "
         "Mo Tc Ru Rh Pd Ag Cd In Sn Sb Te I Xe").split()


def _fortran_e(x: float, w: int, d: int) -> str:
    """Fortran Ew.d edit descriptor (0.dddddE+xx); 'D' exponents are accepted on input (Gaussian writes them)."""
    if x == 0.0:
        return ("0." + "0" * d + "E+00").rjust(w)
    m, ex = ("%.*E" % (d - 1, x)).split("E")
    digits = m.replace(".", "").replace("-", "")
    return (("-" if x < 0 else "") + "0." + digits + "E%+03d" % (int(ex) + 1)).rjust(w)


def read_wfn(path: str) -> Molwfn:
    """LOAD file.wfn.  Columns follow the reference's formats: 101 (4X,A4,10X,3(I5,15X)), 106 (2X,A2,20X,3F12.8,10X,F5.1),
    102 (20X,20I3), 103 (10X,5E14.7), 104 (35X,F12.7,15X,F12.6), 105 (5E16.8)."""
    with open(path) as fh:
        lines = fh.read().split("\n")
    num = lambda s: float(s.replace("D", "E").replace("d", "E"))
    hdr = lines[1]
    nmoocc, npri, nat = int(hdr[18:23]), int(hdr[38:43]), int(hdr[58:63])
    xyz, z = np.zeros((nat, 3)), np.zeros(nat, np.int32)
    for i in range(nat):
        l = lines[2 + i]
        xyz[i] = [num(l[24 + 12 * k:36 + 12 * k]) for k in range(3)]
        z[i] = int(round(num(l[70:75])))              # the nuclear charge column (ECP files: element symbol wins, :681-682)
    pos = 2 + nat

    def fixed(n, skip, width, per, conv):
        nonlocal pos
        out = []
        while len(out) < n:
            l = lines[pos][skip:]
            pos += 1
            out.extend(conv(l[k * width:(k + 1) * width]) for k in range(min(per, n - len(out))))
        return out

    icenter = np.array(fixed(npri, 20, 3, 20, int), np.int32)
    itype = np.array(fixed(npri, 20, 3, 20, int), np.int32)
    if (itype > 56).any():
        raise ValueError("read_wfn: primitive type not supported")
    e = np.array(fixed(npri, 10, 14, 5, num))
    while not lines[pos].startswith("MO"):              # ECP records, :707-711
        pos += 1
    occ, ene, cmo = np.zeros(nmoocc), np.zeros(nmoocc), np.zeros((npri, nmoocc))
    for i in range(nmoocc):
        l = lines[pos]
        pos += 1
        occ[i], ene[i] = num(l[35:47]), num(l[62:74])
        cmo[:, i] = fixed(npri, 0, 16, 5, num)
    ioc = np.rint(occ)
    nalpha, found = -1, False
    for i in range(1, nmoocc):                          # alpha/beta boundary: orbital energy drops by > 1e-3 (:741-745)
        if ene[i] < ene[i - 1] - 1e-3:
            nalpha, found = i, True
    if (np.abs(ioc - occ) > 1e-10).any():
        wfntyp, nalpha = WFN_FRAC, 0
    elif not (ioc == 1).any():
        wfntyp = WFN_RHF
    elif not (ioc == 2).any():
        wfntyp = WFN_UHF
        if not found:
            nalpha = nmoocc
    else:
        raise ValueError("read_wfn: restricted-open wfn files not supported")
    f = Molwfn(natom=nat, xyz=xyz, z=z, icenter=icenter, itype=itype, e=e, occ=occ, cmo=np.ascontiguousarray(cmo),
               nmoocc=nmoocc, wfntyp=wfntyp, nalpha=nalpha)
    prune_primitives(f)
    complete_struct(f)
    return f


def write_wfn(f: Molwfn, path: str, energies=None, title: str = "synthetic") -> None:
    """Write an AIMPAC .wfn the reference's read_wfn accepts (single-precision text: E14.7 exponents, E16.8 coefficients)."""
    ene = np.arange(f.nmoocc, dtype=float) * 0.01 - 10.0 if energies is None else np.asarray(energies, float)
    with open(path, "w") as fh:
        fh.write(title + "\n")
        fh.write("GAUSSIAN".ljust(18) + "%5d" % f.nmoocc + " MOL ORBITALS  " + "%5d" % f.npri + " PRIMITIVES    "
                 + "%5d" % f.natom + " NUCLEI\n")
        for i in range(f.natom):
            el = _ELEM[int(f.z[i]) - 1]
            fh.write("  %-2s%4d    (CENTRE%3d) " % (el, i + 1, i + 1) + "".join("%12.8f" % v for v in f.xyz[i])
                     + "  CHARGE =%5.1f\n" % float(f.z[i]))
        for name, arr in (("CENTRE ASSIGNMENTS  ", f.icenter), ("TYPE ASSIGNMENTS    ", f.itype)):
            for k in range(0, f.npri, 20):
                fh.write(name + "".join("%3d" % v for v in arr[k:k + 20]) + "\n")
        for k in range(0, f.npri, 5):
            fh.write("EXPONENTS ".ljust(10) + "".join(_fortran_e(float(v), 14, 7) for v in f.e[k:k + 5]) + "\n")
        for i in range(f.nmoocc):
            fh.write(("MO%5d" % (i + 1)).ljust(16) + "MO 0.0".ljust(8) + "OCC NO = ".rjust(11) + "%12.7f" % f.occ[i]
                     + "  ORB. ENERGY =" + "%12.6f" % ene[i] + "\n")
            c = f.cmo[:, i]
            for k in range(0, f.npri, 5):
                fh.write("".join(_fortran_e(float(v), 16, 8) for v in c[k:k + 5]) + "\n")
        fh.write("END DATA\n")
