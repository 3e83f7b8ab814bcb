"""TEST INFRASTRUCTURE (oracle): the text the reference's formatted writes produce for cube bodies and .dat lines.

CPU restatement of
  writegrid_cube   src/crystalmod@proc.f90:5026-5036   '(6(1x,e22.14))' | '(1p,6(1x,e12.5))'   g(ix,iy,iz), iz fastest
  write_cube_body  src/nci@proc.f90:867-883             '(6(1x,1p,e13.5e3))'                    c(k,j,i), k fastest
  .dat lines       src/nci@proc.f90:593-595             '(1p,E15.7,E15.7)'
following the Fortran 2008 rules for Ew.d / Ew.dEe editing with a kP scale factor (10.7.2.3.3): significand 0.d1..dd
(k = 0) or d1.d2..d(d+1) (k = 1), exponent field E+zz, or +zzz without the E when |exp| > 99 and no Ee is given, value
right-justified in w columns.  Decimal digits come from Python's correctly rounded '%e' (round-half-even on the exact
binary value), which is what libgfortran's write_float does in the default rounding mode.  Pure Python: small cases only.
Parity unpinned against the reference binary (no Fortran compiler here); pinned by the hand-checked fields in
tests/test_oracle.py."""
import math

FORMATS = {0: dict(w=22, d=14, k=0, e=2, lead=1, per_line=6),
           1: dict(w=12, d=5, k=1, e=2, lead=1, per_line=6),
           2: dict(w=13, d=5, k=1, e=3, lead=1, per_line=6),
           3: dict(w=15, d=7, k=1, e=2, lead=0, per_line=2)}


def fortran_e(x: float, w: int, d: int, k: int = 0, e: int = 2) -> str:
    """One Ew.d (e = 2) or Ew.dEe field with scale factor k in {0, 1}."""
    if math.isnan(x):
        return "NaN".rjust(w)
    if math.isinf(x):
        return ("-Infinity" if x < 0 else "Infinity").rjust(w)
    nd = d if k == 0 else d + 1
    neg = math.copysign(1.0, x) < 0
    if x == 0.0:
        digits, e10 = "0" * nd, -1 if k == 0 else 0
    else:
        s = "%.*e" % (nd - 1, abs(x))
        mant, ex = s.split("e")
        digits, e10 = mant.replace(".", ""), int(ex)
    expo = e10 + 1 if k == 0 else e10
    body = ("0." + digits) if k == 0 else (digits[0] + "." + digits[1:])
    if e == 3:
        es = "E%+04d" % expo
    elif abs(expo) <= 99:
        es = "E%+03d" % expo
    else:
        es = "%+04d" % expo
    out = ("-" if neg else "") + body + es
    return out.rjust(w) if len(out) <= w else "*" * w


def field(x: float, fmt: int) -> str:
    f = FORMATS[fmt]
    return " " * f["lead"] + fortran_e(x, f["w"], f["d"], f["k"], f["e"])


def record(vals, fmt: int) -> str:
    """One formatted write of `vals`: per_line fields per line, a new line at the end of the record."""
    per = FORMATS[fmt]["per_line"]
    out = []
    for i, v in enumerate(vals):
        out.append(field(float(v), fmt))
        if (i + 1) % per == 0 or i == len(vals) - 1:
            out.append("\n")
    return "".join(out)


def cube_body(g, fmt: int, order: int) -> bytes:
    """g: order 0 -> array indexed [iz][iy][ix] (x fastest in memory = lf(ix,iy,iz)), records over (ix, iy);
    order 1 -> array indexed [i][j][k] (k fastest = c(k,j,i)), records over (i, j)."""
    out = []
    if order == 0:
        n3, n2, n1 = g.shape
        for ix in range(n1):
            for iy in range(n2):
                out.append(record(g[:, iy, ix], fmt))
    else:
        n1, n2, n3 = g.shape
        for i in range(n1):
            for j in range(n2):
                out.append(record(g[i, j, :], fmt))
    return "".join(out).encode()


def dat_text(pairs) -> bytes:
    return "".join(record(p, 3) for p in pairs).encode()
