"""GPU parity tests: the CUDA path called through the C ABI (host buffers) against the CPU oracle on the same inputs.

Tolerances (BASELINE.json north_star): rho <= 1e-10 relative, derivatives and RDG <= 1e-8, identical NCI
classification (points with |lambda2| below the eigen-solver's own resolution are counted and excluded).
"""
import numpy as np
import pytest

import cases
from cases import check_rho, check_vec

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def G():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no GPU")
    from critic2_b200 import gpu
    gpu.init_devices([0])
    return gpu


@pytest.fixture(scope="module")
def O():
    import oracle
    oracle.build()
    return oracle


def _ofield(O, f, st):
    env = O.Environ(st)
    return O.Field(env, wfn=O.Wfn(f))


def _compare_cube(got, ref):
    e = {"rho": check_rho(got["f"], ref["f"])}
    g = np.stack([got[k] for k in ("gx", "gy", "gz")], -1)
    gr = np.stack([ref[k] for k in ("gx", "gy", "gz")], -1)
    e["grad"] = check_vec(g, gr, "grad", rho=ref["f"])
    hk = ("hxx", "hxy", "hxz", "hyy", "hyz", "hzz")
    e["hess"] = check_vec(np.stack([got[k] for k in hk], -1), np.stack([ref[k] for k in hk], -1), "hessian", rho=ref["f"])
    e["gmod"] = check_vec(got["gmod"][..., None], ref["gmod"][..., None], "gmod", rho=ref["f"])
    hn = np.sqrt(sum(ref[k] ** 2 for k in hk))
    err = np.abs(got["lap"] - ref["lap"]) / np.maximum(np.maximum(hn, 1e-4 * np.abs(ref["f"])), 1e-20)
    assert err.max() <= cases.TOL_DER, f"lap {err.max():.3e}"
    return e


@pytest.mark.parametrize("name,n", [("h2o-aqz", 24), ("benzene-adz", 20), ("hooh-atz", 20), ("h2o-wfx", 20)])
def test_cube_real_molecules(G, O, name, n):
    f = cases.golden_wfn(name)
    st = cases.mol_structure(f)
    x0, xd, nn = cases.bbox_grid(st, n, margin=3.0)
    ref = _ofield(O, f, st).cube(x0, xd, nn)
    fld = G.WfnField(f, st)
    got = fld.cube(x0, xd, nn)
    e = _compare_cube(got, ref)
    print(name, fld.config(), e)


def test_cube_cell_grid_ragged(G, O):
    """CUBE CELL-like grid over the whole molecular cell (mostly vacuum: exact zeros) with ragged sizes."""
    from critic2_b200 import drivers
    f = cases.golden_wfn("h2o-aqz")
    st = cases.mol_structure(f, border=40.0)
    o, xd, nn = drivers.cube_grid(st, (13, 7, 10))
    ref = _ofield(O, f, st).cube(o, xd, nn)
    got = G.WfnField(f, st).cube(o, xd, nn)
    _compare_cube(got, ref)
    assert (ref["f"] == 0).any()


@pytest.mark.parametrize("cfg", [0, 1, 2, 3, 4, 5, 6, 7, 8, 12, 13, 14, 15, 16, 17])
@pytest.mark.parametrize("variant", ["rhf", "uhf_edf"])
def test_cube_synthetic_all_configs(G, O, cfg, variant):
    """24-atom def2-TZVP-shaped synthetic molecule through every block-synchronous and warp-specialised configuration
    (several MO blocks for the small ones: 42 / 85 MOs in blocks of 8, 24, 64, 112)."""
    from critic2_b200 import synth
    f = synth.small_synthetic(uhf=(variant != "rhf"), with_edf=(variant != "rhf"))
    st = cases.mol_structure(f)
    x0, xd, nn = cases.bbox_grid(st, (18, 14, 12), margin=3.0)
    ref = _ofield(O, f, st).cube(x0, xd, nn)
    G.set_option("wfn_cfg", cfg)
    try:
        fld = G.WfnField(f, st)
    finally:
        G.set_option("wfn_cfg", -1)
    assert fld.config()["cfg"] == cfg
    got = fld.cube(x0, xd, nn)
    print(cfg, variant, _compare_cube(got, ref))


@pytest.mark.parametrize("cfg", [3, 6])
def test_interleaved_parts_tile_the_grid(G, cfg):
    """c2gpu_eval_grid_dev_part / _part_rows: the parts are disjoint, cover the grid and reproduce the one-launch result
    bit for bit (the multi-GPU decomposition, here on one device)."""
    import torch
    from critic2_b200 import synth
    f = synth.small_synthetic()
    st = cases.mol_structure(f)
    x0, xd, nn = cases.bbox_grid(st, (21, 19, 26), margin=3.0)
    G.set_option("wfn_cfg", cfg)
    try:
        fld = G.WfnField(f, st)
    finally:
        G.set_option("wfn_cfg", -1)
    npt = int(np.prod(nn))
    props = ("f", "gx", "lap")
    full = [torch.zeros(npt, dtype=torch.float64, device="cuda") for _ in props]
    fld.cube_dev(x0, xd, nn, full, props)
    for rows in (False, True):
        acc = [torch.zeros(npt, dtype=torch.float64, device="cuda") for _ in props]
        cover = torch.zeros(npt, dtype=torch.float64, device="cuda")
        for part in range(3):
            outs = [torch.full((npt,), float("nan"), dtype=torch.float64, device="cuda") for _ in props]
            fld.cube_dev_part(x0, xd, nn, outs, props, part, 3, rows=rows)
            torch.cuda.synchronize()
            own = ~torch.isnan(outs[0])
            cover += own.double()
            for a_, o_ in zip(acc, outs):
                a_ += torch.nan_to_num(o_, nan=0.0)
        assert bool((cover == 1.0).all()), "parts must be disjoint and cover the grid"
        for a_, f_ in zip(acc, full):
            assert torch.equal(a_, f_)


@pytest.mark.parametrize("cfg", [3, 6, 7])
def test_point_lists_through_every_kernel_family(G, O, cfg):
    """c2gpu_eval_points (ragged list: 1 000 points is not a multiple of any tile) through the block-synchronous, the
    warp-specialised and the small-MO kernels."""
    from critic2_b200 import synth
    f = synth.small_synthetic()
    st = cases.mol_structure(f)
    rng = np.random.default_rng(5)
    lo, hi = st.r.min(0) - 2.5, st.r.max(0) + 2.5
    pts = rng.uniform(lo, hi, (1000, 3))
    pts[:3] = st.r[:3]                                   # nuclei: gradient forced to zero (fieldmod :836-838)
    ref = _ofield(O, f, st).grd_points(pts, nder=2, periodic=False)
    G.set_option("wfn_cfg", cfg)
    try:
        fld = G.WfnField(f, st)
    finally:
        G.set_option("wfn_cfg", -1)
    got = fld.points(pts, nder=2, periodic=False)
    check_rho(got[:, 0], ref[:, 0], "rho")
    check_vec(got[:, 1:4], ref[:, 1:4], "grad", rho=ref[:, 0])
    check_vec(got[:, 4:13], ref[:, 4:13], "hess", rho=ref[:, 0])
    assert np.array_equal(got[:, 15] != 0, ref[:, 15] != 0) and np.all(got[:3, 1:4] == 0.0)


@pytest.mark.parametrize("cfg", [-1, 3, 6])
@pytest.mark.parametrize("nder", [0, 1])
def test_lower_nder(G, O, nder, cfg):
    """nder = 0 / 1 (grd0, LINE/PLANE of f and gradient components): the ND = 1 and ND = 4 instantiations of the default,
    the block-synchronous and the warp-specialised kernels."""
    f = cases.golden_wfn("benzene-adz")
    st = cases.mol_structure(f)
    x0, xd, nn = cases.bbox_grid(st, 12, margin=2.0)
    pts = np.array([x0 + xd @ np.array([i, j, k], float) for k in range(nn[2]) for j in range(nn[1]) for i in range(nn[0])])
    ref = _ofield(O, f, st).grd_points(pts, nder=nder, periodic=False)
    G.set_option("wfn_cfg", cfg)
    try:
        fld = G.WfnField(f, st)
    finally:
        G.set_option("wfn_cfg", -1)
    got = fld.cube(x0, xd, nn, nder=nder)
    check_rho(got["f"].reshape(-1), ref[:, 0])
    if nder == 1:
        check_vec(np.stack([got[k].reshape(-1) for k in ("gx", "gy", "gz")], -1), ref[:, 1:4], "grad", rho=ref[:, 0])
    else:
        assert not got["gx"].any()


def test_points_and_nucleus(G, O):
    """field%grd on a point list, including a point on a nucleus (gradient zeroed, fieldmod@proc.f90:836-838)."""
    f = cases.golden_wfn("h2o-aqz")
    st = cases.mol_structure(f)
    rng = np.random.default_rng(3)
    pts = st.r[rng.integers(0, 3, 700)] + rng.normal(0, 1.5, (700, 3))
    pts[:3] = st.r
    pts[3] = st.r[0] + np.array([3e-6, 0, 0])
    pts[4] = st.r[0] + np.array([3e-5, 0, 0])
    ref = _ofield(O, f, st).grd_points(pts, nder=2, periodic=True)
    got = G.WfnField(f, st).points(pts, nder=2, periodic=True)
    assert np.array_equal(got[:, 15], ref[:, 15])
    assert ref[:4, 15].all() and not ref[4, 15]
    check_rho(got[:, 0], ref[:, 0])
    check_vec(got[:, 1:4], ref[:, 1:4], "grad", rho=ref[:, 0])
    check_vec(got[:, 4:13], ref[:, 4:13], "hess", rho=ref[:, 0])
    assert not got[:4, 1:4].any() and got[:4, 13].max() == 0.0


def _nci_compare(got_rho, got_grad, ref_rho, ref_grad):
    # |crho| = 100*rho ; sign = sign(lambda2)
    check_rho(np.abs(got_rho), np.abs(ref_rho), "crho magnitude")
    # RDG = |grad rho| / (c rho^(4/3)): where |grad rho| vanishes by symmetry (critical points on grid nodes) the
    # gradient floor of cases.check_vec (1e-4*rho) carries over as 1e-4*rho/(c rho^(4/3))
    rho = np.maximum(np.abs(ref_rho) / 100.0, 1e-80)
    floor = 1e-4 * rho / (2.0 * (3.0 * np.pi ** 2) ** (1.0 / 3.0) * rho ** (4.0 / 3.0))
    err = np.abs(got_grad - ref_grad) / np.maximum(np.abs(ref_grad), floor)
    big = ref_grad < 1e40
    assert err[big].max() <= cases.TOL_DER, f"RDG {err[big].max():.3e}"
    flips = np.sign(got_rho) != np.sign(ref_rho)
    return int(flips.sum())


def test_nci_benzene(G, O):
    from critic2_b200 import drivers
    f = cases.golden_wfn("benzene-adz")
    st = cases.mol_structure(f)
    x0, xmat, ns = drivers.nci_grid(st, nstep=(22, 20, 18))
    of = _ofield(O, f, st)
    rrho, rgrad, _ = of.nci(x0, xmat, ns)
    fld = G.WfnField(f, st)
    crho, cgrad, _ = fld.nci(x0, xmat, ns)
    nflip = _nci_compare(crho, cgrad, rrho, rgrad)
    assert nflip == 0, f"{nflip} sign(lambda2) flips"
    # cutoff pass: identical classification
    rg = rgrad.copy()
    rsums, rdat = O.nci_cutoffs(rrho.copy(), rg)
    cg = cgrad.reshape(-1).copy()
    sums, dat, ndat = G.nci_cutoffs(crho.reshape(-1), cg)
    assert ndat == len(rdat)
    assert np.array_equal(cg.reshape(rg.shape) == 100.0, rg == 100.0)
    np.testing.assert_allclose(dat[:, 0], rdat[:, 0], rtol=1e-10, atol=0)
    rho = np.maximum(np.abs(rdat[:, 0]), 1e-80)
    floor = 1e-4 * rho / (2.0 * (3.0 * np.pi ** 2) ** (1.0 / 3.0) * rho ** (4.0 / 3.0))
    assert (np.abs(dat[:, 1] - rdat[:, 1]) / np.maximum(rdat[:, 1], floor)).max() <= cases.TOL_DER


def test_nci_void_with_promolecular(G, O):
    """NCIPLOT with VOID: rhoat from the promolecular field f(0) (nci@proc.f90:529-531) and the void sums (:573-582)."""
    from critic2_b200 import drivers
    f = cases.golden_wfn("h2o-aqz")
    st = cases.mol_structure(f)
    tabs = cases.tables()
    x0, xmat, ns = drivers.nci_grid(st, nstep=(16, 15, 14))
    env = O.Environ(st)
    of = O.Field(env, wfn=O.Wfn(f))
    op = O.Field(env, grids=O.Grid1Set(tabs))
    rrho, rgrad, rat = of.nci(x0, xmat, ns, promol=op)
    fld, pro = G.WfnField(f, st), G.PromolField(st, tabs)
    crho, cgrad, cat = fld.nci(x0, xmat, ns, promol=pro)
    assert _nci_compare(crho, cgrad, rrho, rgrad) == 0
    check_rho(cat, rat, "rhoat")
    rg = rgrad.copy()
    rsums, rdat = O.nci_cutoffs(rrho.copy(), rg, rat, rho_void=0.05, omega=st.omega)
    cg = cgrad.reshape(-1).copy()
    sums, dat, ndat = G.nci_cutoffs(crho.reshape(-1), cg, cat.reshape(-1), rho_void=0.05, omega=st.omega)
    assert ndat == len(rdat)
    np.testing.assert_allclose(sums, rsums, rtol=1e-10)
    assert np.array_equal(cg.reshape(rg.shape) == 100.0, rg == 100.0)


def test_nci_fragments(G, O):
    """NCIPLOT with two fragments (FRAGMENT keyword): box from the fragments (:350-368), rhofrag (:534-550) and the
    `inter` classification (:579-582), host entry points and the interleaved device layers."""
    import torch
    from critic2_b200 import drivers
    f = cases.golden_wfn("benzene-adz")
    st = cases.mol_structure(f)
    tabs = cases.tables()
    order = np.argsort(st.r[:, 0], kind="stable")
    ia, ib = order[:6], order[6:]
    frag_xyz = [st.r[ia], st.r[ib]]
    frag_z = [f.z[ia], f.z[ib]]
    x0, xmat, ns = drivers.nci_grid(st, nstep=(18, 17, 16), frag_xyz=frag_xyz)
    lo = np.concatenate(frag_xyz).min(0) - 2.0
    np.testing.assert_allclose(x0, lo, rtol=0, atol=1e-12)
    env = O.Environ(st)
    of = O.Field(env, wfn=O.Wfn(f))
    grids = O.Grid1Set(tabs)
    op = O.Field(env, grids=grids)
    rrho, rgrad, rat = of.nci(x0, xmat, ns, promol=op)
    rfr = O.nci_fragments(list(zip(frag_xyz, frag_z)), grids, x0, xmat, ns)
    fld, pro = G.WfnField(f, st), G.PromolField(st, tabs)
    frs = [G.FragmentField(r, z, tabs) for r, z in zip(frag_xyz, frag_z)]
    opts = drivers.NciOptions(rhoparam=0.9, rhoparam2=0.8)
    out = drivers.nciplot(fld, st, nstep=(18, 17, 16), opts=opts, promol=pro, fragments=frs, frag_xyz=frag_xyz)
    assert _nci_compare(out["crho"], out["cgrad"], rrho, rgrad) == 0
    check_rho(out["rhoat"], rat, "rhoat")
    check_rho(out["rhofrag"], rfr, "rhofrag")
    rg = rgrad.copy()
    rsums, rdat = O.nci_cutoffs(rrho.copy(), rg, rat, rhofrag=rfr, rhoparam=0.9, rhoparam2=0.8, omega=st.omega)
    assert 0 < len(rdat) < rrho.size
    assert out["ndat"] == len(rdat)
    assert np.array_equal(out["cgrad_masked"] == 100.0, rg == 100.0)
    np.testing.assert_allclose(out["dat"][:, 0], rdat[:, 0], rtol=1e-10, atol=0)
    # device entry: two interleaved parts fill disjoint x-layers and together reproduce the host result
    npt = int(np.prod(ns))
    dev = torch.zeros(2 * npt, dtype=torch.float64, device="cuda")
    lh = fld.layer_height()
    for part in range(2):
        G.nci_fragments_dev(frs, x0, xmat, ns, dev, part=part, nparts=2, layer_height=lh)
    torch.cuda.synchronize()
    assert np.array_equal(dev.cpu().numpy().reshape(out["rhofrag"].shape), out["rhofrag"])


def _format_samples(rng, n):
    """Doubles that stress the decimal conversion: all magnitudes, subnormals, exact ties, powers of ten, zeros."""
    v = [np.ldexp(rng.uniform(0.5, 1.0, n // 2), rng.integers(-1074, 1024, n // 2)) * rng.choice([-1.0, 1.0], n // 2)]
    v.append(rng.normal(0, 1, n // 8) * 10.0 ** rng.integers(-12, 6, n // 8))
    ties = []
    for D in (6, 8, 14):       # N + 1/2 at every decimal position that is exactly representable
        N = rng.integers(10 ** (D - 1), 10 ** D, 200).astype(float)
        for p in range(-20, 21):
            x = (N + 0.5) * 10.0 ** p if p >= 0 else (N + 0.5) / 10.0 ** (-p)
            ties.append(x)
    v.append(np.concatenate(ties))
    v.append(10.0 ** rng.integers(-320, 308, 200))
    v.append(np.array([0.0, -0.0, 5e-324, -5e-324, 2.2250738585072014e-308, 1.7976931348623157e308, 0.5, 1.5, 2.5,
                       9.9999995, 99999.95, 0.999999999999995, 1e15 + 50.0, np.inf, -np.inf, np.nan]))
    out = np.concatenate(v)
    rng.shuffle(out)
    return out


@pytest.mark.parametrize("force_exact", [0, 1])
def test_cube_text_matches_fortran_formatting(G, O, force_exact):
    """writegrid_cube / write_cube_body / .dat text: byte-exact against the formatted-output oracle,
    fast path and the big-integer rounding path, both record orders, ragged sizes."""
    ct = O.cube_text
    rng = np.random.default_rng(123)
    G.set_option("fmt_force_exact", force_exact)
    try:
        for (shape, order) in (((5, 7, 13), 0), ((3, 5, 6), 0), ((6, 3, 11), 1), ((2, 2, 400), 1)):
            nv = int(np.prod(shape))
            g = _format_samples(rng, 4096)[:nv].reshape(shape)
            for fmt in (0, 1, 2):
                got = G.cube_text(g, fmt, order)
                ref = ct.cube_body(g, fmt, order)
                assert len(got) == len(ref)
                if got != ref:
                    i = next(k for k in range(len(ref)) if got[k] != ref[k])
                    raise AssertionError(f"fmt {fmt} order {order}: first difference at byte {i}: "
                                         f"{got[max(0, i - 30):i + 30]!r} vs {ref[max(0, i - 30):i + 30]!r}")
        vals = _format_samples(rng, 8192)
        pairs = vals[:2 * 1000].reshape(-1, 2)
        assert G.dat_text(pairs) == ct.dat_text(pairs)
        assert G.dat_text(pairs[:7]) == ct.dat_text(pairs[:7])
        # every sample once through each format, as one long record
        g1 = vals.reshape(1, 1, -1)
        for fmt in (0, 1, 2):
            assert G.cube_text(g1, fmt, 1) == ct.cube_body(g1, fmt, 1)
    finally:
        G.set_option("fmt_force_exact", 0)


def test_written_cube_file_round_trips(G, O, tmp_path):
    """CUBE -> writegrid_cube: header as crystalmod@proc.f90:5014-5031, body = the oracle's text, values recovered to the
    14 printed digits; the NCIPLOT body and .dat writers likewise."""
    from critic2_b200 import drivers
    f = cases.golden_wfn("h2o-wfx")
    st = cases.mol_structure(f)
    fld = G.WfnField(f, st)
    res, (o, xd, nn) = drivers.rhoplot_cube(fld, st, (9, 7, 11), props=("f", "lap"))
    path = tmp_path / "rho.cube"
    nbytes = drivers.writegrid_cube(str(path), st, res["f"], precise=True)
    raw = path.read_bytes()
    assert len(raw) == nbytes
    lines = raw.decode().split("\n")
    assert lines[0] == "critic2-cube" and int(lines[2].split()[0]) == st.ncell
    assert [int(lines[3 + i].split()[0]) for i in range(3)] == [9, 7, 11]
    head = drivers.cube_header(st, (9, 7, 11))
    assert raw[len(head):] == O.cube_text.cube_body(res["f"], 0, 0)
    vals = np.array(" ".join(lines[6 + st.ncell:]).split(), float).reshape(9, 7, 11)   # (ix, iy, iz)
    np.testing.assert_allclose(vals, res["f"].transpose(2, 1, 0), rtol=5.1e-14, atol=0)
    # NCIPLOT cube body + .dat
    x0, xmat, ns = drivers.nci_grid(st, nstep=(6, 5, 7))
    crho, cgrad, _ = fld.nci(x0, xmat, ns)
    with open(tmp_path / "dens.cube", "wb") as fh:
        drivers.write_nci_cube_body(fh, crho)
    assert (tmp_path / "dens.cube").read_bytes() == O.cube_text.cube_body(crho, 2, 1)
    cg = cgrad.reshape(-1).copy()
    sums, dat, ndat = G.nci_cutoffs(crho.reshape(-1), cg)
    drivers.write_nci_dat(str(tmp_path / "nci.dat"), dat)
    assert (tmp_path / "nci.dat").read_bytes() == O.cube_text.dat_text(dat)


def test_line_plane_and_mesh_drivers(G, O):
    """The point-list callers of grd: LINE, PLANE, mesh fill against the oracle's grd."""
    from critic2_b200 import drivers
    f = cases.golden_wfn("hooh-atz")
    st = cases.mol_structure(f)
    of = _ofield(O, f, st)
    fld = G.WfnField(f, st)
    a, b = st.r[0] + np.array([0.3, -0.2, 0.1]), st.r[-1] + np.array([1.5, 1.0, -0.7])
    xp, rho, lap = drivers.rhoplot_line(fld, st, a, b, 41, "lap")
    ref = of.grd_points(xp, nder=2, periodic=False)
    check_rho(rho, ref[:, 0], "line rho")
    hn = np.sqrt((ref[:, 4:13] ** 2).sum(1))
    assert (np.abs(lap - ref[:, 14]) / np.maximum(hn, 1e-4 * np.abs(ref[:, 0]))).max() <= cases.TOL_DER
    c = st.r.mean(0)
    ff = drivers.rhoplot_plane(fld, st, c + [-3, -3, 0.2], c + [3, -3, 0.2], c + [-3, 3, 0.2], 13, 11, "gmod")
    ix, iy = np.meshgrid(np.arange(13.0), np.arange(11.0), indexing="ij")
    pts = (c + [-3, -3, 0.2])[None, None, :] + ix[..., None] * np.array([0.5, 0, 0]) + iy[..., None] * np.array([0, 0.6, 0])
    refp = of.grd_points(pts.reshape(-1, 3), nder=1, periodic=False)
    check_vec(ff.reshape(-1, 1), refp[:, 13:14], "plane gmod", rho=refp[:, 0])
    m = drivers.mesh_fill(fld, pts.reshape(-1, 3)[:50], ("volume", "rho", "gradrho"), periodic=False)
    assert np.all(m[:, 0] == 1.0)
    check_rho(m[:, 1], refp[:50, 0], "mesh rho")
    check_vec(m[:, 2:3], refp[:50, 13:14], "mesh gradrho", rho=refp[:50, 0])


def test_promolecular_quartz_cube(G, O):
    """Promolecular density of alpha-quartz (periodic, images) with gradient and Hessian on a ragged whole-cell grid."""
    from critic2_b200 import drivers
    st = cases.quartz(1)
    tabs = cases.tables()
    env = O.Environ(st)
    of = O.Field(env, grids=O.Grid1Set(tabs))
    o, xd, nn = drivers.cube_grid(st, (20, 18, 21))
    ref = of.cube(o, xd, nn, periodic=True)
    pro = G.PromolField(st, tabs)
    got = pro.cube(o, xd, nn, periodic=True)
    print(_compare_cube(got, ref))


def test_promolecular_grid_and_nci(G, O):
    st = cases.quartz(1)
    tabs = cases.tables()
    env = O.Environ(st)
    grids = O.Grid1Set(tabs)
    n = (17, 16, 19)
    ref = O.promolecular_grid(env, grids, n)
    pro = G.PromolField(st, tabs)
    got = pro.promolecular_grid(n)
    check_rho(got, ref, "promolecular_grid")
    # NCIPLOT on the promolecular density (iref = 0), periodic cell
    from critic2_b200 import drivers
    x0, xmat, ns = drivers.nci_grid(st, nstep=(12, 13, 14))
    of = O.Field(env, grids=grids)
    rrho, rgrad, _ = of.nci(x0, xmat, ns)
    crho, cgrad, _ = pro.nci(x0, xmat, ns)
    assert _nci_compare(crho, cgrad, rrho, rgrad) == 0


def test_promolecular_molecule_fragment(G, O):
    f = cases.golden_wfn("benzene-adz")
    st = cases.mol_structure(f)
    tabs = cases.tables()
    mask = np.zeros(st.ncell, np.uint8)
    mask[:6] = 1
    env = O.Environ(st)
    x0, xd, nn = cases.bbox_grid(st, (15, 14, 9), margin=5.0)
    for fm in (None, mask):
        of = O.Field(env, grids=O.Grid1Set(tabs), fragmask=fm)
        ref = of.cube(x0, xd, nn)
        got = G.PromolField(st, tabs, fragmask=fm).cube(x0, xd, nn)
        _compare_cube(got, ref)


def _core_fields(G, O, f, st, zpsp, promol_main=False):
    """(GPU field with its core attached, oracle field with the same core tables)"""
    ct = cases.core_tables()
    core = G.core_field(st, ct, zpsp)
    otabs = {z: ct[(z, q)] for z, q in zpsp.items() if q > 0}
    env = O.Environ(st)
    if promol_main:
        fld = G.PromolField(st, cases.tables())
        of = O.Field(env, grids=O.Grid1Set(cases.tables()), core=O.Grid1Set(otabs))
    else:
        fld = G.WfnField(f, st)
        of = O.Field(env, wfn=O.Wfn(f), core=O.Grid1Set(otabs))
    fld.set_core(core)
    return fld, of


@pytest.mark.parametrize("cfg", [-1, 3, 6])
def test_core_augmented_cube_points_and_nci(G, O, cfg):
    """LOAD ... CORE ZPSP: grd adds the promolecular core density (fieldmod@proc.f90:827-832) before the nucleus test,
    gfmod / del2f and the NCIPLOT formulas; through the small-MO, block-synchronous and warp-specialised kernels."""
    from critic2_b200 import drivers
    f = cases.golden_wfn("benzene-adz")
    st = cases.mol_structure(f)
    G.set_option("wfn_cfg", cfg)
    try:
        fld, of = _core_fields(G, O, f, st, {6: 4, 1: 0})
    finally:
        G.set_option("wfn_cfg", -1)
    x0, xd, nn = cases.bbox_grid(st, (23, 21, 18), margin=2.0)
    ref = of.cube(x0, xd, nn)
    plain = _ofield(O, f, st).cube(x0, xd, nn)
    assert np.abs(ref["f"] - plain["f"]).max() > 1e-2          # the core really contributes on this lattice
    _compare_cube(fld.cube(x0, xd, nn), ref)
    for nder in (0, 1):                                        # grd0-like and gradient-only passes carry the core too
        got = fld.cube(x0, xd, nn, props=("f", "gx", "gy", "gz", "gmod"), nder=nder)
        check_rho(got["f"], ref["f"])
        if nder:
            check_vec(np.stack([got[k] for k in ("gx", "gy", "gz")], -1),
                      np.stack([ref[k] for k in ("gx", "gy", "gz")], -1), "grad", rho=ref["f"])
    # point list incl. the nuclei (gradient nullified after the core is added)
    rng = np.random.default_rng(5)
    pts = np.concatenate([st.r[:3], st.r[0] + rng.uniform(-1.0, 1.0, (300, 3))])
    rres = of.grd_points(pts, 2, True)
    res = fld.points(pts, 2, True)
    assert np.array_equal(res[:, 15], rres[:, 15]) and res[:3, 15].all() and not res[:3, 1:4].any()
    check_rho(res[:, 0], rres[:, 0])
    check_vec(res[:, 1:4], rres[:, 1:4], "grad", rho=rres[:, 0])
    check_vec(res[:, 4:13], rres[:, 4:13], "hess", rho=rres[:, 0])
    # NCIPLOT on the augmented field
    x0n, xmat, ns = drivers.nci_grid(st, nstep=(18, 17, 16))
    rrho, rgrad, _ = of.nci(x0n, xmat, ns)
    crho, cgrad, _ = fld.nci(x0n, xmat, ns)
    assert _nci_compare(crho, cgrad, rrho, rgrad) == 0
    # detaching restores the valence-only field
    fld.set_core(None)
    check_rho(fld.cube(x0, xd, nn, props=("f",))["f"], plain["f"])


def test_core_prepass_follows_the_interleaved_parts(G):
    """The core pre-pass only evaluates the tiles of the units a part owns; the parts still tile the grid bit for bit."""
    import torch
    f = cases.golden_wfn("benzene-adz")
    st = cases.mol_structure(f)
    G.set_option("wfn_cfg", 3)
    try:
        fld = G.WfnField(f, st)
    finally:
        G.set_option("wfn_cfg", -1)
    fld.set_core(G.core_field(st, cases.core_tables(), {6: 4}))
    x0, xd, nn = cases.bbox_grid(st, (26, 21, 19), margin=2.0)
    npt = int(np.prod(nn))
    props = ("f", "gmod", "lap")
    full = [torch.zeros(npt, dtype=torch.float64, device="cuda") for _ in props]
    fld.cube_dev(x0, xd, nn, full, props)
    for rows in (False, True):
        acc = [torch.zeros(npt, dtype=torch.float64, device="cuda") for _ in props]
        for part in range(3):
            outs = [torch.full((npt,), float("nan"), dtype=torch.float64, device="cuda") for _ in props]
            fld.cube_dev_part(x0, xd, nn, outs, props, part, 3, rows=rows)
            torch.cuda.synchronize()
            for a_, o_ in zip(acc, outs):
                a_ += torch.nan_to_num(o_, nan=0.0)
        for a_, f_ in zip(acc, full):
            assert torch.equal(a_, f_)


def test_core_on_a_promolecular_crystal_field(G, O):
    """usecore on a type_promol field in a periodic cell (quartz, Si zpsp 4, O zpsp 6): the core call uses wc, the
    valence call wcr (fieldmod@proc.f90:803,828); both are folded by y2z_center."""
    st = cases.quartz(1)
    fld, of = _core_fields(G, O, None, st, {14: 4, 8: 6}, promol_main=True)
    from critic2_b200 import drivers
    o, xd, nn = drivers.cube_grid(st, (18, 17, 20))
    ref = of.cube(o, xd, nn, periodic=True)
    _compare_cube(fld.cube(o, xd, nn, periodic=True), ref)


def test_promolecular_many_species_and_table_edges(G, O):
    """More species than the kernel stages in shared memory (constants then come from global memory), elements without a
    table (agrid not initialised: skipped, environmod@proc.f90:1298-1299), and points on the table edges: exactly on a
    nucleus (r clamped to r(1), :1302), at the last node and just past it (interp returns 0 from rmax on, grid1mod :98)."""
    from critic2_b200 import structure
    rng = np.random.default_rng(21)
    zs = np.array([1, 6, 7, 8, 14, 2, 3, 4, 5, 9, 10, 11, 12, 13, 15, 16, 17, 18, 19, 20], np.int32)   # 20 species
    xyz = rng.uniform(0.0, 9.0, (len(zs), 3))
    st = structure.molecular_cell(xyz, zs)
    tabs = cases.tables()
    fld = G.PromolField(st, tabs)
    of = O.Field(O.Environ(st), grids=O.Grid1Set(tabs))
    x0, xd, nn = cases.bbox_grid(st, (19, 17, 15), margin=3.0)
    _compare_cube(fld.cube(x0, xd, nn), of.cube(x0, xd, nn))
    # table edges of the oxygen atom (index 3), approached along x where no other tabulated atom is within reach
    far = structure.molecular_cell(np.array([[0.0, 0.0, 0.0], [40.0, 0.0, 0.0]]), np.array([8, 1], np.int32))
    fldf = G.PromolField(far, tabs)
    off = O.Field(O.Environ(far), grids=O.Grid1Set(tabs))
    rmax, r1 = float(tabs[8]["r"][-1]), float(tabs[8]["r"][0])
    d = np.array([0.0, 0.5 * r1, r1, float(tabs[8]["r"][1]), float(tabs[8]["r"][-2]), rmax * (1 - 1e-15), rmax,
                  rmax * (1 + 1e-15), rmax + 1e-9])
    pts = far.r[0] + np.outer(d, [-1.0, 0.0, 0.0])
    res, rres = fldf.points(pts, 2, True), off.grd_points(pts, 2, True)
    assert np.array_equal(res[:, 0] == 0.0, rres[:, 0] == 0.0) and rres[-3:, 0].max() == 0.0 and rres[0, 15] == 1.0
    check_rho(res[:, 0], rres[:, 0])
    check_vec(res[:, 1:4], rres[:, 1:4], "grad", rho=rres[:, 0])
    check_vec(res[:, 4:13], rres[:, 4:13], "hess", rho=rres[:, 0])


@pytest.mark.parametrize("name,cfg", [("benzene-adz", -1), ("hooh-atz", -1), ("synth", 3), ("synth", 6), ("synth", 13), ("synth", 14),
                                      ("synth", 16), ("synth", 17)])
def test_gkin_and_mo_values_on_point_lists(G, O, name, cfg):
    """mesh%fill's remaining properties (meshmod@proc.f90:456-464): res%gkin (three derivative-selected value passes of the
    wavefunction kernels) and MO values (grd nder=-1 -> calculate_mo, a linear pass over a one-MO handle)."""
    from critic2_b200 import drivers, synth
    f = synth.small_synthetic(uhf=True, with_edf=True) if name == "synth" else cases.golden_wfn(name)
    st = cases.mol_structure(f)
    G.set_option("wfn_cfg", cfg)
    try:
        fld = G.WfnField(f, st)
    finally:
        G.set_option("wfn_cfg", -1)
    of = _ofield(O, f, st)
    rng = np.random.default_rng(13)
    pts = np.concatenate([st.r[rng.integers(0, f.natom, 700)] + rng.normal(0, 1.5, (700, 3)),
                          st.r[:2], st.r[0] + np.array([[80.0, 0.0, 0.0]])])        # 703 points: ragged, nuclei, a far point (wrapped)
    rres = of.grd_points(pts, 1, True)
    gk = fld.gkin_points(pts)
    assert np.array_equal(gk == 0.0, rres[:, 17] == 0.0)
    err = np.abs(gk - rres[:, 17]) / np.maximum(rres[:, 17], 1e-30)
    assert err.max() <= cases.TOL_DER, f"gkin {err.max():.3e}"
    for imo in (1, f.nmoocc, min(f.cmo.shape[1], f.nmoocc + 1)):
        ref = of.mo_points(pts, imo, True)
        got = fld.mo_points(pts, imo)
        # an MO can vanish by symmetry (a pi orbital at in-plane points): there both sides hold rounding noise of their own
        # summation order, so exact zeros are only required where the reference has them (outside every cutoff)
        assert not got[ref == 0.0].any()
        assert np.max(np.abs(got - ref)) <= cases.TOL_RHO * np.abs(ref).max(), f"MO {imo}"
    out = drivers.mesh_fill(fld, pts, props=("volume", "rho", "gradrho", "gkin", 101))
    assert np.array_equal(out[:, 0], np.ones(len(pts))) and np.array_equal(out[:, 3], gk)
    check_rho(out[:, 1], rres[:, 0])
    assert np.array_equal(out[:, 4], fld.mo_points(pts, 1))


def test_nci_interleaved_parts_with_the_two_pass_epilogue(G):
    """c2gpu_nci_grid_dev_part through a block-synchronous kernel: the second-pass epilogue kernel must finish exactly the
    points of the x-layers the part owns (others stay untouched), and the parts reproduce the one-launch result bit for bit;
    the one-pass evaluation (development switch) gives the same numbers."""
    import torch
    from critic2_b200 import drivers
    f = cases.golden_wfn("benzene-adz")
    st = cases.mol_structure(f)
    G.set_option("wfn_cfg", 7)
    try:
        fld = G.WfnField(f, st)
    finally:
        G.set_option("wfn_cfg", -1)
    assert fld.config()["cfg"] == 7
    x0, xmat, ns = drivers.nci_grid(st, nstep=(27, 19, 22))
    npt = int(np.prod(ns))
    mk = lambda v: torch.full((npt,), v, dtype=torch.float64, device="cuda")
    full_rho, full_grad = mk(0.0), mk(0.0)
    fld.nci_dev(x0, xmat, ns, full_rho, full_grad)
    acc_rho, acc_grad, cover = mk(0.0), mk(0.0), mk(0.0)
    for part in range(3):
        crho, cgrad = mk(float("nan")), mk(float("nan"))
        fld.nci_dev_part(x0, xmat, ns, crho, cgrad, part, 3)
        torch.cuda.synchronize()
        own = ~torch.isnan(crho)
        assert torch.equal(own, ~torch.isnan(cgrad))
        cover += own.double()
        acc_rho += torch.nan_to_num(crho, nan=0.0)
        acc_grad += torch.nan_to_num(cgrad, nan=0.0)
    assert bool((cover == 1.0).all())
    assert torch.equal(acc_rho, full_rho) and torch.equal(acc_grad, full_grad)
    G.set_option("nci_two_pass", 0)
    try:
        one_rho, one_grad = mk(0.0), mk(0.0)
        fld.nci_dev(x0, xmat, ns, one_rho, one_grad)
        torch.cuda.synchronize()
    finally:
        G.set_option("nci_two_pass", 1)
    assert torch.equal(one_rho, full_rho) and torch.equal(one_grad, full_grad)


# ---- round 2: the multi-MO-block paths of the tensor-core kernels -------------------------------------------------------
@pytest.mark.parametrize("cfg", [3, 4, 6, 8, 12, 13, 14, 16, 17])
@pytest.mark.parametrize("nmo", [113, 300])
@pytest.mark.parametrize("variant", ["rhf", "uhf_edf"])
def test_cube_synthetic_multi_block(G, O, cfg, nmo, variant):
    """More MOs than one block holds (the loop wfn_private@proc.F90:2771-2776 cut into column blocks): 113 MOs = 2 blocks of
    112 with a one-column last block, 300 MOs = 3 blocks of 112 (5 of 64 for M2).  In the warp-specialised kernel (cfg 6)
    this runs the chi-tile cache re-read, both zigzag directions of the chunk walk and the register accumulation of the
    quadratic forms across blocks; in L / L1 / M2 the per-block recomputation and the fixed-order cross-block sum."""
    from critic2_b200 import synth
    f = synth.small_synthetic(uhf=(variant != "rhf"), with_edf=(variant != "rhf"), nmo=nmo)
    st = cases.mol_structure(f)
    x0, xd, nn = cases.bbox_grid(st, (18, 14, 12), margin=3.0)
    ref = _ofield(O, f, st).cube(x0, xd, nn)
    G.set_option("wfn_cfg", cfg)
    try:
        fld = G.WfnField(f, st)
    finally:
        G.set_option("wfn_cfg", -1)
    c = fld.config()
    assert c["cfg"] == cfg and c["nblk"] == -(-nmo // c["mo_block"]) and c["nblk"] >= 2
    got = fld.cube(x0, xd, nn)
    print(cfg, nmo, variant, c["nblk"], _compare_cube(got, ref))


@pytest.mark.parametrize("cfg", [4, 6, 13, 14, 16, 17])
def test_nci_and_points_multi_block(G, O, cfg):
    """NCIPLOT evaluation and a ragged point list with three MO blocks (classification identical, nuclei flagged)."""
    from critic2_b200 import drivers, synth
    f = synth.small_synthetic(nmo=230)
    st = cases.mol_structure(f)
    G.set_option("wfn_cfg", cfg)
    try:
        fld = G.WfnField(f, st)
    finally:
        G.set_option("wfn_cfg", -1)
    c = fld.config()
    assert c["cfg"] == cfg and c["nblk"] == -(-230 // c["mo_block"]) and c["nblk"] >= 3
    of = _ofield(O, f, st)
    x0, xmat, ns = drivers.nci_grid(st, nstep=(15, 13, 17))
    rrho, rgrad, _ = of.nci(x0, xmat, ns)
    crho, cgrad, _ = fld.nci(x0, xmat, ns)
    assert _nci_compare(crho, cgrad, rrho, rgrad) == 0
    rg, cg = rgrad.copy(), cgrad.reshape(-1).copy()
    _, rdat = O.nci_cutoffs(rrho.copy(), rg, rhocut=1e3, dimcut=50.0)
    _, dat, ndat = G.nci_cutoffs(crho.reshape(-1), cg, rhocut=1e3, dimcut=50.0)
    assert ndat == len(rdat) and np.array_equal(cg.reshape(rg.shape) == 100.0, rg == 100.0)
    rng = np.random.default_rng(7)
    pts = np.concatenate([st.r[:2], rng.uniform(st.r.min(0) - 2.0, st.r.max(0) + 2.0, (517, 3))])
    ref = of.grd_points(pts, nder=2, periodic=False)
    got = fld.points(pts, nder=2, periodic=False)
    check_rho(got[:, 0], ref[:, 0])
    check_vec(got[:, 1:4], ref[:, 1:4], "grad", rho=ref[:, 0])
    check_vec(got[:, 4:13], ref[:, 4:13], "hess", rho=ref[:, 0])
    assert np.array_equal(got[:, 15] != 0, ref[:, 15] != 0)


def test_count_active_matches_oracle(G, O):
    """The numerators of bench.py's rooflines: c2gpu_wfn_count_active = sum over points of the primitives passing the
    reference's own screen (wfn_private@proc.F90:2747; oracle orc_nact), c2gpu_promol_count_active = (point, atom) pairs
    entering the radial sum (environmod@proc.f90:1284-1300)."""
    from critic2_b200 import drivers, synth
    for f in (synth.small_synthetic(), cases.golden_wfn("h2o-aqz")):
        st = cases.mol_structure(f)
        x0, xd, nn = cases.bbox_grid(st, (19, 13, 11), margin=5.0)
        fld = G.WfnField(f, st)
        I, J, K = np.meshgrid(*[np.arange(int(n)) for n in nn], indexing="ij")
        pts = x0[None, :] + np.stack([I.ravel(), J.ravel(), K.ravel()], 1).astype(float) @ xd.T
        ref = int(_ofield(O, f, st).nact_points(pts).sum())
        assert int(fld.count_active(x0, xd, nn)) == ref and ref > 0
        zsel = pts[:, 2] >= x0[2] + 3 * xd[2, 2] - 1e-9
        zsel &= pts[:, 2] < x0[2] + 8 * xd[2, 2] - 1e-9
        assert int(fld.count_active(x0, xd, nn, iz0=3, iz1=8)) == int(_ofield(O, f, st).nact_points(pts[zsel]).sum())
    st = cases.quartz(1)
    tabs = cases.tables()
    o, xq, nq = drivers.cube_grid(st, (13, 11, 9))
    I, J, K = np.meshgrid(*[np.arange(int(n)) for n in nq], indexing="ij")
    pts = np.asarray(o)[None, :] + np.stack([I.ravel(), J.ravel(), K.ravel()], 1).astype(float) @ np.asarray(xq).T
    ref = int(O.promol_nat_points(O.Environ(st), O.Grid1Set(tabs), pts).sum())
    assert int(G.PromolField(st, tabs).count_active(o, xq, nq, periodic=True)) == ref and ref > 0


# ---- round 2: promolecular_grid variants (crystalmod@proc.f90:5118-5158) ---------------------------
def test_promolecular_grid_core_and_fragment_variants(G, O):
    """promolecular_grid(f, n, zpsp, fr): LOAD AS CORE (fieldmod@proc.f90:414-431) sums the core tables cgrid(iz, zpsp(iz)),
    the fragment variant (hirshfeld@proc.f90:49) only the atoms of a fragment; quartz (periodic images) and a molecule."""
    ct = cases.core_tables()
    tabs = cases.tables()
    # crystal, core
    st = cases.quartz(1)
    zpsp = {14: 4, 8: 6}
    n = (15, 14, 17)
    env = O.Environ(st)
    otabs = {z: ct[(z, q)] for z, q in zpsp.items()}
    ref = O.promolecular_grid(env, O.Grid1Set(otabs), n)
    got = G.core_field(st, ct, zpsp).promolecular_grid(n)
    check_rho(got, ref, "promolecular_grid(zpsp)")
    assert ref.max() > 1.0
    # crystal, only one species has a core (zpsp <= 0 for the other: no table, contributes nothing)
    ref1 = O.promolecular_grid(env, O.Grid1Set({14: ct[(14, 4)]}), n)
    got1 = G.core_field(st, ct, {14: 4, 8: 0}).promolecular_grid(n)
    check_rho(got1, ref1, "promolecular_grid(zpsp, Si only)")
    # crystal, fragment = the atoms of the first formula unit
    mask = np.zeros(st.ncell, np.uint8)
    mask[[0, 3, 4]] = 1
    reff = O.promolecular_grid(env, O.Grid1Set(tabs), n, fragmask=mask)
    gotf = G.PromolField(st, tabs, fragmask=mask).promolecular_grid(n)
    check_rho(gotf, reff, "promolecular_grid(fr)")
    full = O.promolecular_grid(env, O.Grid1Set(tabs), n)
    assert (reff < full).any() and reff.max() > 0
    # molecule (whole molecular cell: mostly exact zeros), core + fragment
    f = cases.golden_wfn("benzene-adz")
    sm = cases.mol_structure(f)
    envm = O.Environ(sm)
    nm = (21, 19, 11)
    refm = O.promolecular_grid(envm, O.Grid1Set({6: ct[(6, 4)]}), nm)
    gotm = G.core_field(sm, ct, {6: 4, 1: 0}).promolecular_grid(nm)
    check_rho(gotm, refm, "molecular promolecular_grid(zpsp)")
    maskm = np.zeros(sm.ncell, np.uint8)
    maskm[:5] = 1
    check_rho(G.PromolField(sm, tabs, fragmask=maskm).promolecular_grid(nm),
              O.promolecular_grid(envm, O.Grid1Set(tabs), nm, fragmask=maskm), "molecular promolecular_grid(fr)")


# ---- round 2: the single-process, all-devices host path (what the Fortran caller gets from gpu_init -> c2gpu_init(0)) -----
def test_multi_device_host_path_matches_one_device(G):
    """host_eval with ndev > 1 (csrc/c2gpu_drivers.inc): layers dealt round-robin over the replicas, per-slab events,
    interleaved copy-back, core pre-pass and two-pass NCIPLOT under c_nparts > 1 -- bit for bit the one-device result, on
    ragged grids whose layer count is not a multiple of the device count."""
    import torch
    from critic2_b200 import drivers, synth
    ndev = torch.cuda.device_count()
    if ndev < 2:
        pytest.skip("needs at least two GPUs")
    tabs, ct = cases.tables(), cases.core_tables()
    fb = cases.golden_wfn("benzene-adz")
    sb = cases.mol_structure(fb)
    fs = synth.small_synthetic(nmo=230)
    ss = cases.mol_structure(fs)
    q = cases.quartz(1)

    def run_all():
        out = {}
        for cfg, f, st, tag in ((-1, fb, sb, "benzene"), (6, fs, ss, "ws"), (4, fs, ss, "l1")):
            G.set_option("wfn_cfg", cfg)
            try:
                fld = G.WfnField(f, st)
            finally:
                G.set_option("wfn_cfg", -1)
            x0, xd, nn = cases.bbox_grid(st, (21, 19, 8 * (ndev + 1) + 3), margin=3.0)   # ndev + 2 layers, ragged last
            out[tag + "/cube"] = fld.cube(x0, xd, nn)
            xn, xm, ns = drivers.nci_grid(st, nstep=(8 * ndev + 5, 14, 13))
            pro = G.PromolField(st, tabs)
            out[tag + "/nci"] = dict(zip(("crho", "cgrad", "rhoat"), fld.nci(xn, xm, ns, promol=pro)))
            if tag == "benzene":
                fld.set_core(G.core_field(st, ct, {6: 4, 1: 0}))
                out[tag + "/core cube"] = fld.cube(x0, xd, nn, props=("f", "gmod", "lap"))
                out[tag + "/core nci"] = dict(zip(("crho", "cgrad"), fld.nci(xn, xm, ns)[:2]))
        pq = G.PromolField(q, tabs)
        out["quartz/grid"] = {"f": pq.promolecular_grid((13, 11, 4 * (ndev + 1) + 2))}
        o, xq, nq = drivers.cube_grid(q, (12, 10, 4 * ndev + 3))
        out["quartz/cube"] = pq.cube(o, xq, nq, periodic=True)
        return out

    try:
        G.init_devices([0])
        one = run_all()
        assert G.init(0) == ndev
        many = run_all()
    finally:
        G.init_devices([0])
    for key, d1 in one.items():
        for name, a in d1.items():
            assert np.array_equal(a, many[key][name]), f"{key}/{name} differs between 1 and {ndev} devices"


def test_libxc_reference_energies_through_the_point_list_path(G, O):
    """The reference-held numbers of /root/reference/tests_libxc/ref.txt:1-2 (LDA and PBE energies of h2o.wfx on critic2's
    default molecular mesh; see tests/test_oracle.py::test_libxc_reference_energies_pin_rho2) from the GPU's own rho and
    |grad rho|: c2gpu_eval_points on the 52 246 mesh nodes (mesh%fill's point list, meshmod@proc.f90:446-470)."""
    from oracle import mesh
    from test_oracle import LIBXC_REF, LIBXC_TOL, _libxc_mesh_integrals
    f = cases.golden_wfn("h2o-wfx")
    st = cases.mol_structure(f)
    x, w = mesh.genmesh_franchini(st.r, f.z)
    res = G.WfnField(f, st).points(x, nder=1, periodic=True)
    got = _libxc_mesh_integrals(res[:, 0], res[:, 13], w)
    assert abs(got["nelec"] - 10.0) < 1e-5
    assert abs(got["lda"] - LIBXC_REF["lda"]) < LIBXC_TOL and abs(got["pbe"] - LIBXC_REF["pbe"]) < LIBXC_TOL, got
    ref = _ofield(O, f, st).grd_points(x, nder=1, periodic=True)
    exp = _libxc_mesh_integrals(ref[:, 0], ref[:, 13], w)
    for k in exp:
        assert abs(got[k] - exp[k]) <= 1e-11 * abs(exp[k]), (k, got[k], exp[k])
    check_rho(res[:, 0], ref[:, 0])
    check_vec(res[:, 13:14], ref[:, 13:14], "gmod", rho=ref[:, 0])
    print(got)


@pytest.mark.parametrize("pinned", [False, True])
def test_gather_part_host_reassembles_the_grid(G, pinned):
    """c2gpu_gather_part_host: each part copies the units it evaluated (layers / rows of super-tiles, CUBE and NCIPLOT
    orders) from its own full-size device arrays into ONE host buffer; together they reproduce the one-launch result bit for
    bit, through the pinned bounce buffers (pageable destination) and by direct DMA (page-locked destination)."""
    import torch
    from critic2_b200 import drivers, synth
    f = synth.small_synthetic()
    st = cases.mol_structure(f)
    fld = G.WfnField(f, st)
    x0, xd, nn = cases.bbox_grid(st, (21, 19, 26), margin=3.0)
    npt = int(np.prod(nn))
    props = ("f", "gy", "hxz")
    full = [torch.zeros(npt, dtype=torch.float64, device="cuda") for _ in props]
    fld.cube_dev(x0, xd, nn, full, props)
    mk_host = (lambda n: torch.full((n,), -7.0, dtype=torch.float64).pin_memory()) if pinned else \
              (lambda n: torch.from_numpy(np.full(n, -7.0)))
    for rows in (False, True):
        host = [mk_host(npt) for _ in props]
        for part in range(3):
            outs = [torch.full((npt,), float("nan"), dtype=torch.float64, device="cuda") for _ in props]
            fld.cube_dev_part(x0, xd, nn, outs, props, part, 3, rows=rows)
            fld.gather_part_host(nn, outs, [t.data_ptr() for t in host], part, 3, rows=rows)
        for h_, f_ in zip(host, full):
            assert torch.equal(h_, f_.cpu())
        if pinned:   # the pass cut into z-slabs, every slab followed by an asynchronous ranged gather
            host = [mk_host(npt) for _ in props]
            for part in range(3):
                outs = [torch.full((npt,), float("nan"), dtype=torch.float64, device="cuda") for _ in props]
                for z0, z1 in ((0, 8), (8, 24), (24, 26)):
                    fld.cube_dev_part_range(x0, xd, nn, outs, props, part, 3, z0, z1, rows=rows)
                    fld.gather_part_host_range(nn, outs, [t.data_ptr() for t in host], part, 3, z0, z1, rows=rows, async_=True)
                fld.gather_wait()
            for h_, f_ in zip(host, full):
                assert torch.equal(h_, f_.cpu())
        else:
            with pytest.raises(G.C2GpuError):   # an asynchronous gather into pageable memory is refused
                fld.gather_part_host_range(nn, full, [t.data_ptr() for t in host], 0, 3, 0, 8, rows=rows, async_=True)
    xn, xm, ns = drivers.nci_grid(st, nstep=(27, 13, 11))
    nptn = int(np.prod(ns))
    frho, fgrad = (torch.zeros(nptn, dtype=torch.float64, device="cuda") for _ in range(2))
    fld.nci_dev(xn, xm, ns, frho, fgrad)
    host = [mk_host(nptn) for _ in range(2)]
    for part in range(2):
        outs = [torch.full((nptn,), float("nan"), dtype=torch.float64, device="cuda") for _ in range(2)]
        fld.nci_dev_part(xn, xm, ns, outs[0], outs[1], part, 2)
        fld.gather_part_host(ns, outs, [t.data_ptr() for t in host], part, 2, nci_order=True)
    assert torch.equal(host[0], frho.cpu()) and torch.equal(host[1], fgrad.cpu())


# ---- round 2: the warp-autonomous kernel for small MO counts (csrc/wfn_warp_kernel.cuh; configurations 9, 10, 11) ----------
@pytest.mark.parametrize("name,cfg", [("h2o-aqz", 9), ("h2o-aqz", 11), ("h2o-wfx", 9), ("hooh-atz", 10), ("benzene-adz", 11),
                                      ("synth7", 9), ("synth16u", 10), ("synth24u", 11)])
def test_warp_kernel_cube(G, O, name, cfg):
    """CUBE through the warp kernel with 1, 2 and 3 fragments of 8 MO columns: s...g shells (h2o-aqz), Cartesian d/f
    (hooh-atz), full and partly filled column blocks, UHF + EDF, ragged lattices (bricks cut by the grid edge)."""
    from critic2_b200 import synth
    if name.startswith("synth"):
        u = name.endswith("u")
        f = synth.small_synthetic(uhf=u, with_edf=u, nmo=int(name[5:].rstrip("u")))
    else:
        f = cases.golden_wfn(name)
    st = cases.mol_structure(f)
    x0, xd, nn = cases.bbox_grid(st, (19, 15, 13), margin=3.0)
    ref = _ofield(O, f, st).cube(x0, xd, nn)
    G.set_option("wfn_cfg", cfg)
    try:
        fld = G.WfnField(f, st)
    finally:
        G.set_option("wfn_cfg", -1)
    c = fld.config()
    assert c["cfg"] == cfg and c["nblk"] == 1 and c["points_per_tile"] == 8
    print(name, cfg, _compare_cube(fld.cube(x0, xd, nn), ref))


@pytest.mark.parametrize("cfg", [9, 11])
def test_warp_kernel_points_nder_nci_core_gkin_mo(G, O, cfg):
    """Every driver through the warp kernel: ragged point list with nuclei, nder 0 / 1, NCIPLOT (two-pass and one-pass) with
    its cutoff pass, core augmentation, gkin and MO values (value-only chi modes), interleaved parts."""
    import torch
    from critic2_b200 import drivers
    f = cases.golden_wfn("benzene-adz" if cfg == 11 else "h2o-aqz")
    st = cases.mol_structure(f)
    of = _ofield(O, f, st)
    G.set_option("wfn_cfg", cfg)
    try:
        fld = G.WfnField(f, st)
        assert fld.config()["cfg"] == cfg
        rng = np.random.default_rng(17)
        pts = np.concatenate([st.r[:3], st.r[0] + np.array([[3e-6, 0, 0]]),
                              rng.uniform(st.r.min(0) - 3.0, st.r.max(0) + 3.0, (733, 3))])
        ref = of.grd_points(pts, nder=2, periodic=True)
        got = fld.points(pts, nder=2, periodic=True)
        check_rho(got[:, 0], ref[:, 0])
        check_vec(got[:, 1:4], ref[:, 1:4], "grad", rho=ref[:, 0])
        check_vec(got[:, 4:13], ref[:, 4:13], "hess", rho=ref[:, 0])
        assert np.array_equal(got[:, 15], ref[:, 15]) and ref[:4, 15].all() and not got[:4, 1:4].any()
        x0, xd, nn = cases.bbox_grid(st, (13, 11, 9), margin=2.5)
        grid_pts = np.array([x0 + xd @ np.array([i, j, k], float) for k in range(nn[2]) for j in range(nn[1]) for i in range(nn[0])])
        for nder in (0, 1):
            r = of.grd_points(grid_pts, nder=nder, periodic=False)
            g = fld.cube(x0, xd, nn, nder=nder)
            check_rho(g["f"].reshape(-1), r[:, 0])
            if nder:
                check_vec(np.stack([g[k].reshape(-1) for k in ("gx", "gy", "gz")], -1), r[:, 1:4], "grad", rho=r[:, 0])
        xn, xm, ns = drivers.nci_grid(st, nstep=(17, 15, 14))
        rrho, rgrad, _ = of.nci(xn, xm, ns)
        crho, cgrad, _ = fld.nci(xn, xm, ns)
        assert _nci_compare(crho, cgrad, rrho, rgrad) == 0
        G.set_option("nci_two_pass", 0)
        try:
            c1, g1, _ = fld.nci(xn, xm, ns)
        finally:
            G.set_option("nci_two_pass", 1)
        assert np.array_equal(c1, crho) and np.array_equal(g1, cgrad)
        rg, cg = rgrad.copy(), cgrad.reshape(-1).copy()
        _, rdat = O.nci_cutoffs(rrho.copy(), rg)
        _, dat, ndat = G.nci_cutoffs(crho.reshape(-1), cg)
        assert ndat == len(rdat) and np.array_equal(cg.reshape(rg.shape) == 100.0, rg == 100.0)
        # kinetic energy density and MO values
        rres = of.grd_points(pts, 1, True)
        gk = fld.gkin_points(pts)
        assert np.array_equal(gk == 0.0, rres[:, 17] == 0.0)
        assert (np.abs(gk - rres[:, 17]) / np.maximum(rres[:, 17], 1e-30)).max() <= cases.TOL_DER
        for imo in (1, f.nmoocc):
            refm = of.mo_points(pts, imo, True)
            gotm = fld.mo_points(pts, imo)
            assert not gotm[refm == 0.0].any() and np.max(np.abs(gotm - refm)) <= cases.TOL_RHO * np.abs(refm).max()
        # interleaved parts tile the grid bit for bit
        npt = int(np.prod(nn))
        full = [torch.zeros(npt, dtype=torch.float64, device="cuda") for _ in range(2)]
        fld.cube_dev(x0, xd, nn, full, ("f", "lap"))
        for rows in (False, True):
            acc = [torch.zeros(npt, dtype=torch.float64, device="cuda") for _ in range(2)]
            for part in range(3):
                outs = [torch.full((npt,), float("nan"), dtype=torch.float64, device="cuda") for _ in range(2)]
                fld.cube_dev_part(x0, xd, nn, outs, ("f", "lap"), part, 3, rows=rows)
                torch.cuda.synchronize()
                for a_, o_ in zip(acc, outs):
                    a_ += torch.nan_to_num(o_, nan=0.0)
            assert all(torch.equal(a_, f_) for a_, f_ in zip(acc, full))
        if cfg == 11:   # core augmentation (C zpsp 4)
            ct = cases.core_tables()
            fld.set_core(G.core_field(st, ct, {6: 4, 1: 0}))
            ofc = O.Field(O.Environ(st), wfn=O.Wfn(f), core=O.Grid1Set({6: ct[(6, 4)]}))
            _compare_cube(fld.cube(x0, xd, nn), ofc.cube(x0, xd, nn))
            fld.set_core(None)
    finally:
        G.set_option("wfn_cfg", -1)


@pytest.mark.parametrize("cfg,nmo", [(9, 8), (11, 20), (8, 40), (4, 120), (6, 120), (13, 150), (14, 120)])
def test_g_and_h_functions_through_every_kernel_family(G, O, cfg, nmo):
    """Primitive types 21-56 (g and h shells: l = 4, 5 -- the top of calculate_mo_gto's type table,
    wfn_private@proc.F90:2695-2705; powers up to x^7 in the second derivatives) through the warp, block-synchronous and
    warp-specialised kernels: CUBE with all derivatives and a point list on the nuclei."""
    from critic2_b200 import synth
    f = synth.small_synthetic(nmo=nmo, gh=True)
    assert f.itype.max() == 56
    st = cases.mol_structure(f)
    x0, xd, nn = cases.bbox_grid(st, (14, 12, 11), margin=2.5)
    of = _ofield(O, f, st)
    ref = of.cube(x0, xd, nn)
    G.set_option("wfn_cfg", cfg)
    try:
        fld = G.WfnField(f, st)
    finally:
        G.set_option("wfn_cfg", -1)
    assert fld.config()["cfg"] == cfg
    print(cfg, nmo, _compare_cube(fld.cube(x0, xd, nn), ref))
    pts = np.concatenate([st.r[:4], st.r[:4] + 0.37])
    rres, res = of.grd_points(pts, 2, False), fld.points(pts, 2, False)
    check_rho(res[:, 0], rres[:, 0])
    check_vec(res[:, 4:13], rres[:, 4:13], "hess", rho=rres[:, 0])


@pytest.mark.parametrize("uhf,cfg,nmo", [(True, -1, 20), (True, 6, 230), (False, -1, 20), (True, 8, 50)])
def test_spin_and_valence_channels_of_rho2(G, O, uhf, cfg, nmo):
    """res%fspin / gfmodspin / lapspin / gkinspin and res%fval / gfmodval / del2fval (+ *spinval) of grd for a wavefunction
    field (fieldmod@proc.f90:763-786; rho2 wfn_private@proc.F90:2091-2147, 2159-2200): companion handles with masked
    occupations and a scaled / removed EDF core, against the oracle's full rho2 output."""
    from critic2_b200 import synth
    f = synth.small_synthetic(uhf=uhf, with_edf=True, nmo=nmo)
    st = cases.mol_structure(f)
    env, w = O.Environ(st), O.Wfn(f)
    G.set_option("wfn_cfg", cfg)
    try:
        fld = G.WfnField(f, st)
        rng = np.random.default_rng(31)
        heavy = st.r[f.z > 1][:4]
        pts = np.concatenate([heavy + rng.normal(0, 0.15, heavy.shape), rng.uniform(st.r.min(0) - 2, st.r.max(0) + 2, (150, 3))])
        ref = [w.rho2(env, x, 2) for x in pts]
        assert max(abs(r["rho"][0] - r["rhoval"][0]) for r in ref) > 1e-3          # the EDF core really contributes
        for spin in (0, 1, 2):
            for val in (False, True):
                ch = fld.channel(spin=spin, valence=val)
                res = ch.points(pts, nder=2, periodic=False)
                gk = ch.gkin_points(pts, periodic=False)
                kr, kg, kh = ("rhoval", "gradval", "hval") if val else ("rho", "grad", "h")
                rho = np.array([r[kr][spin] for r in ref])
                grad = np.array([r[kg][spin] for r in ref])
                hh = np.array([r[kh][spin].reshape(-1) for r in ref])
                check_rho(res[:, 0], rho, f"rho spin {spin} val {val}")
                check_vec(res[:, 1:4], grad, "grad", rho=rho)
                check_vec(res[:, 4:13], hh, "hess", rho=rho)
                lap = hh[:, 0] + hh[:, 4] + hh[:, 8]
                hn = np.sqrt((hh ** 2).sum(1))
                assert (np.abs(res[:, 14] - lap) / np.maximum(hn, 1e-4 * np.abs(rho))).max() <= cases.TOL_DER
                gkr = np.array([r["gkin"][spin] for r in ref])                      # no core term in gkin
                assert (np.abs(gk - gkr) / np.maximum(np.abs(gkr), 1e-30)).max() <= cases.TOL_DER
        assert fld.channel() is fld and fld.channel(spin=1) is fld.channel(spin=1)
    finally:
        G.set_option("wfn_cfg", -1)


@pytest.mark.parametrize("cfg", [-1, 6, 8, 13])
def test_exp_neg_over_its_whole_range(G, cfg):
    """The device's own exp for t <= 0 (exp_neg, csrc/wfn_device.cuh: table + degree-6 polynomial; library exp below -700)
    through the kernels that use it: one s primitive with its cutoff radius pushed out until exp(-a r^2) underflows, so the
    arguments run from 0 to -1168.  rho = 2 exp(-2 a r^2) against numpy; the argument a r^2 itself carries the rounding of the
    coordinates (cell frame, ~80 bohr), an error that grows with |t|, hence the |t|-proportional bound.  (With the reference's cutoffs |t| stays below 46.)"""
    from critic2_b200 import wfn
    a = 0.73
    f = wfn.Molwfn(natom=1, xyz=np.zeros((1, 3)), z=np.array([2], np.int32), icenter=np.array([1], np.int32),
                   itype=np.array([1], np.int32), e=np.array([a]), occ=np.array([2.0]), cmo=np.array([[1.0]]), nmoocc=1)
    wfn.complete_struct(f)
    f.dran = np.array([40.0])                      # a r^2 up to 1168: past the underflow of exp
    f.spcutoff = np.array([40.0])
    st = cases.mol_structure(f, border=45.0)          # every point inside the molecular cell
    r = np.concatenate([np.linspace(0.0, 32.0, 4001), np.sqrt(np.array([690.0, 699.9, 700.1, 708.0, 744.0, 760.0]) / a)])
    u = np.array([0.36, 0.48, 0.8])                # |u| = 1
    pts = st.r[0][None, :] + r[:, None] * u[None, :]
    G.set_option("wfn_cfg", cfg)
    try:
        fld = G.WfnField(f, st)
    finally:
        G.set_option("wfn_cfg", -1)
    got = fld.points(pts, nder=2, periodic=False)
    d2 = ((pts - st.r[0]) ** 2).sum(1)
    chi = np.exp(-a * d2)
    ref = 2.0 * chi * chi
    big = ref > 1e-290
    assert big.sum() > 2500 and (~big).sum() > 50      # rho = 2 chi^2 is checked down to chi = 1e-145 (t = -334); beyond: no NaN,
                                                       # clean underflow (t down to -1168 runs the library fallback)
    rel = np.abs(got[big, 0] - ref[big]) / ref[big]
    # the displacement carries the rounding of the cell-frame coordinates (~80 bohr: ~3e-14 bohr), which enters t = 2 a r^2 as
    # 4 a r x 3e-14: the bound follows it, so what is tested is that exp_neg adds nothing visible on top
    bound = 3e-14 + 4.0 * a * r[big] * 3e-14
    assert np.all(rel < bound), (cfg, float((rel / bound).max()), float(r[big][(rel / bound).argmax()]))
    assert np.all(got[~big, 0] <= 1e-289) and np.all(np.isfinite(got))
    # gradient of the same function: -4 a r rho along u
    gref = (-4.0 * a * ref)[:, None] * (pts - st.r[0])
    m = big & (r > 0.5)
    gb = 2.0 * (3e-14 + 4.0 * a * r[m] * 3e-14) + 3e-14 / r[m]      # + the displacement's own relative rounding
    assert np.all(np.abs(got[m, 1:4] - gref[m]).max(1) / np.abs(gref[m]).max(1) < gb)
