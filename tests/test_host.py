"""CPU tests of the host side: the C-ABI library loads and exports what include/c2gpu.h declares, the loaders'
tables agree with the oracle's independent complete_struct, the driver grid definitions follow the reference, and the
multi-GPU slab logic works over a world_size-2 gloo group.  No compute call is made on the (absent) GPU."""
import ctypes
import os
import re
import socket
import sys

import numpy as np
import pytest

import cases

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_abi_library_exports_every_declared_symbol():
    from critic2_b200 import gpu
    L = ctypes.CDLL(gpu.LIB_PATH)
    hdr = open(os.path.join(ROOT, "include", "c2gpu.h")).read()
    names = sorted(set(re.findall(r"^\s*(?:int|long)\s+(c2gpu_\w+)\s*\(", hdr, re.M)))
    assert len(names) >= 20
    for n in names:
        assert hasattr(L, n), f"{n} declared in include/c2gpu.h but not exported"
    # the Fortran binding names the same symbols
    f90 = open(os.path.join(ROOT, "fortran", "gpu_interface.f90")).read()
    for n in re.findall(r'name="(c2gpu_\w+)"', f90):
        assert n in names, f"fortran/gpu_interface.f90 binds unknown symbol {n}"


def test_no_cpu_fallback_without_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from critic2_b200 import gpu
    f = cases.golden_wfn("h2o-wfx")
    st = cases.mol_structure(f)
    with pytest.raises(gpu.C2GpuError):
        gpu.WfnField(f, st)
    with pytest.raises(gpu.C2GpuError):
        gpu.nci_cutoffs(np.zeros(8), np.zeros(8))


def test_product_never_imports_the_oracle():
    for dp, _, files in os.walk(os.path.join(ROOT, "critic2_b200")):
        for fn in files:
            if fn.endswith((".py", ".cu", ".cuh", ".h", ".inc")):
                src = open(os.path.join(dp, fn)).read()
                assert not re.search(r"^\s*(import|from)\s+oracle\b", src, re.M), fn
                assert "liboracle" not in src and "oracle.h" not in src, fn


def test_complete_struct_tables_match_oracle():
    import oracle
    for name in ("h2o-aqz", "benzene-adz", "hooh-atz"):
        f = cases.golden_wfn(name)
        dran, spc, lmax, icord, lo, hi = oracle.Wfn(f).tables()
        np.testing.assert_array_equal(dran, f.dran)
        np.testing.assert_array_equal(spc, f.spcutoff)
        np.testing.assert_array_equal(lmax, f.lmax)
        np.testing.assert_array_equal(icord, f.icord)
        np.testing.assert_array_equal(lo, f.iprilo)
        np.testing.assert_array_equal(hi, f.iprihi)


def test_wfx_roundtrip_and_synthetic_shapes(tmp_path):
    from critic2_b200 import synth, wfn
    f = synth.small_synthetic(with_edf=True)
    p = str(tmp_path / "s.wfx")
    wfn.write_wfx(f, p)
    g = wfn.read_wfx(p)
    assert g.npri == f.npri and g.nmoocc == f.nmoocc and g.nedf == f.nedf
    np.testing.assert_allclose(g.cmo, f.cmo, rtol=1e-14)
    np.testing.assert_allclose(g.e, f.e, rtol=1e-14)
    np.testing.assert_array_equal(g.itype, f.itype)
    f4 = synth.config4()
    assert (f4.natom, f4.npri, f4.nmoocc) == (500, 12600, 780)


def test_driver_grid_definitions():
    from critic2_b200 import drivers
    f = cases.golden_wfn("benzene-adz")
    st = cases.mol_structure(f)
    o, xd, nn = drivers.cube_grid(st, (10, 12, 14))
    np.testing.assert_allclose(o, 0)                      # CUBE CELL: x0 = cell origin
    np.testing.assert_allclose(xd @ nn.astype(float), st.m_x2c @ np.ones(3))   # end point excluded
    x0, xmat, ns = drivers.nci_grid(st)                   # xinc 0.1 bohr default
    assert np.allclose(np.diag(xmat), 0.1) and (ns > 50).all()
    lo = st.m_x2c @ st.molborder
    np.testing.assert_allclose(x0, lo, rtol=1e-12)
    q = cases.quartz(1)
    x0, xmat, ns = drivers.nci_grid(q, nstep=(10, 10, 11))
    np.testing.assert_allclose(xmat[:, 2] * 11, q.m_x2c[:, 2])
    assert len(q.environment_atoms()[0]) > 9 * 27


def test_rank_layers_cover_the_grid():
    from critic2_b200 import sharding
    for nz, lh, world in ((400, 8, 8), (101, 8, 3), (5, 4, 2), (512, 8, 8)):
        seen = np.zeros(nz, int)
        for r in range(world):
            for c0, c1 in sharding.rank_layers(nz, lh, r, world):
                seen[c0:c1] += 1
        assert (seen == 1).all()
        assert sum(sharding.rank_points(nz, lh, r, world, 7) for r in range(world)) == nz * 7


def _gloo_worker(rank, world, port, q):
    import torch
    import torch.distributed as dist
    sys.path.insert(0, ROOT)
    from critic2_b200 import sharding
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    # the gather of the one-process-per-GPU model: every rank copies the blocks it owns into ONE shared host buffer
    # (SharedPinned; here without cudaHostRegister) -- layers and rows of super-tiles
    nz, ny, nx, lh = 21, 13, 6, 4
    truth = np.arange(nz * ny * nx, dtype=np.float64) * 0.5 + 1.0
    sp = sharding.SharedPinned(rank, world, dist, register=False)
    gathered = []
    for k, rows in enumerate((False, True)):
        shared = sp.get(truth.nbytes, f"t{k}")
        mine = np.full(truth.shape, 7.0)                              # garbage outside the rank's blocks
        sharding.copy_owned_blocks(mine, truth, nz, ny, lh, rank, world, rows=rows)   # "evaluate" the own blocks
        sharding.copy_owned_blocks(shared, mine, nz, ny, lh, rank, world, rows=rows)  # the gather
        dist.barrier()
        gathered.append(bool(np.array_equal(shared, truth)))
    sums = torch.tensor([1.0 + rank, 2.0, 3.0], dtype=torch.float64)
    sharding.reduce_void_sums(sums, dist)
    ok = bool(torch.equal(sums, torch.tensor([3.0, 4.0, 6.0], dtype=torch.float64)))
    ok = ok and all(gathered)
    dist.barrier()
    sp.close()
    q.put((rank, ok))
    dist.destroy_process_group()


def test_rank_rows_tile_the_grid():
    from critic2_b200 import sharding
    for (nslow, nmid, lh, world) in ((400, 400, 8, 8), (26, 19, 8, 3), (5, 3, 8, 4), (64, 70, 4, 5)):
        seen = np.zeros((nslow, nmid), int)
        sizes = []
        for r in range(world):
            blocks = sharding.rank_rows(nslow, nmid, lh, r, world)
            sizes.append(sum((c1 - c0) * (b1 - b0) for c0, c1, b0, b1 in blocks))
            for c0, c1, b0, b1 in blocks:
                seen[c0:c1, b0:b1] += 1
        assert (seen == 1).all()
        if nslow == 400:   # the bench grid: 2 500 rows over 8 ranks, against 7 vs 6 whole layers
            assert max(sizes) / (sum(sizes) / world) < 1.002
            lay = [sharding.rank_points(nslow, lh, r, world, nmid) for r in range(world)]
            assert max(lay) / (sum(lay) / world) > 1.1


def test_two_rank_gather_and_reduction_over_gloo():
    """world_size-2 gloo run of the N > 1 host logic: shared-buffer gather (layers and rows) + void-sum reduction"""
    import torch.multiprocessing as mp
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_gloo_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    assert all(ok for _, ok in res), res


def test_chk_nci_checkpoint_round_trip(tmp_path, monkeypatch):
    """writechk / readchk (nci@proc.f90:962-1059): gfortran unformatted records, including the subrecord split of a
    record longer than the 2^31-9 byte limit (exercised with a small limit) and the compatibility checks."""
    from critic2_b200 import drivers
    rng = np.random.default_rng(2)
    ns = (5, 4, 3)
    x0, xmat = rng.normal(size=3), rng.normal(size=(3, 3))
    crho, cgrad, rhoat = (rng.normal(size=ns) for _ in range(3))
    rhofrag = rng.normal(size=(2,) + ns)
    p = str(tmp_path / "a.chk_nci")
    drivers.write_chk_nci(p, x0, xmat, ns, crho, cgrad, rhoat, rhofrag)
    raw = open(p, "rb").read()
    # first record: three 4-byte values between two 4-byte markers of 12
    assert np.frombuffer(raw[:20], np.int32).tolist() == [12, 1, 1, 2, 12]
    assert len(raw) == (12 + 8) + 4 * (24 + 8) + (12 + 8) + (2 * 480 + 8) + (480 + 8) + (960 + 8)
    got = drivers.read_chk_nci(p, x0, xmat, ns, want_rhoat=True, nfrag=2)
    for a, b in zip(got, (crho, cgrad, rhoat, rhofrag)):
        assert np.array_equal(a, b)
    assert drivers.read_chk_nci(p, x0 + 1e-9, xmat, ns, want_rhoat=True, nfrag=2) is None
    assert drivers.read_chk_nci(p, x0, xmat, (5, 4, 4), want_rhoat=True, nfrag=2) is None
    assert drivers.read_chk_nci(p, x0, xmat, ns, want_rhoat=True, nfrag=1) is None
    monkeypatch.setattr(drivers, "_SUBREC", 200)          # force subrecords: 960-byte record -> 5 pieces
    drivers.write_chk_nci(p, x0, xmat, ns, crho, cgrad)
    raw = open(p, "rb").read()
    off = 20 + 4 * 32 + 20
    assert np.frombuffer(raw[off:off + 4], np.int32)[0] == -200
    got = drivers.read_chk_nci(p, x0, xmat, ns)
    assert np.array_equal(got[0], crho) and np.array_equal(got[1], cgrad) and got[2] is None


def test_kernel_configuration_choice_follows_the_measured_cost_model():
    """wfn_choose_config (no device needed): the warp kernel with 1 / 2 / 3 column fragments up to 24 MOs, then the cheapest of M2 (64-column blocks, two CTAs
    per SM), L1 (112-column blocks, two CTAs per SM) and the two warp-specialised kernels by the measured per-block costs (csrc/wfn_launch.cu)."""
    import ctypes as C
    from critic2_b200 import gpu
    L = C.CDLL(gpu.LIB_PATH)
    pick = lambda n: L.c2gpu_wfn_choose_config(n)
    assert [pick(n) for n in (1, 5, 8)] == [9, 9, 9]
    assert [pick(n) for n in (9, 16, 17, 21, 24)] == [10, 10, 11, 11, 11]
    assert [pick(n) for n in (25, 32, 33, 48, 49, 64)] == [16, 16, 17, 17, 8, 8]   # M2 with 32- / 48- / 64-column blocks
    # WS64g (64-column blocks, seven generator warps) from 65 MOs on since its generators evaluate shell-wise
    assert [pick(n) for n in (65, 100, 112, 113, 128, 129, 150, 192, 193, 224, 225, 256, 300, 320, 330, 350, 384)] == [13] * 17
    assert [pick(n) for n in (385, 448, 512, 570)] == [13] * 4          # the cost model holds up to nine 64-column blocks
    assert [pick(n) for n in (513, 560, 577, 780, 896)] == [6] * 5       # incl. BASELINE config 4
    assert [pick(n) for n in (897, 1120, 1560, 3120)] == [14] * 4        # dense consumer loop from nine blocks; BASELINE config 5


def test_aimpac_wfn_round_trip(tmp_path):
    """read_wfn (wfn_private@proc.F90:648-771): the fixed-column formats 101-106, the wavefunction-type rules and the
    alpha/beta boundary from the orbital-energy drop, through a write/read round trip at the file's text precision."""
    from critic2_b200 import synth, wfn
    f = cases.golden_wfn("h2o-wfx")
    p = str(tmp_path / "h2o.wfn")
    wfn.write_wfn(f, p)
    lines = open(p).read().split("\n")
    assert lines[1] == "GAUSSIAN              5 MOL ORBITALS     59 PRIMITIVES        3 NUCLEI"
    assert lines[2][2:4] == "O " and float(lines[2][70:75]) == 8.0 and len(lines[2]) == 75
    g = wfn.read_wfn(p)
    assert (g.npri, g.nmoocc, g.wfntyp) == (f.npri, f.nmoocc, wfn.WFN_RHF)
    assert np.array_equal(g.icenter, f.icenter) and np.array_equal(g.itype, f.itype) and np.array_equal(g.z, f.z)
    np.testing.assert_allclose(g.e, f.e, rtol=1e-6)
    np.testing.assert_allclose(g.cmo, f.cmo, rtol=0, atol=1e-7 * np.abs(f.cmo).max())
    np.testing.assert_allclose(g.xyz, f.xyz, atol=1e-8)
    np.testing.assert_allclose(g.dran, f.dran, rtol=1e-6)
    # unrestricted: occupations 1, alpha block first; the beta block starts where the orbital energy drops (:741-745)
    u = synth.small_synthetic(uhf=True)
    ene = np.concatenate([np.linspace(-5.0, -0.3, u.nalpha), np.linspace(-4.9, -0.4, u.nmoocc - u.nalpha)])
    q = str(tmp_path / "u.wfn")
    wfn.write_wfn(u, q, energies=ene)
    v = wfn.read_wfn(q)
    assert v.wfntyp == wfn.WFN_UHF and v.nalpha == u.nalpha and v.npri == u.npri


def test_fortran_shim_matches_the_c_header():
    """fortran/gpu_interface.f90 cannot be compiled here (no Fortran compiler), so at least its bind(c) interfaces are
    checked against include/c2gpu.h: every bound name is declared in the header with the same number of arguments, and
    pass-by-value dummies are exactly the header's non-pointer parameters."""
    import re
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    hdr = open(os.path.join(root, "include", "c2gpu.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    cdecl = {}
    for m in re.finditer(r"\b(?:int|long)\s+(c2gpu_\w+)\s*\(([^;]*?)\)\s*;", hdr, flags=re.S):
        args = [a.strip() for a in m.group(2).replace("\n", " ").split(",")]
        args = [] if args == ["void"] else args
        cdecl[m.group(1)] = ["*" in a or "[" in a for a in args]          # True = pointer / array parameter
    src = open(os.path.join(root, "fortran", "gpu_interface.f90")).read()
    src = re.sub(r"&\s*\n\s*&?", "", src)                                   # join continuation lines
    blocks = re.findall(r"function\s+(c2gpu_\w+)\s*\(([^)]*)\)\s*bind\(c,name=\"(\w+)\"\)(.*?)end function", src, flags=re.S)
    assert len(blocks) >= 15
    for fname, arglist, cname, body in blocks:
        assert fname == cname and cname in cdecl, cname
        names = [a.strip() for a in arglist.split(",") if a.strip()]
        assert len(names) == len(cdecl[cname]), f"{cname}: {len(names)} Fortran dummies vs {len(cdecl[cname])} C parameters"
        byval = set()
        for line in body.split("\n"):
            if "value" in line.split("::")[0]:
                byval.update(n.strip().split("(")[0] for n in line.split("::")[1].split(","))
        for n, is_ptr in zip(names, cdecl[cname]):
            # type(c_ptr), value carries a C pointer by value: allowed for pointer parameters (handles, nullable arrays)
            if not is_ptr:
                assert n in byval, f"{cname}: {n} must be passed by value"


def test_fortran_shim_exports_every_name_the_integration_notes_use():
    """Every gpu_* entity INTEGRATION.md's call-site patches use must be PUBLIC in the (private-by-default) module
    gpu_interface, every public name must be defined there, and every c2gpu_* the wrappers call must have a bind(c)
    interface -- the checks a Fortran compiler would make at `use gpu_interface`."""
    import re
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    src = open(os.path.join(root, "fortran", "gpu_interface.f90")).read()
    src_nc = "\n".join(line.split("!")[0] for line in src.split("\n"))           # comments stripped
    src_j = re.sub(r"&\s*\n\s*&?", "", src_nc)
    public = set()
    for m in re.finditer(r"^\s*public\s*::\s*(.*)$", src_j, flags=re.M):
        public.update(n.strip() for n in m.group(1).split(","))
    defined = set(re.findall(r"^\s*(?:subroutine|function)\s+(gpu_\w+)", src_j, flags=re.M))
    defined |= set(re.findall(r"::\s*(gpu_\w+)\s*=", src_j))                      # module variables (gpu_enabled)
    assert public <= defined, public - defined
    notes = open(os.path.join(root, "INTEGRATION.md")).read()
    used = set(re.findall(r"\bgpu_[a-z_0-9]+", notes)) - {"gpu_interface"}
    used = {u for u in used if not u.endswith(".f90")}
    assert used <= public, f"used in INTEGRATION.md but not public: {sorted(used - public)}"
    bound = set(re.findall(r"bind\(c,name=\"(c2gpu_\w+)\"\)", src_j))
    called = set(re.findall(r"\b(c2gpu_\w+)\s*\(", src_j))
    assert called <= bound, called - bound
