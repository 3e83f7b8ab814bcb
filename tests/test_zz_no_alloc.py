"""c2gpu_nci_cutoffs_dev_noalloc: the allocation-free variant of the NCIPLOT cutoff pass (static __device__ scratch,
no counterpart in the reference) against the existing c2gpu_nci_cutoffs_dev through
the C ABI.  Same kernels on the same grid, so the results must be bit-identical (well inside the 1e-10 / 1e-8 tolerances of
the other GPU tests).  Collected last on purpose: a failure here hides no other test.  Runs alone in
a fresh process as well as after the other tests."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _inputs(torch, n, seed):
    """crho = sign(lambda2) * 100 rho, cgrad = RDG, rhoat: magnitudes around the NCIPLOT thresholds so that every branch of
    the pass (.dat membership, cgrad = 100 masking, VOID sums) fires on a good share of the points"""
    g = torch.Generator(device="cuda").manual_seed(seed)
    u = lambda: torch.rand(n, dtype=torch.float64, device="cuda", generator=g)
    crho = (u() - 0.5) * 60.0                  # |crho| < rhocut*100 = 20 on ~2/3 of the points, > rhoplot*100 = 10 on ~2/3
    cgrad = u() * 4.0                          # dimcut = 2
    rhoat = u() * 0.1                          # rho_void = 0.05
    return crho, cgrad, rhoat


def _run(L, gpu, fn, crho, cgrad, rhoat, rho_void, want_dat=True):
    import torch
    n = crho.numel()
    cg = cgrad.clone()
    dat = torch.full((2 * n,), -1.0, dtype=torch.float64, device="cuda") if want_dat else None
    sums = np.zeros(3)
    nd = C.c_long(-1)
    sr = np.array([-1e30, 1e30])
    rc = fn(n, C.c_void_p(crho.data_ptr()), C.c_void_p(cg.data_ptr()),
            C.c_void_p(rhoat.data_ptr()) if rhoat is not None else None, 0.2, 2.0, 0.1, gpu._d(sr), 0, rho_void, gpu._d(sums),
            C.byref(nd), C.c_void_p(dat.data_ptr()) if want_dat else None, n if want_dat else 0,
            C.c_void_p(torch.cuda.current_stream().cuda_stream))
    gpu._check(rc)
    torch.cuda.synchronize()
    return sums.copy(), int(nd.value), cg, dat


def test_noalloc_cutoff_pass_matches_the_allocating_entry_on_two_calls_in_a_row():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no GPU")
    from critic2_b200 import gpu
    torch.cuda.set_device(0)
    L = gpu.lib()
    # two calls in a row on the shared static scratch: a large ragged grid, then a smaller one (which finds the first
    # call's per-block entries in the array and must not be affected by them), then the large one again
    for k, (n, seed, rho_void) in enumerate(((211 * 157 * 31 + 5, 1, 0.05), (77 * 41 + 3, 2, -1.0), (211 * 157 * 31 + 5, 3, 0.05))):
        crho, cgrad, rhoat = _inputs(torch, n, seed)
        ra = rhoat if rho_void > 0 else None
        ref = _run(L, gpu, L.c2gpu_nci_cutoffs_dev, crho, cgrad, ra, rho_void)
        got = _run(L, gpu, L.c2gpu_nci_cutoffs_dev_noalloc, crho, cgrad, ra, rho_void)
        assert got[1] == ref[1] and 0 < ref[1] < n, (k, got[1], ref[1])
        assert np.array_equal(got[0], ref[0]), (k, got[0], ref[0])              # void sums (raw): identical
        if rho_void > 0:
            assert ref[0][2] > 0
        assert torch.equal(got[2], ref[2]) and bool((ref[2] == 100.0).any())     # masked cgrad
        assert torch.equal(got[3][:2 * ref[1]], ref[3][:2 * ref[1]])             # (rho, s) pairs in the reference's order
        assert bool((got[3][2 * ref[1]:] == -1.0).all())                        # nothing written past the pair list
        # independent check of the classification against numpy (nci@proc.f90:593-598)
        cr, cg0 = crho.cpu().numpy(), cgrad.cpu().numpy()
        inter = np.ones(n, bool) if rho_void <= 0 else rhoat.cpu().numpy() < rho_void
        assert ref[1] == int(np.count_nonzero((np.abs(cr) < 20.0) & (cg0 < 2.0) & inter))
    # without a pair buffer
    crho, cgrad, rhoat = _inputs(torch, 5000, 9)
    a = _run(L, gpu, L.c2gpu_nci_cutoffs_dev, crho, cgrad, None, -1.0, want_dat=False)
    b = _run(L, gpu, L.c2gpu_nci_cutoffs_dev_noalloc, crho, cgrad, None, -1.0, want_dat=False)
    assert a[1] == b[1] and torch.equal(a[2], b[2])


def test_noalloc_cutoff_pass_rejects_grids_beyond_its_static_scratch():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no GPU")
    from critic2_b200 import gpu
    torch.cuda.set_device(0)
    L = gpu.lib()
    small = torch.zeros(16, dtype=torch.float64, device="cuda")
    sums, nd, sr = np.zeros(3), C.c_long(0), np.array([-1e30, 1e30])
    too_many = 134217728 + 1                                                  # C2GPU_NOALLOC_MAX_POINTS + 1
    rc = L.c2gpu_nci_cutoffs_dev_noalloc(too_many, C.c_void_p(small.data_ptr()), C.c_void_p(small.data_ptr()), None, 0.2, 2.0,
                                         0.1, gpu._d(sr), 0, -1.0, gpu._d(sums), C.byref(nd), None, 0, None)
    assert rc == 1                                                            # rejected before anything is launched
    buf = C.create_string_buffer(512)
    L.c2gpu_last_error(buf, 512)
    msg = buf.value.decode()
    assert "exceed the static scratch limit" in msg and "134217728" in msg
    torch.cuda.synchronize()
    assert not small.any()
    # empty grid: nothing to do, status 0
    assert L.c2gpu_nci_cutoffs_dev_noalloc(0, None, None, None, 0.2, 2.0, 0.1, gpu._d(sr), 0, -1.0, gpu._d(sums), C.byref(nd),
                                           None, 0, None) == 0 and nd.value == 0
