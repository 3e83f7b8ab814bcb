"""Device-resident timings of BASELINE.json configs 1, 2, 3 and 5 (config 4 is bench.py's headline line).

One entry per config (a list of dicts; printed as JSON lines when run standalone): points/s, the algorithmic work of SURVEY.md 8(d) counted on the device with the reference's
own screens (n_act / n_at), and the fraction of the bounding roofline.  Inputs are resident in HBM, CUDA events on the
launching stream, W warm-up passes, an L2 flush between passes.  Reads only tests/golden and the synthetic generators
(nothing under /root/reference) and never touches oracle/: the CPU baseline of each config is supplied by bench.py
(`python bench.py --aux configs`, and the "aux" array of bench.py's headline line), which is the one place allowed to run
the oracle; the same object computes the `parity` block of every entry from the GPU values handed to it."""
import argparse
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from critic2_b200 import drivers, gpu, structure, synth  # noqa: E402
import cases  # noqa: E402

FP64_DMMA_TFLOPS, FP64_DFMA_TFLOPS = 37.1, 34.0    # profiles/r1_fp64_peak_microbench.log
HBM_GBS = 6552.0                                   # MEASURED_PEAKS.json hbm_gbs (copy bandwidth)


def main(argv=None, cpu=None):
    """cpu: optional object with cube(f, st, x0, xd, nn, props), nci(f, st, x0, xmat, ns), promol(st, tables, x0, xd, nn, nder)
    returning (cpu_baseline, parity) (bench.py supplies it).  Returns the list of entries."""
    global a, flush, stream, dev
    entries = []
    ap = argparse.ArgumentParser()
    ap.add_argument("--cfgs", default="1,2,3,5")
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--force", default="", help="kernel configuration per BASELINE config, e.g. 1:0,2:7 (development A/B)")
    ap.add_argument("--cfg5-parts", type=int, default=8, help="time part 0 of this many interleaved parts of config 5")
    a = ap.parse_args(argv)

    dev = torch.device("cuda", torch.cuda.current_device())
    gpu.init_devices([dev.index])
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    stream = torch.cuda.current_stream().cuda_stream


    def timed(fn):
        for _ in range(a.warmup):
            flush.zero_()
            fn()
        torch.cuda.synchronize()
        ms = []
        for _ in range(a.steps):
            flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            fn()
            e1.record()
            torch.cuda.synchronize()
            ms.append(e0.elapsed_time(e1))
        return float(np.mean(ms))


    def emit(d):
        entries.append(d)


    def zeros(n, k):
        return [torch.zeros(n, dtype=torch.float64, device=dev) for _ in range(k)]


    cfgs = [int(c) for c in a.cfgs.split(",")]
    force = {int(k): int(v) for k, v in (kv.split(":") for kv in a.force.split(",") if kv)}

    def wfn_field(which, f, st):
        gpu.set_option("wfn_cfg", force.get(which, -1))
        try:
            return gpu.WfnField(f, st)
        finally:
            gpu.set_option("wfn_cfg", -1)

    if 1 in cfgs:   # h2o-aqz, CUBE CELL 128^3, f + gradient + Laplacian (one nder=2 evaluation)
        f = cases.golden_wfn("h2o-aqz")
        st = structure.molecular_cell(f.xyz, f.z)
        fld = wfn_field(1, f, st)
        x0, xd, nn = drivers.cube_grid(st, [128] * 3)
        props = ["f", "gx", "gy", "gz", "lap"]
        npts = int(np.prod(nn))
        outs = zeros(npts, len(props))
        nact = fld.count_active(x0, xd, nn)
        flops = 20.0 * nact * f.nmoocc + 80.0 * nact + 60.0 * f.nmoocc * npts
        ms = timed(lambda: fld.cube_dev(x0, xd, nn, outs, props, stream=stream))
        base, par = cpu.cube(f, st, x0, xd, nn, props, got={p_: o_.cpu().numpy() for p_, o_ in zip(props, outs)}) if cpu else (None, None)
        emit({"cfg": 1, "workload": "h2o-aqz.fchk (238 primitives, 5 MOs), CUBE CELL 128^3, f+grad+lap", "points": npts,
              "ms": ms, "points_per_s": npts / ms * 1e3, "mean_n_act": nact / npts, "kernel": fld.config(),
              "cpu_baseline": base, "parity": par,
              "roofline": {"bound": "fp64 (primitive evaluation: DFMA + exp; the 5-column contraction is negligible)",
                           "achieved_tflops": flops / ms / 1e9, "peak_tflops": FP64_DFMA_TFLOPS,
                           "frac": flops / ms / 1e9 / FP64_DFMA_TFLOPS,
                           "hbm_gbs": 8.0 * len(props) * npts / ms / 1e6, "hbm_frac": 8.0 * len(props) * npts / ms / 1e6 / HBM_GBS}})
        fld.close()
        del outs

    if 2 in cfgs:   # benzene-adz NCIPLOT 200^3 + cutoff pass
        f = cases.golden_wfn("benzene-adz")
        st = structure.molecular_cell(f.xyz, f.z)
        fld = wfn_field(2, f, st)
        x0, xmat, ns = drivers.nci_grid(st, nstep=[200] * 3)
        npts = int(np.prod(ns))
        crho, cgrad = zeros(npts, 2)
        # n_act on the same lattice (count_active takes x-fastest grids; the count is order independent)
        nact = fld.count_active(x0, xmat, ns)
        flops = 20.0 * nact * f.nmoocc + 80.0 * nact + (60.0 * f.nmoocc + 150.0) * npts
        ms = timed(lambda: fld.nci_dev(x0, xmat, ns, crho, cgrad, stream=stream))
        lib = gpu.lib()
        import ctypes as C
        sums = np.zeros(3)
        nd = C.c_long(0)
        sr = np.array([-1e30, 1e30])
        dat = torch.zeros(2 * npts, dtype=torch.float64, device=dev)
        cg2 = cgrad.clone()

        def cut():
            cg2.copy_(cgrad)
            gpu._check(lib.c2gpu_nci_cutoffs_dev(npts, C.c_void_p(crho.data_ptr()), C.c_void_p(cg2.data_ptr()), None, 0.2, 2.0,
                                                 0.1, gpu._d(sr), 0, -1.0, gpu._d(sums), C.byref(nd),
                                                 C.c_void_p(dat.data_ptr()), npts, C.c_void_p(stream)))
        ms_cut = timed(cut)
        ms_copy = timed(lambda: cg2.copy_(cgrad))
        ndat_full = int(nd.value)
        base, par = None, None
        if cpu:   # every 8th x-plane: the GPU's values and its cutoff pass there against the CPU oracle on the same sub-lattice
            n1, plane = int(ns[0]), int(ns[1]) * int(ns[2])
            ii = np.arange(0, n1, 8)
            idx = torch.from_numpy((ii[:, None] * plane + np.arange(plane)[None, :]).reshape(-1)).to(dev)
            g_rho, g_grad = crho.index_select(0, idx).contiguous(), cgrad.index_select(0, idx).contiguous()
            gm = g_grad.clone()
            s2 = np.zeros(3)
            nd2 = C.c_long(0)
            gpu._check(lib.c2gpu_nci_cutoffs_dev(idx.numel(), C.c_void_p(g_rho.data_ptr()), C.c_void_p(gm.data_ptr()), None, 0.2,
                                                 2.0, 0.1, gpu._d(sr), 0, -1.0, gpu._d(s2), C.byref(nd2),
                                                 C.c_void_p(dat.data_ptr()), idx.numel(), C.c_void_p(stream)))
            torch.cuda.synchronize()
            xm = np.array(xmat, float).copy()
            xm[:, 0] *= 8.0
            par, base = cpu.nci_parity(f, st, x0, xm, np.array([len(ii), ns[1], ns[2]], np.int32),
                                       [g_rho.cpu().numpy(), g_grad.cpu().numpy(), None], gm.cpu().numpy(), s2, int(nd2.value),
                                       -1.0, f"{len(ii)} of {n1} x-planes")
        emit({"cfg": 2, "workload": "benzene-adz.fchk (324 primitives, 21 MOs), NCIPLOT 200^3 (crho, cgrad) + cutoff pass",
              "points": npts, "ms": ms, "points_per_s": npts / ms * 1e3, "mean_n_act": nact / npts, "kernel": fld.config(),
              "ndat": ndat_full, "cpu_baseline": base, "parity": par,
              "roofline": {"bound": "fp64 (primitive evaluation)", "achieved_tflops": flops / ms / 1e9,
                           "peak_tflops": FP64_DFMA_TFLOPS, "frac": flops / ms / 1e9 / FP64_DFMA_TFLOPS},
              "cutoff_pass": {"ms": ms_cut - ms_copy, "bound": "hbm", "bytes": 24.0 * npts,
                              "achieved_gbs": 24.0 * npts / max(ms_cut - ms_copy, 1e-6) / 1e6, "peak_gbs": HBM_GBS,
                              "note": "3 streaming kernels: 2 reads of crho/cgrad + masked cgrad write; .dat pairs extra"}})
        fld.close()
        del crho, cgrad, dat, cg2

    if 3 in cfgs:   # quartz 4x4x4 promolecular, CUBE CELL 256^3: rho only (promolecular_grid) and rho+grad+Hessian
        st = cases.quartz(4)
        fld = gpu.PromolField(st, cases.tables())
        nn = np.array([256] * 3, np.int32)
        npts = int(np.prod(nn))
        x0, xd, _ = drivers.cube_grid(st, nn)
        nat = fld.count_active(x0, xd, nn, periodic=True)
        out = zeros(npts, 1)[0]
        ms0 = timed(lambda: fld.promolecular_grid_dev(nn, out, stream=stream))
        outs = zeros(npts, 10)
        p10 = gpu.PROPS[:4] + gpu.PROPS[5:11]
        ms2 = timed(lambda: fld.cube_dev(x0, xd, nn, outs, p10, periodic=True, stream=stream))
        for nder, ms, nout in ((0, ms0, 1), (2, ms2, 10)):
            flops = 120.0 * nat
            base, par = None, None
            if cpu:
                arrs = [out] if nder == 0 else outs

                def got(flat, arrs=arrs):
                    ix = torch.from_numpy(np.ascontiguousarray(flat)).to(dev)
                    return torch.stack([a_.index_select(0, ix) for a_ in arrs], 1).cpu().numpy()
                base, par = cpu.promol(st, cases.tables(), x0, xd, nn, nder, got=got)
            emit({"cfg": 3, "workload": f"alpha-quartz 4x4x4 (576 atoms in cell, {fld.nenv} with images), promolecular, 256^3, nder={nder}",
                  "points": npts, "ms": ms, "points_per_s": npts / ms * 1e3, "mean_n_at": nat / npts,
                  "cpu_baseline": base, "parity": par,
                  "roofline": {"bound": "fp64 alu/sfu (log, sqrt, reciprocals, Lagrange weights); HBM floor = output stream",
                               "achieved_tflops": flops / ms / 1e9, "peak_tflops": FP64_DFMA_TFLOPS,
                               "frac": flops / ms / 1e9 / FP64_DFMA_TFLOPS, "hbm_gbs": 8.0 * nout * npts / ms / 1e6,
                               "hbm_frac": 8.0 * nout * npts / ms / 1e6 / HBM_GBS}})
        fld.close()

    if 5 in cfgs:   # 2000 atoms, 50 400 primitives, 3 120 MOs: NCIPLOT 512^3 + promolecular rhoat, 1 of 8 interleaved parts
        f = synth.config5()
        st = structure.molecular_cell(f.xyz, f.z)
        fld = gpu.WfnField(f, st)
        pro = gpu.PromolField(st, cases.tables())
        x0, xmat, ns = drivers.nci_grid(st, nstep=[512] * 3)
        npts = int(np.prod(ns))
        crho, cgrad, rhoat = zeros(npts, 3)
        parts = a.cfg5_parts
        lh = fld.layer_height()
        nlay = (int(ns[0]) + lh - 1) // lh
        mine = [L for L in range(0, nlay, parts)]
        mypts = sum(min(int(ns[0]), (L + 1) * lh) - L * lh for L in mine) * int(ns[1]) * int(ns[2])
        # n_act of this part: x-layers of the NCI lattice = z-layers of the transposed x-fastest lattice
        xmat_t = xmat[:, ::-1].copy()
        ns_t = ns[::-1].copy()
        nact = sum(fld.count_active(x0, xmat_t, ns_t, iz0=L * lh, iz1=min(int(ns[0]), (L + 1) * lh)) for L in mine)
        flops = 20.0 * nact * f.nmoocc + 80.0 * nact + (60.0 * f.nmoocc + 150.0) * mypts
        a.warmup, a.steps = 1, 2
        ms = timed(lambda: fld.nci_dev_part(x0, xmat, ns, crho, cgrad, 0, parts, rhoat=rhoat, promol=pro, stream=stream))
        ms_w = timed(lambda: fld.nci_dev_part(x0, xmat, ns, crho, cgrad, 0, parts, stream=stream))
        emit({"cfg": 5, "workload": f"synthetic C640N80O80H1200 (50400 primitives, 3120 MOs), wfn NCIPLOT + promolecular rhoat, "
                                    f"512^3, part 0 of {parts} interleaved x-layers (one GPU's share of an {parts}-GPU run)",
              "points": mypts, "ms": ms, "points_per_s": mypts / ms * 1e3, "ms_wfn_only": ms_w, "mean_n_act": nact / mypts,
              "kernel": fld.config(),
              "roofline": {"bound": "tensor (FP64 DMMA)", "achieved_tflops": flops / ms_w / 1e9, "peak_tflops": FP64_DMMA_TFLOPS,
                           "frac": flops / ms_w / 1e9 / FP64_DMMA_TFLOPS},
              "whole_grid_estimate_s_on_8_gpus": ms / 1e3})
        fld.close()
        pro.close()
    torch.cuda.empty_cache()
    return entries


if __name__ == "__main__":
    for d_ in main():
        print(json.dumps(d_), flush=True)
