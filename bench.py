#!/usr/bin/env python
"""bench.py -- grid points/s of the critic2 field-evaluation hot path on B200 (BASELINE.json metric).

Workload (N = any): BASELINE.json configs[3] -- synthetic 500-atom C160N20O20H300 molecule, def2-TZVP-shaped basis
(12 600 primitives), 780 random MOs, rho + grad + Hessian (10 doubles per point) on a 400^3 grid.  One "step" is one
full pass over the grid; with N GPUs the 50 z-layers of 8 planes are dealt round-robin to the ranks (strong scaling,
no data-path collective; the gather to the writer rank is part of the e2e leg only).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--grid 400]

`value`    : points/s, device-resident outputs, CUDA events on the launching stream, max over ranks.
`e2e`      : the same pass through the C-ABI host entry (c2gpu_eval_grid: host descriptors in, pageable host arrays
             out, device->host copies inside the timed region).
`roofline` : algorithmic FP64 flops (sum_p 2*10*n_act(p)*nmo + 80*n_act(p) + 60*nmo, n_act counted
             exactly on the device with the reference's own screen) / kernel time, against the FP64 tensor-core (DMMA)
             peak measured on this pool's B200 (profiles/r1_fp64_peak_microbench.log: MEASURED_PEAKS.json has no FP64).
`cpu_baseline` / --impl reference: the CPU oracle (C + OpenMP restatement of the reference path; the Fortran reference
             cannot be built: no Fortran compiler) on a bounded sub-lattice of the same grid, all host cores.
`parity`   : the e2e result at the CPU sample's points against the oracle's values (rho <= 1e-10, derivatives <= 1e-8 of
             the vector / tensor norm, identical exact-zero masks); the run FAILS (exit 1) when it is violated.
`aux`      : BASELINE configs 1, 2, 3 (N = 1) and 5 (N = 8) with their own roofline, cpu_baseline and parity blocks.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

FP64_PEAK_TFLOPS = 37.1   # mma.sync m8n8k4 f64 chains, profiles/r1_fp64_peak_microbench.log (DFMA chains: 34.0)
PROPS10 = ("f", "gx", "gy", "gz", "hxx", "hxy", "hxz", "hyy", "hyz", "hzz")
TOL_RHO, TOL_DER = 1e-10, 1e-8     # BASELINE.json north_star


def workload_name(n):
    """config.workload of BOTH arms (the driver compares the strings)"""
    return (f"cfg4: synthetic C160N20O20H300 (seed 4), def2-TZVP-shaped 12600 primitives, 780 MOs, rho+grad+Hessian "
            f"(10 outputs/point), {n}^3 grid")


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=2)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--grid", dest="n", type=int, default=400, help="grid points per axis (400 = the BASELINE config)")
    ap.add_argument("--cfg", type=int, default=-1, help="force a wavefunction kernel configuration (development)")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--cpu-stride", type=int, default=8, help="sub-lattice stride of the CPU sample")
    ap.add_argument("--aux", default="", choices=["", "configs", "format"],
                    help="run ONLY a secondary measurement with its CPU baselines: BASELINE configs 1,2,3,5 "
                         "(tools/bench_configs.py) or the cube-text formatter (tools/bench_format.py); extra arguments go to the tool")
    ap.add_argument("--no-aux", action="store_true", help="skip the aux entries (configs 1,2,3 at N=1; 5 at N=8) of the line")
    ap.add_argument("--aux-cfg5", action="store_true", help="run the config-5 aux entry at any N (default: N = 8 only)")
    ap.add_argument("--e2e-calls", type=int, default=3, help="timed end-to-end calls (after one warm-up call)")
    return ap.parse_known_args()


def workload(n):
    from critic2_b200 import structure, synth
    f = synth.config4()
    st = structure.molecular_cell(f.xyz, f.z)
    lo, hi = st.r.min(0) - 6.0, st.r.max(0) + 6.0
    nn = np.array([n, n, n], np.int32)
    xd = np.diag((hi - lo) / nn)
    return f, st, lo, xd, nn


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled during the timed region (B200_PROFILING.md)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows, self.proc = [], None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "200", "-i", str(index)], stdout=subprocess.PIPE, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=3)
        except Exception:
            self.proc.kill()
        sm, mx, reasons, pw = [], [], set(), []
        names = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx.append(float(r[1])); pw.append(float(r[2]))
                for nm, v in zip(names, r[3:7]):
                    if v.lower().startswith("active"):
                        reasons.add(nm)
            except Exception:
                pass
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(pw) if pw else None, "samples": len(sm), "reasons": sorted(reasons)}


def sample_index(nn, stride):
    """flat x-fastest indices and (i, j, k) of the sub-lattice (stride // 2 + stride * m) of an nn grid"""
    idx = [np.arange(stride // 2, int(n), stride) for n in nn]
    I, J, K = np.meshgrid(*idx, indexing="ij")
    ijk = np.stack([I.ravel(), J.ravel(), K.ravel()], 1)
    flat = ijk[:, 0] + int(nn[0]) * (ijk[:, 1] + int(nn[1]) * ijk[:, 2])
    return flat, ijk


def cpu_sample(f, st, lo, xd, nn, stride, reps=1):
    """The CPU oracle (OpenMP, all cores) on the sub-lattice (stride,stride,stride) of the workload grid, nder = 2."""
    import oracle
    oracle.build()
    _, ijk = sample_index(nn, stride)
    pts = lo[None, :] + ijk.astype(float) @ xd.T
    fld = oracle.Field(oracle.Environ(st), wfn=oracle.Wfn(f))
    cores = os.cpu_count() or 1
    t0 = time.perf_counter()
    for _ in range(reps):
        res = fld.grd_points(pts, nder=2, periodic=False, nthreads=cores)
    dt = (time.perf_counter() - t0) / reps
    return pts, res, dt, cores


# ---- parity blocks -------------------------------------------------------------------------------------------------------
def _vec_err(got, ref, rho):
    """max over points of |got - ref| / max(|ref vector|, 1e-4 rho): the measure of tests/cases.py check_vec"""
    nrm = np.sqrt((ref ** 2).sum(-1))
    floor = np.maximum(1e-4 * np.abs(rho), 1e-20)
    return float((np.abs(got - ref).max(-1) / np.maximum(nrm, floor)).max())


def parity_fields(rho, rho_ref, vecs, sample):
    """rho: (n,), vecs: [(name, got (n,k), ref (n,k))] -> the `parity` object of a bench line"""
    rho, rho_ref = np.asarray(rho), np.asarray(rho_ref)
    zero_eq = bool(np.array_equal(rho == 0.0, rho_ref == 0.0))
    e_rho = float((np.abs(rho - rho_ref) / np.maximum(np.abs(rho_ref), 1e-30)).max())
    out = {"points": int(rho.size), "sample": sample, "rho_max_rel": e_rho, "zero_mask_equal": zero_eq,
           "nonzero_points": int(np.count_nonzero(rho_ref))}
    worst = 0.0
    for name, got, ref in vecs:
        e = _vec_err(np.asarray(got), np.asarray(ref), rho_ref)
        out[f"{name}_max_rel"] = e
        worst = max(worst, e)
    out["deriv_max_rel"] = worst
    out["tol"] = {"rho": TOL_RHO, "deriv": TOL_DER}
    out["ok"] = bool(zero_eq and e_rho <= TOL_RHO and worst <= TOL_DER)
    return out


def parity_cube10(vals, res, sample):
    """vals: (n, 10) GPU f, g(3), h(xx,xy,xz,yy,yz,zz); res: oracle grd rows (f, gf(3), hf(3,3), ...)"""
    hcols = [4, 5, 6, 8, 9, 12]
    return parity_fields(vals[:, 0], res[:, 0], [("grad", vals[:, 1:4], res[:, 1:4]), ("hess", vals[:, 4:10], res[:, hcols])],
                         sample)


class CpuLegs:
    """CPU baselines and parity checks of the secondary measurements: the oracle on all host cores, bounded samples.
    bench.py is the only non-test code that runs the oracle; tools/bench_*.py receive this object and never import it."""

    def __init__(self):
        import oracle
        oracle.build()
        self.o = oracle
        self.cores = os.cpu_count() or 1

    def _res(self, npts, dt, sample):
        return {"value": npts / dt, "unit": "points/s", "cores": self.cores, "kind": "port", "sample": sample}

    def cube(self, f, st, x0, xd, nn, props, got=None):
        """full-grid CUBE on the CPU; got: {prop: GPU array} -> (cpu_baseline, parity)"""
        fld = self.o.Field(self.o.Environ(st), wfn=self.o.Wfn(f))
        t0 = time.perf_counter()
        ref = fld.cube(x0, xd, nn, periodic=False, nthreads=self.cores)          # all 12 properties of one nder=2 pass
        base = self._res(int(np.prod(nn)), time.perf_counter() - t0, "full grid, one nder=2 evaluation per point")
        par = None
        if got is not None:
            r = {k: v.reshape(-1) for k, v in ref.items()}
            g = {k: np.asarray(v).reshape(-1) for k, v in got.items()}
            hk = ("hxx", "hxy", "hxz", "hyy", "hyz", "hzz")
            hn = np.sqrt(sum(r[k] ** 2 for k in hk))
            par = parity_fields(g["f"], r["f"], [("grad", np.stack([g[k] for k in ("gx", "gy", "gz")], -1),
                                                  np.stack([r[k] for k in ("gx", "gy", "gz")], -1))], "every point of the grid")
            e_lap = float((np.abs(g["lap"] - r["lap"]) / np.maximum(np.maximum(hn, 1e-4 * np.abs(r["f"])), 1e-20)).max())
            par["lap_max_rel_of_hessian_norm"] = e_lap
            par["deriv_max_rel"] = max(par["deriv_max_rel"], e_lap)
            par["ok"] = bool(par["ok"] and e_lap <= TOL_DER)
        return base, par

    def nci_parity(self, f, st, x0, xmat, ns, got, got_masked, got_sums, got_ndat, rho_void, sample,
                   rhocut=0.2, dimcut=2.0, rhoplot=0.1):
        """NCIPLOT on the (sub-)lattice x0, xmat, ns on the CPU against the GPU's values there.  got: [crho, cgrad, rhoat or
        None] (flat, z fastest) BEFORE the cutoff pass; got_masked / got_sums / got_ndat: what the GPU cutoff pass made of
        them.  Classification = .dat membership count, the cgrad = 100 mask and sign(lambda2)."""
        o = self.o
        env = o.Environ(st)
        fld = o.Field(env, wfn=o.Wfn(f))
        pro = o.Field(env, grids=o.Grid1Set(_tables())) if got[2] is not None else None
        t0 = time.perf_counter()
        rrho, rgrad, rat = fld.nci(x0, xmat, ns, promol=pro, nthreads=self.cores)
        base = self._res(int(np.prod(ns)), time.perf_counter() - t0, sample + ", NCIPLOT evaluation loop")
        rrho, rgrad = rrho.reshape(-1), rgrad.reshape(-1)
        crho, cgrad = np.asarray(got[0]).reshape(-1), np.asarray(got[1]).reshape(-1)
        rho = np.maximum(np.abs(rrho) / 100.0, 1e-80)
        floor = 1e-4 * rho / (2.0 * (3.0 * np.pi ** 2) ** (1.0 / 3.0) * rho ** (4.0 / 3.0))
        big = rgrad < 1e40
        par = {"points": int(rrho.size), "sample": sample,
               "rho_max_rel": float((np.abs(np.abs(crho) - np.abs(rrho)) / np.maximum(np.abs(rrho), 1e-30)).max()),
               "zero_mask_equal": bool(np.array_equal(crho == 0.0, rrho == 0.0)),
               "rdg_max_rel": float((np.abs(cgrad - rgrad) / np.maximum(np.abs(rgrad), floor))[big].max()),
               "sign_lambda2_flips": int(np.count_nonzero(np.sign(crho) != np.sign(rrho)))}
        ok = par["zero_mask_equal"] and par["rho_max_rel"] <= TOL_RHO and par["rdg_max_rel"] <= TOL_DER and \
            par["sign_lambda2_flips"] == 0
        if rat is not None:
            cat = np.asarray(got[2]).reshape(-1)
            rat = rat.reshape(-1)
            par["rhoat_max_rel"] = float((np.abs(cat - rat) / np.maximum(np.abs(rat), 1e-30)).max())
            ok = ok and par["rhoat_max_rel"] <= TOL_RHO and bool(np.array_equal(cat == 0.0, rat == 0.0))
        rg = rgrad.copy().reshape(tuple(int(v) for v in ns))
        rsums, rdat = o.nci_cutoffs(rrho.copy().reshape(rg.shape), rg, None if rat is None else rat.reshape(rg.shape),
                                    rhocut=rhocut, dimcut=dimcut, rhoplot=rhoplot, rho_void=rho_void, omega=float(rrho.size))
        par["dat_count"] = [int(got_ndat), int(len(rdat))]
        par["masked_points"] = [int(np.count_nonzero(np.asarray(got_masked) == 100.0)), int(np.count_nonzero(rg == 100.0))]
        par["classification_identical"] = bool(got_ndat == len(rdat) and
                                               np.array_equal(np.asarray(got_masked).reshape(-1) == 100.0, rg.reshape(-1) == 100.0))
        if rho_void > 0:      # oracle sums carry omega / npts = 1 here: raw sums on both sides
            par["void_sums_max_rel"] = float((np.abs(np.asarray(got_sums) - rsums) / np.maximum(np.abs(rsums), 1e-300)).max())
            ok = ok and par["void_sums_max_rel"] <= TOL_RHO
        par["deriv_max_rel"] = par["rdg_max_rel"]
        par["tol"] = {"rho": TOL_RHO, "deriv": TOL_DER}
        par["ok"] = bool(ok and par["classification_identical"])
        return par, base

    def promol(self, st, tables, x0, xd, nn, nder, got=None, zstride=16):
        """every zstride-th z-plane on the CPU; got: callable(flat indices) -> (n, 1 or 10) GPU values there"""
        fld = self.o.Field(self.o.Environ(st), grids=self.o.Grid1Set(tables))
        kz = np.arange(0, int(nn[2]), zstride)
        I, J, K = np.meshgrid(np.arange(int(nn[0])), np.arange(int(nn[1])), kz, indexing="ij")
        ijk = np.stack([I.ravel(), J.ravel(), K.ravel()], 1)
        pts = np.asarray(x0, float)[None, :] + ijk.astype(float) @ np.asarray(xd, float).T
        t0 = time.perf_counter()
        res = fld.grd_points(pts, nder=nder, periodic=True, nthreads=self.cores)
        sample = f"{len(kz)} of {int(nn[2])} z-planes, nder={nder}"
        base = self._res(len(pts), time.perf_counter() - t0, sample)
        par = None
        if got is not None:
            flat = ijk[:, 0] + int(nn[0]) * (ijk[:, 1] + int(nn[1]) * ijk[:, 2])
            vals = got(flat)
            par = parity_cube10(vals, res, sample) if nder == 2 else parity_fields(vals[:, 0], res[:, 0], [], sample)
        return base, par

    def text(self, sample, fmt):
        t0 = time.perf_counter()
        self.o.cube_text_c(sample, fmt, 1, nthreads=self.cores)
        return {"value": sample.size / (time.perf_counter() - t0), "unit": "values/s", "cores": self.cores, "kind": "port",
                "sample": f"{sample.shape[0]}^3 values, C oracle (glibc %.*E, OpenMP)"}


def _tables():
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import cases
    return cases.tables()


def run_aux(a, rest):
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    cpu = None if a.no_cpu else CpuLegs()
    if a.aux == "configs":
        import bench_configs
        for d in bench_configs.main(rest, cpu):
            print(json.dumps(d), flush=True)
    else:
        import bench_format
        bench_format.main(rest, cpu)


def run_reference(a):
    """--impl reference: the reference path's CPU implementation (oracle port, OpenMP on all host cores)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    f, st, lo, xd, nn = workload(a.n)
    times = []
    for i in range(a.warmup + a.steps):
        # warm-up steps use a coarser sample so the whole run stays within minutes
        stride = a.cpu_stride * 2 if i < a.warmup else a.cpu_stride
        pts, _, dt, cores = cpu_sample(f, st, lo, xd, nn, stride)
        if i >= a.warmup:
            times.append((len(pts), dt))
    npts = sum(t[0] for t in times)
    tot = sum(t[1] for t in times)
    v = npts / tot
    sample = f"sub-lattice stride {a.cpu_stride} of the {a.n}^3 grid ({times[0][0]} points per step), nder=2"
    line = {"impl": "reference", "metric": "grid points/sec (rho+grad+Hessian)", "value": v, "unit": "points/s",
            "n_gpus": a.gpus, "steps": a.steps, "warmup": a.warmup, "ms_per_step": tot / a.steps * 1e3,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": workload_name(a.n), "sampled": sample},
            "cpu_baseline": {"value": v, "unit": "points/s", "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": v, "unit": "points/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def run_ours(a):
    import torch
    import torch.distributed as dist
    from critic2_b200 import gpu, sharding

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the hot path has no CPU fallback")
    torch.cuda.set_device(local)
    shared = None
    # stdout carries exactly one JSON line: everything native libraries print there (NCCL writes its version banner to
    # stdout whatever NCCL_DEBUG_FILE says) is sent to stderr; the line itself goes to the saved descriptor
    sys.stdout.flush()
    json_out = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    if world > 1:
        os.environ.setdefault("NCCL_DEBUG", "WARN")      # the caller's level (e.g. INFO for a rank check) is kept
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        shared = sharding.SharedPinned(rank, world, dist)
    gpu.init_devices([local])
    dev = torch.device("cuda", local)
    peak_now = gpu.fp64_peak() if rank == 0 else None       # (DMMA, DFMA) TFLOP/s measured on this GPU, now

    f, st, lo, xd, nn = workload(a.n)
    if a.cfg >= 0:
        gpu.set_option("wfn_cfg", a.cfg)
    fld = gpu.WfnField(f, st)
    cfg = fld.config()
    npts_total = int(nn[0]) * int(nn[1]) * int(nn[2])
    outs = [torch.zeros(npts_total, dtype=torch.float64, device=dev) for _ in PROPS10]
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)   # > 126 MB L2
    stream = torch.cuda.current_stream().cuda_stream

    # exact algorithmic work of this rank's layers (the reference's own per-point screen, counted on the device)
    lh = fld.layer_height()
    nlayers = (int(nn[2]) + lh - 1) // lh
    nact, mypts = 0.0, 0
    for L in range(rank, nlayers, world):
        z0, z1 = L * lh, min(int(nn[2]), (L + 1) * lh)
        nact += fld.count_active(lo, xd, nn, iz0=z0, iz1=z1)
        mypts += (z1 - z0) * int(nn[0]) * int(nn[1])
    if world > 1:   # rows of super-tiles are interleaved (finer than layers): this GPU's share = total / world (+-0.2 %)
        tt = torch.tensor([nact, float(mypts)], dtype=torch.float64, device=dev)
        dist.all_reduce(tt)
        nact, mypts = float(tt[0].item()) / world, float(tt[1].item()) / world
    flops = 2.0 * 10 * nact * f.nmoocc + 80.0 * nact + 60.0 * f.nmoocc * mypts

    def sync():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    for _ in range(a.warmup):
        flush.zero_()
        fld.cube_dev_part(lo, xd, nn, outs, PROPS10, rank, world, stream=stream, rows=world > 1)
    sync()
    sampler = ClockSampler(local) if rank == 0 else None
    l0 = gpu.launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    kev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(a.steps)]
    e0.record()
    for i in range(a.steps):
        flush.zero_()
        kev[i][0].record()
        fld.cube_dev_part(lo, xd, nn, outs, PROPS10, rank, world, stream=stream, rows=world > 1)
        kev[i][1].record()
    e1.record()
    sync()
    launches = gpu.launch_count() - l0
    clocks = sampler.stop() if sampler else None
    ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
    kms = float(np.mean([s.elapsed_time(e) for s, e in kev]))
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms_total = float(ms.item())
    value = npts_total * a.steps / (ms_total * 1e-3)

    # ---- e2e: results in HOST memory.  N = 1: the C-ABI host entry c2gpu_eval_grid (host descriptors in, fresh pageable
    # host arrays out).  N > 1 (one process per GPU): c2gpu_eval_grid_dev_part_rows + c2gpu_gather_part_host, every rank
    # copying the rows it owns over its own PCIe link into ONE shared page-locked host buffer (the writer rank's arrays) --
    # no inter-GPU traffic, 1/N of the bytes per link. ----------------------------------------------------------------------
    e2e, host_vals, e2e_sp = None, None, None
    flat_idx, _ = sample_index(nn, a.cpu_stride)
    if not a.no_e2e:
        tes = []
        if world == 1:
            for it in range(1 + a.e2e_calls):               # first call = warm-up (device output arrays kept by the handle)
                t0 = time.perf_counter()
                host = fld.cube(lo, xd, nn, props=PROPS10)
                if it > 0:
                    tes.append(time.perf_counter() - t0)
                d2h = sum(v.nbytes for v in host.values())
                if it == a.e2e_calls:
                    chk = float(host["f"].sum())
                    host_vals = np.stack([host[p].reshape(-1)[flat_idx] for p in PROPS10], 1)
                del host
            note = ("field handle (LOAD) resident; inputs per step are the grid descriptor; outputs to fresh pageable host "
                    f"arrays; mean of {a.e2e_calls} calls after one warm-up call")
        else:
            hbuf = shared.get(len(PROPS10) * npts_total * 8, "cfg4")
            hptr = [hbuf.ctypes.data + k * npts_total * 8 for k in range(len(PROPS10))]
            d2h = len(PROPS10) * npts_total * 8
            # two z-slabs when the per-rank copy is long enough (2-3 ranks: >= 70 ms) to pay for the tail of a second launch
            # (super-tiles take ~14 ms each on this molecule); measured at N = 2: 6135 -> 6120 ms; one slab from 4 ranks up
            nsl = 2 if world <= 3 else 1
            zcut = [min(int(nn[2]), ((int(nn[2]) * k // nsl + lh - 1) // lh) * lh) for k in range(nsl + 1)]
            slabs = [(zcut[k], zcut[k + 1]) for k in range(nsl) if zcut[k + 1] > zcut[k]]
            for it in range(1 + a.e2e_calls):
                sync()
                t0 = time.perf_counter()
                for z0, z1 in slabs:      # the copy of slab s runs behind the kernel of slab s + 1
                    fld.cube_dev_part_range(lo, xd, nn, outs, PROPS10, rank, world, z0, z1, stream=stream, rows=True)
                    fld.gather_part_host_range(nn, outs, hptr, rank, world, z0, z1, rows=True, async_=True, stream=stream)
                fld.gather_wait()
                dist.barrier()
                if it > 0:
                    tes.append(time.perf_counter() - t0)
            hv = hbuf.reshape(len(PROPS10), npts_total)
            if rank == 0:
                chk = float(hv[0].sum())
                host_vals = np.stack([hv[k][flat_idx] for k in range(len(PROPS10))], 1)
            note = (f"per-rank rows of super-tiles in {len(slabs)} z-slabs (c2gpu_eval_grid_dev_part_range), each followed by an "
                    f"asynchronous c2gpu_gather_part_host_range into one shared page-locked host buffer ({world} PCIe links in "
                    f"parallel, no inter-GPU traffic); mean of {a.e2e_calls} steps after one warm-up step")
        tt = torch.tensor([float(np.mean(tes)), float(np.min(tes)), float(np.max(tes))], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        if rank == 0:
            e2e = {"value": npts_total / float(tt[0].item()), "unit": "points/s", "h2d_bytes_per_step": 3 * 8 + 9 * 8 + 3 * 4,
                   "d2h_bytes_per_step": int(d2h), "ms": float(tt[0].item()) * 1e3, "ms_min": float(tt[1].item()) * 1e3,
                   "ms_max": float(tt[2].item()) * 1e3, "calls": a.e2e_calls, "checksum_rho": chk, "note": note}

        # ---- the same pass through the single-process host entry over all N devices (what the Fortran caller gets from
        # gpu_init -> c2gpu_init(0)): rank 0 alone drives every GPU, the other ranks wait on the host ------------------------
        if world > 1:
            outs = None
            torch.cuda.empty_cache()
            store = dist.distributed_c10d._get_default_store()
            sync()
            if rank == 0:
                try:
                    gpu.init(world)
                    fsp = gpu.WfnField(f, st)
                    tsp = []
                    for it in range(1 + a.e2e_calls):
                        t0 = time.perf_counter()
                        host = fsp.cube(lo, xd, nn, props=PROPS10)
                        if it > 0:
                            tsp.append(time.perf_counter() - t0)
                        same = bool(np.array_equal(np.stack([host[p].reshape(-1)[flat_idx] for p in PROPS10], 1), host_vals))
                        del host
                    e2e_sp = {"value": npts_total / float(np.mean(tsp)), "unit": "points/s", "ms": float(np.mean(tsp)) * 1e3,
                              "ms_min": float(np.min(tsp)) * 1e3, "calls": a.e2e_calls, "devices": world,
                              "equals_per_process_result_at_sample": same,
                              "note": "c2gpu_init(N) + c2gpu_eval_grid from ONE process over all N devices, pageable host "
                                      "arrays out; the other ranks idle"}
                    fsp.close()
                except Exception as ex:      # reported, not fatal: the headline numbers stand
                    e2e_sp = {"error": str(ex)}
                finally:
                    gpu.init_devices([local])
                    store.set("c2gpu_sp_done", "1")
            else:
                import datetime
                store.wait(["c2gpu_sp_done"], datetime.timedelta(minutes=20))

    cpu, parity = None, None
    if rank == 0 and not a.no_cpu:
        pts, res, dt, cores = cpu_sample(f, st, lo, xd, nn, a.cpu_stride)
        sample = f"sub-lattice stride {a.cpu_stride} of the {a.n}^3 grid ({len(pts)} points, {dt:.1f} s), nder=2"
        cpu = {"value": len(pts) / dt, "unit": "points/s", "cores": cores, "kind": "port", "sample": sample}
        if host_vals is None and world == 1:     # --no-e2e: the device-resident arrays of the timed region
            idx = torch.from_numpy(flat_idx).to(dev)
            host_vals = torch.stack([o.index_select(0, idx) for o in outs], 1).cpu().numpy()
        if host_vals is not None:
            parity = parity_cube10(host_vals, res, sample)
            parity["source"] = "e2e host arrays" if e2e else "device arrays of the timed region"

    # ---- aux: the other BASELINE configs (1, 2, 3 on one GPU; 5 on eight) ------------------------------------------------
    aux = []
    if not a.no_aux:
        sys.path.insert(0, os.path.join(ROOT, "tools"))
        fld.close()
        outs = flush = None
        torch.cuda.empty_cache()
        legs = CpuLegs() if (rank == 0 and not a.no_cpu) else None
        if world == 1:
            import bench_configs
            aux += bench_configs.main(["--cfgs", "1,2,3"], legs)
        if world == 8 or a.aux_cfg5:
            import bench_cfg5
            d5 = bench_cfg5.run(rank, world, dev, dist if world > 1 else None, steps=1, warmup=1, cpu=legs,
                                shared_host=shared.get if shared else None)
            if rank == 0:
                aux.append(d5)

    ok = True
    if rank == 0:
        ach = flops / (kms * 1e-3) / 1e12
        line = {"metric": "grid points/sec (rho+grad+Hessian)", "value": value, "unit": "points/s", "n_gpus": world,
                "steps": a.steps, "warmup": a.warmup, "ms_per_step": ms_total / a.steps, "higher_is_better": True,
                "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                "config": {"workload": workload_name(a.n),
                           "parallelism": (f"z-layers of {lh} planes on 1 GPU" if world == 1 else
                                           f"rows of super-tiles ({lh} planes x {lh} y-lines) round-robin over {world} GPUs"),
                           "l2": f"256 MiB flush write between steps; outputs {80.0 * npts_total / 1e9:.2f} GB/step vs 126 MB L2",
                           "kernel": cfg},
                "roofline": {"bound": "tensor", "achieved": ach, "peak": FP64_PEAK_TFLOPS, "unit": "TFLOP/s",
                             "frac": ach / FP64_PEAK_TFLOPS,
                             "peak_measured_in_run": {"dmma_tflops": peak_now[0], "dfma_tflops": peak_now[1],
                                                      "how": "c2gpu_fp64_peak: independent mma.sync m8n8k4 f64 / DFMA chains, "
                                                             "best of 3, on this GPU before the timed region"},
                             "traffic": TRAFFIC.get((a.n, world, cfg["cfg"])), "traffic_unit": "bytes/launch",
                             "traffic_source": TRAFFIC_SOURCE if (a.n, world, cfg["cfg"]) in TRAFFIC else None,
                             "kernel": f"{'wfn_ws_kernel' if cfg['cfg'] in (6, 12, 13, 14) else 'wfn_grid_kernel'} cfg {cfg['cfg']} (FP64 DMMA, {cfg['threads']} threads, "
                                       f"P={cfg['points_per_tile']}, NB={cfg['mo_block']})", "kernel_ms": kms,
                             "algorithmic_flops_per_launch": flops, "mean_n_act": nact / max(mypts, 1),
                             "peak_source": "measured FP64 mma.sync (DMMA) peak, profiles/r1_fp64_peak_microbench.log; "
                                            "MEASURED_PEAKS.json holds no FP64 figure"},
                "cpu_baseline": cpu, "parity": parity, "e2e": e2e, "gpu_launches": int(launches), "clocks": clocks}
        if e2e_sp is not None:
            line["e2e_single_process"] = e2e_sp
        if aux:
            line["aux"] = aux
        print(json.dumps(line), file=json_out, flush=True)
        bad = [("headline", parity)] + [(f"cfg{d.get('cfg')}", d.get("parity")) for d in aux]
        for name, p in bad:
            if p is not None and not p.get("ok", True):
                ok = False
                print(f"bench.py: PARITY FAILURE in {name}: {json.dumps(p)}", file=sys.stderr, flush=True)
    if shared:
        shared.close()
    if world > 1:
        dist.destroy_process_group()
    if not ok:
        raise SystemExit(1)


# dram__bytes_read.sum + dram__bytes_write.sum of ONE whole-grid launch on one GPU from an `ncu --set full` capture (not
# measured in the run): keyed by (grid, n_gpus, kernel cfg); chi-tile cache spills, not algorithmic bytes
TRAFFIC = {(400, 1, 6): 1.069e13}
TRAFFIC_SOURCE = "ncu, profiles/r2_ncu_dram_wfn_ws_kernel_400.txt"


if __name__ == "__main__":
    args, rest = parse()
    if args.aux:
        run_aux(args, rest)
    elif args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)
