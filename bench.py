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
                    help="secondary measurements with their CPU baselines: BASELINE configs 1,2,3,5 (tools/bench_configs.py) "
                         "or the cube-text formatter (tools/bench_format.py); extra arguments go to the tool")
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


def cpu_sample(f, st, lo, xd, nn, stride, reps=1):
    """The CPU oracle (OpenMP, all cores) on the sub-lattice (stride,stride,stride) of the workload grid, nder = 2."""
    import oracle
    oracle.build()
    idx = [np.arange(stride // 2, int(n), stride) for n in nn]
    I, J, K = np.meshgrid(*idx, indexing="ij")
    pts = lo[None, :] + np.stack([I.ravel(), J.ravel(), K.ravel()], 1).astype(float) @ xd.T
    fld = oracle.Field(oracle.Environ(st), wfn=oracle.Wfn(f))
    cores = os.cpu_count() or 1
    t0 = time.perf_counter()
    for _ in range(reps):
        res = fld.grd_points(pts, nder=2, periodic=False, nthreads=cores)
    dt = (time.perf_counter() - t0) / reps
    return pts, res, dt, cores


class CpuLegs:
    """CPU baselines of the secondary measurements (--aux): the oracle on all host cores, bounded samples.  bench.py is
    the only non-test code that runs the oracle; tools/bench_*.py receive this object and never import it."""

    def __init__(self):
        import oracle
        oracle.build()
        self.o = oracle
        self.cores = os.cpu_count() or 1

    def _res(self, npts, dt, sample):
        return {"value": npts / dt, "unit": "points/s", "cores": self.cores, "kind": "port", "sample": sample}

    def cube(self, f, st, x0, xd, nn, props):
        fld = self.o.Field(self.o.Environ(st), wfn=self.o.Wfn(f))
        t0 = time.perf_counter()
        fld.cube(x0, xd, nn, props=tuple(props), periodic=False, nthreads=self.cores)
        return self._res(int(np.prod(nn)), time.perf_counter() - t0, "full grid, one nder=2 evaluation per point")

    def nci(self, f, st, x0, xmat, ns):
        fld = self.o.Field(self.o.Environ(st), wfn=self.o.Wfn(f))
        sub = np.array([max(1, int(ns[0]) // 8), int(ns[1]), int(ns[2])], np.int32)     # every 8th x-plane
        xm = np.array(xmat, float).copy()
        xm[:, 0] *= float(ns[0]) / float(sub[0])
        t0 = time.perf_counter()
        fld.nci(x0, xm, sub, nthreads=self.cores)
        return self._res(int(np.prod(sub)), time.perf_counter() - t0, f"{sub[0]} of {int(ns[0])} x-planes, NCIPLOT evaluation loop")

    def promol(self, st, tables, x0, xd, nn, nder):
        fld = self.o.Field(self.o.Environ(st), grids=self.o.Grid1Set(tables))
        sub = np.array([int(nn[0]), int(nn[1]), max(1, int(nn[2]) // 16)], np.int32)  # every 16th z-plane
        xs = np.array(xd, float).copy()
        xs[:, 2] *= float(nn[2]) / float(sub[2])
        idx = [np.arange(int(n)) for n in sub]
        I, J, K = np.meshgrid(*idx, indexing="ij")
        pts = np.asarray(x0, float)[None, :] + np.stack([I.ravel(), J.ravel(), K.ravel()], 1).astype(float) @ xs.T
        t0 = time.perf_counter()
        fld.grd_points(pts, nder=nder, periodic=True, nthreads=self.cores)
        return self._res(len(pts), time.perf_counter() - t0, f"{sub[2]} of {int(nn[2])} z-planes, nder={nder}")

    def text(self, sample, fmt):
        t0 = time.perf_counter()
        self.o.cube_text_c(sample, fmt, 1, nthreads=self.cores)
        return {"value": sample.size / (time.perf_counter() - t0), "unit": "values/s", "cores": self.cores, "kind": "port",
                "sample": f"{sample.shape[0]}^3 values, C oracle (glibc %.*E, OpenMP)"}


def run_aux(a, rest):
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    cpu = None if a.no_cpu else CpuLegs()
    if a.aux == "configs":
        import bench_configs
        bench_configs.main(rest, cpu)
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
            "config": {"workload": f"cfg4: synthetic C160N20O20H300, 12600 primitives, 780 MOs, rho+grad+Hessian, {a.n}^3 grid",
                       "sampled": sample},
            "cpu_baseline": {"value": v, "unit": "points/s", "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": v, "unit": "points/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def run_ours(a):
    import torch
    import torch.distributed as dist
    from critic2_b200 import gpu

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the hot path has no CPU fallback")
    torch.cuda.set_device(local)
    if world > 1:
        # rank 0 prints exactly one JSON line on stdout: NCCL's log (its version banner is printed from level VERSION up,
        # WARN included) goes to stderr
        os.environ["NCCL_DEBUG"] = "WARN"
        os.environ["NCCL_DEBUG_FILE"] = "/dev/stderr"
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    gpu.init_devices([local])
    dev = torch.device("cuda", local)

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

    def step():
        flush.zero_()
        fld.cube_dev_part(lo, xd, nn, outs, PROPS10, rank, world, stream=stream, rows=world > 1)

    def sync():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    for _ in range(a.warmup):
        step()
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

    # ---- e2e through the C-ABI host entry (rank-local share at N>1, then NCCL reduce to the writer rank) ----------
    e2e = None
    if not a.no_e2e:
        if world == 1:
            host = fld.cube(lo, xd, nn, props=PROPS10)     # warm-up call (device output arrays are kept by the handle)
            del host
            t0 = time.perf_counter()
            host = fld.cube(lo, xd, nn, props=PROPS10)
            te = time.perf_counter() - t0
            d2h = sum(v.nbytes for v in host.values())
            chk = float(host["f"].sum())
            del host
        else:
            pin = torch.empty(npts_total, dtype=torch.float64).pin_memory() if rank == 0 else None
            for o in outs:
                o.zero_()
            sync()
            t0 = time.perf_counter()
            fld.cube_dev_part(lo, xd, nn, outs, PROPS10, rank, world, stream=stream, rows=True)
            d2h = 0
            for o in outs:   # layers are disjoint across ranks: a sum-reduce to the writer rank is an exact gather
                dist.reduce(o, dst=0, op=dist.ReduceOp.SUM)
                if rank == 0:
                    pin.copy_(o)
                    d2h += pin.numel() * 8
            torch.cuda.synchronize()
            dist.barrier()
            te = time.perf_counter() - t0
            chk = float(outs[0].sum().item())
        tt = torch.tensor([te], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        e2e = {"value": npts_total / float(tt.item()), "unit": "points/s", "h2d_bytes_per_step": 3 * 8 + 9 * 8 + 3 * 4,
               "d2h_bytes_per_step": int(d2h), "ms": float(tt.item()) * 1e3, "checksum_rho": chk,
               "note": "field handle (LOAD) resident; inputs per step are the grid descriptor; outputs to fresh pageable host arrays; one warm-up call"
                       if world == 1 else "per-rank layers + NCCL reduce to rank 0 + D2H to pinned host on rank 0"}

    cpu = None
    if rank == 0 and not a.no_cpu:
        pts, _, dt, cores = cpu_sample(f, st, lo, xd, nn, a.cpu_stride)
        cpu = {"value": len(pts) / dt, "unit": "points/s", "cores": cores, "kind": "port",
               "sample": f"sub-lattice stride {a.cpu_stride} of the {a.n}^3 grid ({len(pts)} points, {dt:.1f} s), nder=2"}

    if rank == 0:
        ach = flops / (kms * 1e-3) / 1e12
        line = {"metric": "grid points/sec (rho+grad+Hessian)", "value": value, "unit": "points/s", "n_gpus": world,
                "steps": a.steps, "warmup": a.warmup, "ms_per_step": ms_total / a.steps, "higher_is_better": True,
                "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                "config": {"workload": f"cfg4: synthetic C160N20O20H300 (seed 4), def2-TZVP-shaped 12600 primitives, 780 MOs, "
                                       f"rho+grad+Hessian (10 outputs/point), {a.n}^3 grid",
                           "parallelism": (f"z-layers of {lh} planes on 1 GPU" if world == 1 else
                                           f"rows of super-tiles ({lh} planes x {lh} y-lines) round-robin over {world} GPUs"),
                           "l2": f"256 MiB flush write between steps; outputs {80.0 * npts_total / 1e9:.2f} GB/step vs 126 MB L2",
                           "kernel": cfg},
                "roofline": {"bound": "tensor", "achieved": ach, "peak": FP64_PEAK_TFLOPS, "unit": "TFLOP/s",
                             "frac": ach / FP64_PEAK_TFLOPS,
                             # dram__bytes_read.sum + dram__bytes_write.sum of one whole-grid launch on one GPU (ncu,
                             # profiles/r1_ncu_dram_wfn_ws_kernel_400.txt): chi-tile cache spills, not algorithmic bytes
                             "traffic": 1.026e13 if (a.n == 400 and world == 1 and cfg["cfg"] == 6) else None,
                             "traffic_unit": "bytes/launch",
                             "kernel": f"{'wfn_ws_kernel' if cfg['cfg'] == 6 else 'wfn_grid_kernel'} cfg {cfg['cfg']} (FP64 DMMA, {cfg['threads']} threads, "
                                       f"P={cfg['points_per_tile']}, NB={cfg['mo_block']})", "kernel_ms": kms,
                             "algorithmic_flops_per_launch": flops, "mean_n_act": nact / max(mypts, 1),
                             "peak_source": "measured FP64 mma.sync (DMMA) peak, profiles/r1_fp64_peak_microbench.log; "
                                            "MEASURED_PEAKS.json holds no FP64 figure"},
                "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": int(launches), "clocks": clocks}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    args, rest = parse()
    if args.aux:
        run_aux(args, rest)
    elif args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)
