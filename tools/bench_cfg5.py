"""BASELINE.json configs[4]: synthetic 2000-atom wavefunction NCIPLOT + promolecular rhoat on 512^3, x-layer-sharded over
the GPUs of one box (one process per GPU under torchrun; also runs on 1 GPU).

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29512 \
        tools/bench_cfg5.py --steps 2 --warmup 1

A step = the whole NCIPLOT data path of nci@proc.f90:499-604: crho/cgrad (wavefunction, nder = 2 + eig) and rhoat
(promolecular) on this rank's interleaved x-layers, the cutoff pass with VOID on the same layers, then the only
communication the path has: NCCL all-reduce of the void sums / .dat count, and the gather to the writer:
every rank copies its layers over its own PCIe link into ONE shared page-locked host buffer (c2gpu_gather_part_host).
`run()` is also called by bench.py (its "aux" entry for config 5 at N = 8), which supplies the CPU legs (the oracle is
never imported here).  Timing = wall clock between barriers (the step ends in host memory), max over ranks."""
import argparse
import ctypes as C
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from critic2_b200 import drivers, gpu, sharding, structure, synth  # noqa: E402
import cases  # noqa: E402

RHO_VOID = 0.05
FP64_DMMA_TFLOPS = 37.1


def run(rank, world, dev, dist=None, grid=512, steps=2, warmup=1, small=False, cpu=None, shared_host=None):
    """cpu: optional object with nci_sample(f, st, x0, xmat, ns, sub) -> (rrho, rgrad, rat, cutoffs(...), baseline dict)
    (bench.py's CpuLegs).  shared_host: callable(nbytes, tag) -> (numpy float64 view of a page-locked host buffer shared by
    all ranks); None: rank-local pinned memory (single process)."""
    f = synth.config4() if small else synth.config5()
    st = structure.molecular_cell(f.xyz, f.z)
    fld = gpu.WfnField(f, st)
    pro = gpu.PromolField(st, cases.tables())
    x0, xmat, ns = drivers.nci_grid(st, nstep=[grid] * 3)
    n1, n2, n3 = (int(v) for v in ns)
    npts = n1 * n2 * n3
    plane = n2 * n3
    lh = fld.layer_height()
    layers = sharding.rank_layers(n1, lh, rank, world)
    crho, cgrad, rhoat = (torch.zeros(npts, dtype=torch.float64, device=dev) for _ in range(3))
    dat = torch.zeros(2 * lh * plane, dtype=torch.float64, device=dev)
    stream = torch.cuda.current_stream().cuda_stream
    lib = gpu.lib()
    sr = np.array([-1e30, 1e30])
    if shared_host is not None:
        host = shared_host(3 * npts * 8, "cfg5")
    else:
        host = torch.empty(3 * npts, dtype=torch.float64).pin_memory().numpy()
    hptr = [host.ctypes.data + k * npts * 8 for k in range(3)]

    def cutoffs(c0, c1, tot):
        n = (c1 - c0) * plane
        off = c0 * plane * 8
        sums = np.zeros(3)
        nd = C.c_long(0)
        gpu._check(lib.c2gpu_nci_cutoffs_dev(n, C.c_void_p(crho.data_ptr() + off), C.c_void_p(cgrad.data_ptr() + off),
                                             C.c_void_p(rhoat.data_ptr() + off), 0.2, 2.0, 0.1, gpu._d(sr), 0, RHO_VOID,
                                             gpu._d(sums), C.byref(nd), C.c_void_p(dat.data_ptr()), n, C.c_void_p(stream)))
        tot[:3] += sums
        tot[3] += nd.value

    def step(ev=None):
        if ev:
            ev[0].record()
        fld.nci_dev_part(x0, xmat, ns, crho, cgrad, rank, world, rhoat=rhoat, promol=pro, stream=stream)
        if ev:
            ev[1].record()
        tot = np.zeros(4)
        for c0, c1 in layers:        # cutoff pass on the layers this rank owns (contiguous: x is the slowest index)
            cutoffs(c0, c1, tot)
        t = torch.from_numpy(tot).to(dev)
        if world > 1:
            sharding.reduce_void_sums(t, dist)                   # sumq, sumqp, sumv (+ .dat count)
        fld.gather_part_host(ns, [crho, cgrad, rhoat], hptr, rank, world, rows=False, nci_order=True, stream=stream)
        return t

    def sync():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    for _ in range(warmup):
        step()
    sync()
    times, kms = [], []
    for _ in range(steps):
        ev = (torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
        sync()
        t0 = time.perf_counter()
        t = step(ev)
        sync()
        times.append(time.perf_counter() - t0)
        kms.append(ev[0].elapsed_time(ev[1]))   # evaluation kernels (wavefunction + the ~1 % promolecular rhoat) of the step
    tm = torch.tensor([float(np.mean(times)), float(np.mean(kms))], dtype=torch.float64, device=dev)
    # algorithmic work of this rank's layers (x-layers of the NCI lattice = z-layers of the transposed x-fastest lattice)
    xmat_t, ns_t = xmat[:, ::-1].copy(), ns[::-1].copy()
    nact = sum(fld.count_active(x0, xmat_t, ns_t, iz0=c0, iz1=c1) for c0, c1 in layers)
    mypts = sum(c1 - c0 for c0, c1 in layers) * plane
    flops = torch.tensor([20.0 * nact * f.nmoocc + 80.0 * nact + (60.0 * f.nmoocc + 150.0) * mypts, nact, mypts],
                         dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tm, op=dist.ReduceOp.MAX)
        dist.all_reduce(flops)
    # ---- parity sample: one x-plane owned by rank 0, every 4th (j, k) -> the CPU oracle through bench.py --------------
    parity, cpu_base = None, None
    if rank == 0 and cpu is not None:
        nl = (n1 + lh - 1) // lh
        L = max(0, (nl // 2) // world * world)
        i = min(n1 - 1, L * lh + lh // 2)
        fld.nci_dev(x0, xmat, ns, crho, cgrad, rhoat=rhoat, promol=pro, i0=L * lh, i1=min(n1, (L + 1) * lh), stream=stream)
        torch.cuda.synchronize()
        sub = 4
        jj, kk = np.arange(sub // 2, n2, sub), np.arange(sub // 2, n3, sub)
        idx = torch.from_numpy((i * plane + jj[:, None] * n3 + kk[None, :]).reshape(-1)).to(dev)
        g = [a_.index_select(0, idx).contiguous() for a_ in (crho, cgrad, rhoat)]
        nsub = idx.numel()
        sums = np.zeros(3)
        nd = C.c_long(0)
        gm = g[1].clone()
        gpu._check(lib.c2gpu_nci_cutoffs_dev(nsub, C.c_void_p(g[0].data_ptr()), C.c_void_p(gm.data_ptr()),
                                             C.c_void_p(g[2].data_ptr()), 0.2, 2.0, 0.1, gpu._d(sr), 0, RHO_VOID,
                                             gpu._d(sums), C.byref(nd), C.c_void_p(dat.data_ptr()), nsub, C.c_void_p(stream)))
        torch.cuda.synchronize()
        # the sampled (j, k) sub-lattice of the plane as a lattice of its own, (j, k) on the two SLOWEST axes: same flat order
        # (k fastest), and the oracle's OpenMP loop over the slowest index gets len(jj) iterations to spread over the cores
        xs = x0 + i * xmat[:, 0] + (sub // 2) * xmat[:, 1] + (sub // 2) * xmat[:, 2]
        xm = np.stack([sub * xmat[:, 1], sub * xmat[:, 2], xmat[:, 0]], 1)
        parity, cpu_base = cpu.nci_parity(f, st, xs, xm, np.array([len(jj), len(kk), 1], np.int32),
                                          [t_.cpu().numpy() for t_ in g], gm.cpu().numpy(), sums, int(nd.value), RHO_VOID,
                                          f"x-plane {i} of {n1}, every {sub}th (j,k): {nsub} points")
    out = None
    if rank == 0:
        omega_n = st.omega / npts
        tv = t.cpu().numpy()
        sec, kernel_ms = float(tm[0].item()), float(tm[1].item())
        fl = flops.cpu().numpy()
        ach = fl[0] / world / (kernel_ms * 1e-3) / 1e12       # per GPU: every rank runs its share in kernel_ms (max)
        hv = np.frombuffer(host, dtype=np.float64)[:npts] if isinstance(host, np.ndarray) else host[:npts]
        out = {
            "cfg": 5,
            "metric": "grid points/sec (wfn NCIPLOT: RDG + sign(lambda2)*rho, promolecular rhoat, cutoff pass, gather to host)",
            "value": npts / sec, "points_per_s": npts / sec, "unit": "points/s", "n_gpus": world, "steps": steps,
            "warmup": warmup, "ms_per_step": sec * 1e3, "scaling": "strong", "dtype": "f64", "data": "synthetic",
            "config": {"workload": ("cfg4 molecule (functional check)" if small else
                                    "cfg5: synthetic C640N80O80H1200 (seed 5), 50400 primitives, 3120 MOs") +
                                   f", NCIPLOT {grid}^3 with VOID {RHO_VOID}",
                       "parallelism": f"x-layers of {lh} planes round-robin over {world} GPU(s); NCCL all-reduce of the void "
                                      "sums; every rank copies its layers of crho/cgrad/rhoat into one shared page-locked "
                                      "host buffer (no inter-GPU traffic)",
                       "kernel": fld.config()},
            "roofline": {"bound": "tensor", "achieved": ach, "peak": FP64_DMMA_TFLOPS, "unit": "TFLOP/s",
                         "frac": ach / FP64_DMMA_TFLOPS, "kernel_ms": kernel_ms, "mean_n_act": float(fl[1] / fl[2]),
                         "note": "evaluation kernels of the step (wavefunction; rhoat adds ~1 %), per GPU, max over ranks"},
            "d2h_bytes_per_step": 3 * npts * 8,
            "void": {"sumq": tv[0] * omega_n, "sumqp": tv[1] * omega_n, "sumv": tv[2] * omega_n, "ndat": int(tv[3])},
            "checksum_crho_host": float(np.abs(hv).sum()),
            "parity": parity, "cpu_baseline": cpu_base}
    fld.close()
    pro.close()
    del crho, cgrad, rhoat, dat
    torch.cuda.empty_cache()
    return out


def main():
    import torch.distributed as dist
    ap = argparse.ArgumentParser()
    ap.add_argument("--grid", type=int, default=512)
    ap.add_argument("--steps", type=int, default=2)
    ap.add_argument("--warmup", type=int, default=1)
    ap.add_argument("--small", action="store_true", help="config-4 molecule instead (quick functional check)")
    a = ap.parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    shared = None
    if world > 1:
        os.environ.setdefault("NCCL_DEBUG", "WARN")
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")   # keep stdout for the JSON line
        dist.init_process_group("nccl", device_id=dev)
        shared = sharding.SharedPinned(rank, world, dist)
    gpu.init_devices([local])
    out = run(rank, world, dev, dist if world > 1 else None, a.grid, a.steps, a.warmup, a.small,
              shared_host=shared.get if shared else None)
    if rank == 0:
        print(json.dumps(out), flush=True)
    if shared:
        shared.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
