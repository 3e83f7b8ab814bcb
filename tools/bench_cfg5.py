"""BASELINE.json configs[4]: synthetic 2000-atom wavefunction NCIPLOT + promolecular rhoat on 512^3, x-layer-sharded over
the GPUs of one box (one process per GPU under torchrun; also runs on 1 GPU).

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29512 \
        tools/bench_cfg5.py --steps 2 --warmup 1

A step = the whole NCIPLOT data path of nci@proc.f90:499-604 on the device: crho/cgrad (wavefunction, nder = 2 + eig) and
rhoat (promolecular) on this rank's interleaved x-layers, the cutoff pass with VOID on the same layers, then the only
communication the path has: NCCL all-reduce of the void sums / .dat count and the gather of the layers to
the writer rank (sum-reduce of disjoint layers).  One JSON line from rank 0; timing = CUDA events, max over ranks."""
import argparse
import ctypes as C
import json
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from critic2_b200 import drivers, gpu, sharding, structure, synth  # noqa: E402
import cases  # noqa: E402

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
if world > 1:
    os.environ["NCCL_DEBUG"] = "WARN"
    os.environ["NCCL_DEBUG_FILE"] = "/dev/stderr"   # the version banner is part of the log: keep stdout for the JSON line
    dist.init_process_group("nccl", device_id=dev)
gpu.init_devices([local])

f = synth.config4() if a.small else synth.config5()
st = structure.molecular_cell(f.xyz, f.z)
fld = gpu.WfnField(f, st)
pro = gpu.PromolField(st, cases.tables())
x0, xmat, ns = drivers.nci_grid(st, nstep=[a.grid] * 3)
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
RHO_VOID = 0.05


def step():
    fld.nci_dev_part(x0, xmat, ns, crho, cgrad, rank, world, rhoat=rhoat, promol=pro, stream=stream)
    tot = np.zeros(4)
    for c0, c1 in layers:        # cutoff pass on the layers this rank owns (contiguous: x is the slowest index)
        n = (c1 - c0) * plane
        off = c0 * plane * 8
        sums = np.zeros(3)
        nd = C.c_long(0)
        gpu._check(lib.c2gpu_nci_cutoffs_dev(n, C.c_void_p(crho.data_ptr() + off), C.c_void_p(cgrad.data_ptr() + off),
                                             C.c_void_p(rhoat.data_ptr() + off), 0.2, 2.0, 0.1, gpu._d(sr), 0, RHO_VOID,
                                             gpu._d(sums), C.byref(nd), C.c_void_p(dat.data_ptr()), n, C.c_void_p(stream)))
        tot[:3] += sums
        tot[3] += nd.value
    t = torch.from_numpy(tot).to(dev)
    if world > 1:
        sharding.reduce_void_sums(t, dist)                       # sumq, sumqp, sumv (+ .dat count)
        for arr in (crho, cgrad, rhoat):
            sharding.gather_layers(arr, n1, lh, rank, world, dist, dst=0)
    return t


def restore():
    """The gather adds the other ranks' layers into rank 0's arrays: zero them again so every step starts alike."""
    if world > 1 and rank == 0:
        for arr in (crho, cgrad, rhoat):
            sharding.zero_foreign_layers(arr, n1, lh, rank, world)


def sync():
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
        torch.cuda.synchronize()


for _ in range(a.warmup):
    step()
    restore()
sync()
ms = []
for i in range(a.steps):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    sync()
    e0.record()
    t = step()
    e1.record()
    sync()
    ms.append(e0.elapsed_time(e1))
    if i + 1 < a.steps:
        restore()
tm = torch.tensor([float(np.mean(ms))], dtype=torch.float64, device=dev)
if world > 1:
    dist.all_reduce(tm, op=dist.ReduceOp.MAX)
if rank == 0:
    omega_n = st.omega / npts
    tv = t.cpu().numpy()
    print(json.dumps({
        "metric": "grid points/sec (wfn NCIPLOT: RDG + sign(lambda2)*rho, promolecular rhoat, cutoff pass, gather)",
        "value": npts / (float(tm.item()) * 1e-3), "unit": "points/s", "n_gpus": world, "steps": a.steps, "warmup": a.warmup,
        "ms_per_step": float(tm.item()), "scaling": "strong", "dtype": "f64", "data": "synthetic",
        "config": {"workload": ("cfg4 molecule (functional check)" if a.small else
                                "cfg5: synthetic C640N80O80H1200 (seed 5), 50400 primitives, 3120 MOs") +
                               f", NCIPLOT {a.grid}^3 with VOID {RHO_VOID}",
                   "parallelism": f"x-layers of {lh} planes round-robin over {world} GPU(s); NCCL all-reduce of the void sums, "
                                  "sum-reduce gather of crho/cgrad/rhoat to rank 0",
                   "kernel": fld.config()},
        "void": {"sumq": tv[0] * omega_n, "sumqp": tv[1] * omega_n, "sumv": tv[2] * omega_n, "ndat": int(tv[3])},
        "checksum_crho_rank0": float(crho.sum().item())}), flush=True)
if world > 1:
    dist.destroy_process_group()
