"""Quick device-resident timing probe of the wavefunction kernel on config-4-like inputs (development tool)."""
import argparse
import sys
import time
import os

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from critic2_b200 import gpu, structure, synth  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--n", type=int, default=400)
ap.add_argument("--z0", type=int, default=192)
ap.add_argument("--z1", type=int, default=208)
ap.add_argument("--cfgs", type=str, default="3,4")
ap.add_argument("--small", action="store_true")
ap.add_argument("--nmo", type=int, default=0, help="MO count of the config-4 molecule (default 780)")
a = ap.parse_args()

gpu.init_devices([0])
f = synth.small_synthetic() if a.small else (synth.synthetic_molecule(nmo=a.nmo) if a.nmo else synth.config4())
st = structure.molecular_cell(f.xyz, f.z)
lo, hi = st.r.min(0) - 6.0, st.r.max(0) + 6.0
nn = np.array([a.n] * 3, np.int32)
xd = np.diag((hi - lo) / nn)
props = gpu.PROPS[:4] + gpu.PROPS[5:11]
npl = a.n * a.n
for cfg in [int(c) for c in a.cfgs.split(",")]:
    gpu.set_option("wfn_cfg", cfg)
    fld = gpu.WfnField(f, st)
    outs = [torch.zeros(a.n ** 3, dtype=torch.float64, device="cuda") for _ in props]
    nact = fld.count_active(lo, xd, nn, iz0=a.z0, iz1=a.z1)
    npts = npl * (a.z1 - a.z0)
    flops = 20.0 * nact * f.nmoocc + 80.0 * nact + 60.0 * f.nmoocc * npts
    s = torch.cuda.current_stream().cuda_stream
    for rep in range(2):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fld.cube_dev(lo, xd, nn, outs, props, iz0=a.z0, iz1=a.z1, stream=s)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        c_ = fld.config()
        print(f"nmo {f.nmoocc} padded {c_['nblk'] * c_['mo_block']} ms/column {ms / (c_['nblk'] * c_['mo_block']):.4f} | "
              f"cfg {cfg} {fld.config()} slab [{a.z0},{a.z1}) {npts} pts mean n_act {nact / npts:.1f} "
              f"{ms:.1f} ms -> {npts / ms * 1e3:.3e} pts/s, {flops / ms / 1e9:.2f} TFLOP/s algorithmic "
              f"({flops / ms / 1e9 / 37.1 * 100:.1f}% of 37.1)", flush=True)
    fld.close()
