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
ap.add_argument("--super", type=int, default=0, help="super-tile edge (option 'super'; development A/B)")
ap.add_argument("--cfg5", action="store_true", help="the 2000-atom molecule of BASELINE config 5")
ap.add_argument("--nmo", type=int, default=0, help="MO count of the config-4 molecule (default 780)")
a = ap.parse_args()

gpu.init_devices([0])
if a.super:
    gpu.set_option("super", a.super)
f = synth.config5() if a.cfg5 else synth.small_synthetic() if a.small else (synth.synthetic_molecule(nmo=a.nmo) if a.nmo else synth.config4())
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
    L = gpu.lib()
    if hasattr(L, "c2gpu_ws_diag_read") and cfg in (6, 12, 13, 14, 18, 19, 20):   # diagnostic build (tools/ws_diag.sh, bit 16)
        import ctypes as C
        nsm = torch.cuda.get_device_properties(0).multi_processor_count
        buf = (C.c_ulonglong * (nsm * 64))()
        L.c2gpu_ws_diag_read(buf, nsm * 64)
        d = np.frombuffer(buf, dtype=np.uint64).reshape(nsm, 8, 8).astype(np.float64)
        for name, sl in (("4-fragment warps", slice(0, 4)), ("3-fragment warps", slice(4, 8))):
            w_, m_, n_, t_, w1_, ep_, st_, ti_ = (d[:, sl, k] for k in range(8))
            print(f"  {name}: of the alive cycles: wait at a sub-tile's first stage {100 * (w1_ / t_).mean():.2f} %, other waits "
                  f"{100 * ((w_ - w1_) / t_).mean():.2f} %, block/sub-tile epilogues {100 * (ep_ / t_).mean():.2f} %, rest (stage "
                  f"hand-over) {100 * ((t_ - w_ - m_ - ep_) / t_).mean():.2f} % = {((t_ - w_ - m_ - ep_) / st_).mean():.0f} cycles per stage; "
                  f"{st_.mean():.0f} stages, {ti_.mean():.0f} sub-tiles, loop cycles per stage {(m_ / st_).mean():.0f}")
            print(f"  {name}: waiting for a full stage {100 * (w_ / t_).mean():.1f} % (by sub-partition "
                  f"{[round(100 * (w_ / t_)[:, j].mean(), 1) for j in range(4)]}), in the MMA loop {100 * (m_ / t_).mean():.1f} %, "
                  f"DMMA x 16 clk / loop cycles {100 * (16 * n_ / m_).mean():.1f} %, / alive cycles {100 * (16 * n_ / t_).mean():.1f} %")
        tot = 16 * (d[:, :4, 2] + d[:, 4:, 2]) / np.maximum(d[:, :4, 3], d[:, 4:, 3])
        print(f"  DMMA pipe cycles per sub-partition / alive cycles: {100 * tot.mean():.1f} % (by sub-partition "
              f"{[round(100 * tot[:, j].mean(), 1) for j in range(4)]})")
    if hasattr(L, "c2gpu_ws_gdiag_read") and cfg in (6, 12, 13, 14, 18, 19, 20):
        gb = (C.c_ulonglong * (nsm * 8))()
        L.c2gpu_ws_gdiag_read(gb, nsm * 8)
        gd = np.frombuffer(gb, dtype=np.uint64).reshape(nsm, 2, 4).astype(np.float64)
        print(f"  generator warps 0/1: waiting for a free cache buffer {100 * (gd[:, :, 0] / gd[:, :, 3]).mean():.1f} %, S2 + records "
              f"{100 * (gd[:, :, 1] / gd[:, :, 3]).mean():.1f} %, chi tiles {100 * (gd[:, :, 2] / gd[:, :, 3]).mean():.1f} % of their cycles")
    fld.close()
