"""Development probe: block-synchronous M/L against the warp-specialised kernels on mid-size MO counts."""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from critic2_b200 import gpu, structure, synth
gpu.init_devices([0])
n = 160
import os as _os
for nmo in [int(v) for v in _os.environ.get('PROBE_NMO', '48,150,200').split(',')]:
    f = synth.synthetic_molecule(lattice=(4, 4, 4), ndrop=4, composition=((6, 24), (7, 3), (8, 3), (1, 30)), seed=7, nmo=nmo) \
        if "nmo" in synth.synthetic_molecule.__code__.co_varnames else synth.small_synthetic(nmo=nmo)
    st = structure.molecular_cell(f.xyz, f.z)
    lo, hi = st.r.min(0) - 5.0, st.r.max(0) + 5.0
    nn = np.array([n] * 3, np.int32); xd = np.diag((hi - lo) / nn)
    props = gpu.PROPS[:4] + gpu.PROPS[5:11]
    outs = [torch.zeros(n ** 3, dtype=torch.float64, device="cuda") for _ in props]
    s = torch.cuda.current_stream().cuda_stream
    ref = None
    for cfg in [int(v) for v in _os.environ.get('PROBE_CFG', '2,3,6').split(',')]:
        gpu.set_option("wfn_cfg", cfg)
        fld = gpu.WfnField(f, st)
        nact = fld.count_active(lo, xd, nn)
        flops = 20.0 * nact * f.nmoocc + 80.0 * nact + 60.0 * f.nmoocc * n ** 3
        for rep in range(3):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); fld.cube_dev(lo, xd, nn, outs, props, stream=s); e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        r = outs[0].cpu().numpy()
        if ref is None: ref = r
        print(f"npri {f.npri} nmo {nmo} cfg {cfg} nblk {fld.config()['nblk']} {ms:8.2f} ms {flops/ms/1e9:6.2f} TFLOP/s ({flops/ms/1e9/37.1*100:4.1f}%) "
              f"maxrel {float(np.abs(r-ref).max()/np.abs(ref).max()):.1e}", flush=True)
        fld.close()
gpu.set_option("wfn_cfg", -1)
