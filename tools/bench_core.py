"""Cost of the core augmentation (f%usecore, fieldmod@proc.f90:827-832) on the bench molecule: the cfg-4 wavefunction
on an n^3 lattice with and without a core field (C zpsp 4, N 5, O 6) attached.  Prints one JSON line."""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import cases  # noqa: E402  (golden core tables only)
from critic2_b200 import gpu, structure, synth  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 200
gpu.init_devices([0])
f = synth.config4()
st = structure.molecular_cell(f.xyz, f.z)
lo, hi = st.r.min(0) - 6.0, st.r.max(0) + 6.0
nn = np.array([n, n, n], np.int32)
xd = np.diag((hi - lo) / nn)
props = gpu.PROPS[:4] + gpu.PROPS[5:11]
outs = [torch.zeros(n ** 3, dtype=torch.float64, device="cuda") for _ in props]
fld = gpu.WfnField(f, st)
core = gpu.core_field(st, cases.core_tables(), {6: 4, 7: 5, 8: 6})
s = torch.cuda.current_stream().cuda_stream
res = {}
for name, c in (("plain", None), ("core", core), ("plain2", None)):
    fld.set_core(c)
    ms = []
    for rep in range(3):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fld.cube_dev(lo, xd, nn, outs, props, stream=s); e1.record(); torch.cuda.synchronize()
        ms.append(e0.elapsed_time(e1))
    res[name] = min(ms)
    res[name + "_checksum"] = float(outs[0].sum())
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
raw = [torch.zeros(n ** 3, dtype=torch.float64, device="cuda") for _ in props]
e0.record(); core.cube_dev(lo, xd, nn, raw, props, stream=s); e1.record(); torch.cuda.synchronize()
res["core_field_alone_ms"] = e0.elapsed_time(e1)
res["grid"] = n
print(json.dumps(res))
