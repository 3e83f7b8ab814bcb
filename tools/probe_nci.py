"""Development probe: benzene NCIPLOT 200^3 (BASELINE config 2) through forced wavefunction kernel configurations."""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, 'tests'))
import cases
from critic2_b200 import gpu, drivers
gpu.init_devices([0])
gpu.set_option("nci_two_pass", int(os.environ.get("NCI_TWO_PASS", "1")))
f = cases.golden_wfn("benzene-adz"); st = cases.mol_structure(f)
x0, xmat, ns = drivers.nci_grid(st, nstep=(200, 200, 200))
n = 200 ** 3
crho = torch.zeros(n, dtype=torch.float64, device="cuda"); cgrad = torch.zeros_like(crho)
s = torch.cuda.current_stream().cuda_stream
ref = None
for cfg in [int(v) for v in os.environ.get("PROBE_CFG", "7,1").split(",")]:
    gpu.set_option("wfn_cfg", cfg)
    fld = gpu.WfnField(f, st)
    for rep in range(4):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fld.nci_dev(x0, xmat, ns, crho, cgrad, stream=s); e1.record(); torch.cuda.synchronize()
    r = crho.cpu().numpy()
    if ref is None: ref = r
    print(f"cfg {cfg} {fld.config()} {e0.elapsed_time(e1):8.3f} ms maxdiff {np.abs(r-ref).max():.2e}", flush=True)
    fld.close()
# the same lattice as a CUBE of the 10 raw properties (cheap point epilogue): the difference to the NCIPLOT pass is the cost
# of the eigenvalue / RDG epilogue on the critical path of the block-synchronous kernels
props = gpu.PROPS[:4] + gpu.PROPS[5:11]
outs = [torch.zeros(n, dtype=torch.float64, device="cuda") for _ in props]
gpu.set_option("wfn_cfg", -1)
fld = gpu.WfnField(f, st)
xd = np.asarray(xmat)
for rep in range(4):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); fld.cube_dev(x0, xd, ns, outs, props, stream=s); e1.record(); torch.cuda.synchronize()
print(f"CUBE 10 properties on the same lattice: {e0.elapsed_time(e1):8.3f} ms")
for rep in range(4):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); fld.cube_dev(x0, xd, ns, outs[:1], props[:1], nder=0, stream=s); e1.record(); torch.cuda.synchronize()
print(f"CUBE rho only (nder 0): {e0.elapsed_time(e1):8.3f} ms")
