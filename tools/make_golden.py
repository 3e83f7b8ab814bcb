"""Generate the committed fixtures under tests/golden/ (run in the authoring container, where /root/reference exists).

* promol_tables.npz : radial promolecular tables for H, C, N, O, Si built by the oracle's read_critic restatement
  (grid1mod@proc.f90:204-332) from /root/reference/dat/wfc/<el>_pbe.wfc.  The .wfc inputs cannot travel to the GPU box.
* fchk_*.npz : the flat molwfn arrays produced by critic2_b200.wfn.read_fchk for the tests_cint inputs of BASELINE
  configs 1 and 2 (h2o-aqz, benzene-adz) plus a Cartesian-d/f case (hooh-atz) and the shipped h2o.wfx.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import oracle  # noqa: E402
from critic2_b200 import wfn  # noqa: E402

REF = "/root/reference"
OUT = os.path.join(ROOT, "tests", "golden")
os.makedirs(OUT, exist_ok=True)

tabs = {}
for z, el in ((1, "h_"), (6, "c_"), (7, "n_"), (8, "o_"), (14, "si")):
    t = oracle.read_critic(f"{REF}/dat/wfc/{el}_pbe.wfc", z, 0)
    for k, v in t.items():
        tabs[f"{z}_{k}"] = np.asarray(v)
    print(el, "ngrid", len(t["r"]), "rmax", t["r"][-1], "a", t["a"], "b", t["b"])
np.savez_compressed(os.path.join(OUT, "promol_tables.npz"), **tabs)

# core tables cgrid(Z, zpsp) (grid1_register_core -> read_db(Z, q), grid1mod@proc.f90:139-164,57-80): the usual
# pseudopotential valences C 4, N 5, O 6, Si 4
ctabs = {}
for z, q, el in ((6, 4, "c_"), (7, 5, "n_"), (8, 6, "o_"), (14, 4, "si")):
    t = oracle.read_critic(f"{REF}/dat/wfc/{el}_pbe.wfc", z, q)
    for k, v in t.items():
        ctabs[f"{z}_{q}_{k}"] = np.asarray(v)
    print("core", el, q, "ngrid", len(t["r"]), "rmax", t["r"][-1])
np.savez_compressed(os.path.join(OUT, "core_tables.npz"), **ctabs)

for name, path, rd in (("h2o-aqz", f"{REF}/tests_cint/h2o-aqz.fchk", wfn.read_fchk),
                       ("benzene-adz", f"{REF}/tests_cint/benzene-adz.fchk", wfn.read_fchk),
                       ("hooh-atz", f"{REF}/tests_cint/hooh-atz.fchk", wfn.read_fchk),
                       ("h2o-wfx", f"{REF}/tests_libxc/h2o.wfx", wfn.read_wfx)):
    f = rd(path)
    np.savez_compressed(os.path.join(OUT, f"wfn_{name}.npz"), xyz=f.xyz, z=f.z, icenter=f.icenter, itype=f.itype, e=f.e,
                        occ=f.occ, cmo=f.cmo, nmoocc=f.nmoocc, wfntyp=f.wfntyp, nalpha=f.nalpha)
    print(name, "npri", f.npri, "nmo", f.nmoocc)
