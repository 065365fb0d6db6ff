"""Where do yielding elements differ from the reference? (diagnostic, GPU)"""
import sys
import numpy as np
sys.path.insert(0, "tests")
from conftest import ref_dump, ref_branches
import refmodel
import test_branches_gpu as tb

def eigvals(s6):
    n = s6.shape[1]
    m = np.zeros((n, 3, 3))
    m[:, 0, 0], m[:, 1, 1], m[:, 2, 2] = s6[0], s6[1], s6[2]
    m[:, 0, 1] = m[:, 1, 0] = s6[3]; m[:, 0, 2] = m[:, 2, 0] = s6[4]; m[:, 1, 2] = m[:, 2, 1] = s6[5]
    return np.linalg.eigvalsh(m)

d = ref_branches("ql")
g = refmodel.model_from_dump(d)
up = {k: d["ql_in_" + k] for k in ("stress", "strain", "strain_rate", "plstrain", "edvoldt", "volume", "volume_old", "temperature")}
g.upload_state(**up)
g.call("update_stress", float(d["s0_dt"]))
got = g.download_state("stress", "delta_plstrain", "plstrain")
E = d["ql_kinds"].size
a = got["stress"].reshape(6, E); b = d["ql_out_stress"].reshape(6, E)
scale = np.abs(b).max()
err = np.abs(a - b).max(axis=0) / scale
ea, eb = eigvals(a), eigvals(b)
eerr = np.abs(ea - eb).max(axis=1) / scale
yl = d["ql_out_delta_plstrain"] > 0
for k in range(6):
    m = d["ql_kinds"] == k
    print("QL kind", k, "n", m.sum(), "yield", (yl & m).sum(), "max comp err %.2e" % err[m].max(), "max eigenvalue err %.2e" % eerr[m].max(),
          "n(err>1e-10)", (err[m] > 1e-10).sum(), "dpl err %.2e" % (np.abs(got["delta_plstrain"] - d["ql_out_delta_plstrain"])[m].max()))
w = np.argsort(err)[-3:]
for e in w:
    print(" elem", e, "kind", d["ql_kinds"][e], "in", d["ql_in_stress"].reshape(6, E)[:, e], "\n   ref", b[:, e], "\n   gpu", a[:, e], "\n   eig ref", eb[e], "gpu", ea[e])
g.close()

d = ref_dump("box-ep", [1, 20, 250, 251, 270], over=tb.YIELDING["two-materials"])
g = refmodel.model_from_dump(d, prefix="s250_")
dt, time = float(d["s250_dt"]), float(d["s250_time"])
g.step(dt, time + dt, 251)
got = g.download_state("stress", "delta_plstrain", "plstrain")
E = int(d["p_nelem"])
a = got["stress"].reshape(6, E); b = d["s251_stress"].reshape(6, E)
scale = np.abs(b).max()
err = np.abs(a - b).max(axis=0) / scale
em = d["elemmarkers"].reshape(E, 2)
yl = d["s251_delta_plstrain"] > 0
print("two-materials s251: yielding", yl.sum(), "max err yield %.2e" % err[yl].max(), "non-yield %.2e" % err[~yl].max())
for name, m in (("mat0 only", em[:, 1] == 0), ("mat1 only", em[:, 0] == 0), ("mixed", (em > 0).all(axis=1))):
    print("  ", name, m.sum(), "yield", (yl & m).sum(), "max err %.2e" % err[m].max(), "n>1e-10", (err[m] > 1e-10).sum())
ea, eb = eigvals(a), eigvals(b)
print("   eigenvalue err %.2e" % (np.abs(ea - eb).max() / scale))
w = np.argsort(err)[-3:]
for e in w:
    print(" elem", e, "markers", em[e], "pls", d["s250_plstrain"][e], "dpl ref", d["s251_delta_plstrain"][e], "gpu", got["delta_plstrain"][e],
          "\n   s250", d["s250_stress"].reshape(6, E)[:, e], "\n   ref", b[:, e], "\n   gpu", a[:, e], "\n   eig ref", eb[e], "gpu", ea[e])
