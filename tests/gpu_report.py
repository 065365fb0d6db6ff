"""Prints, per case and field, whether the CUDA path is bit-exact with the reference or the
relative error (max|a-b|/max|b|).  Run on the GPU box; output is committed under profiles/."""
import sys, os
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import refmodel
from conftest import ref_dump
from test_parity_gpu import FIELDS, CASES


def line(tag, got, d, prefix):
    parts = []
    for k in FIELDS:
        if prefix + k not in d:
            continue
        if np.array_equal(got[k], d[prefix + k]):
            parts.append("%s=EXACT" % k)
        else:
            nd = int((got[k] != d[prefix + k]).sum())
            parts.append("%s=%.1e(%d/%d differ)" % (k, refmodel.rel_err(got[k], d[prefix + k]), nd, got[k].size))
    print("%-28s %s" % (tag, " ".join(parts)))


for case in CASES:
    d = ref_dump(case, [1, 20])
    g = refmodel.model_from_dump(d)
    dt = float(d["s0_dt"])
    g.step(dt, dt, 1)
    line(case + " step1", g.download_state(*FIELDS), d, "s1_")
    dt2, t, s = g.run(19, dt, dt, 1)
    line(case + " step20", g.download_state(*FIELDS), d, "s20_")
    print("   dt after 20 steps: gpu %.17g ref %.17g" % (dt2, float(d["s20_dt"])))
    g.close()

d = ref_dump("box-ep", [250, 251, 270])
g = refmodel.model_from_dump(d, prefix="s250_")
dt = float(d["s250_dt"]); time = float(d["s250_time"])
g.step(dt, time + dt, 251)
print("yielding elements at step 251 (reference):", int((d["s251_delta_plstrain"] > 0).sum()))
line("box-ep yielding step251", g.download_state(*FIELDS), d, "s251_")
g.run(19, dt, time + dt, 251)
line("box-ep yielding step270", g.download_state(*FIELDS), d, "s270_")
g.close()
