"""Per-step field differences of one cfg variant against the reference (diagnostic, GPU)."""
import sys
import numpy as np
sys.path.insert(0, "tests")
from conftest import ref_dump
import refmodel
import test_branches_gpu as tb
from test_parity_gpu import FIELDS
name = sys.argv[1]
steps = [int(x) for x in sys.argv[2].split(",")]
d = ref_dump("box-ep", steps, over=tb.VARIANTS[name][0])
g = refmodel.model_from_dump(d)
dt = float(d["s0_dt"]); t = 0.0; s = 0
for k in steps:
    dt, t, s = g.run(k - s, dt, t, s)
    got = g.download_state(*FIELDS)
    rep = []
    for f in FIELDS:
        e = refmodel.rel_err(got[f], d["s%d_%s" % (k, f)])
        if e: rep.append("%s:%.1e" % (f, e))
    print("step", k, "dt", dt, float(d["s%d_dt" % k]), "time", t, float(d["s%d_time" % k]), " ".join(rep) or "all bit-exact")
