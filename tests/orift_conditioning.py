"""How well-conditioned is examples/oblique-rift-3d.cfg at its configured length?  (CPU only.)

The reference library runs tests/cfg/orift-full.cfg twice: as it is, and with every stress
entry of the INITIAL state moved by one ulp (np.nextafter).  Both runs are the reference's own
code on the same machine; the relative difference between them (compare.py's measure,
max|a-b|/max|a|) at the two output times is what ANY implementation that is not bit-identical
in libm (the EVP viscosity calls pow/exp; the Mohr-Coulomb return mapping goes through
atan2/cos/sin in 3x3-C, built -ffast-math) will show against the reference at best.

    python tests/orift_conditioning.py out.json            # ~6 min of CPU
"""
import json
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
FIELDS = ["coord", "vel", "temperature", "stress", "strain", "strain_rate", "plstrain", "viscosity"]
STEPS = [300, 19000, 38000]

WORKER = r"""
import sys, numpy as np
sys.path.insert(0, %r)
from refharness import Ref
perturb, out = int(sys.argv[1]), sys.argv[2]
r = Ref(%r, workdir=sys.argv[3])
if perturb:
    s = r.get("stress")
    r.set("stress", np.nextafter(s, np.inf))
save, done = {}, 0
for n in %r:
    r.step(n - done); done = n
    for f in %r:
        save["s%%d_%%s" %% (n, f)] = r.get(f)
np.savez(out, **save)
""" % (HERE, os.path.join(HERE, "cfg", "orift-full.cfg"), STEPS, FIELDS)


def main():
    import tempfile
    import numpy as np
    out = sys.argv[1]
    wd = tempfile.mkdtemp(prefix="orift_cond_")
    procs = []
    for p in (0, 1):
        d = os.path.join(wd, str(p)); os.makedirs(d)
        env = dict(os.environ, OMP_NUM_THREADS=str(max(1, (os.cpu_count() or 2) // 2)))
        procs.append(subprocess.Popen([sys.executable, "-c", WORKER, str(p), os.path.join(wd, "r%d.npz" % p), d],
                                      env=env, stdout=subprocess.DEVNULL))
    for p in procs:
        assert p.wait() == 0
    a, b = np.load(os.path.join(wd, "r0.npz")), np.load(os.path.join(wd, "r1.npz"))
    res = {"config": "tests/cfg/orift-full.cfg (examples/oblique-rift-3d.cfg, configured length)",
           "perturbation": "every entry of the initial stress moved by one ulp", "measure": "max|a-b| / max|a| (compare.py)",
           "steps": {}}
    for n in STEPS:
        res["steps"][str(n)] = {}
        for f in FIELDS:
            x, y = a["s%d_%s" % (n, f)], b["s%d_%s" % (n, f)]
            m = float(np.abs(x).max())
            res["steps"][str(n)][f] = float(np.abs(x - y).max() / (m if m > 0 else 1.0))
    json.dump(res, open(out, "w"), indent=1)
    print(json.dumps(res, indent=1))


if __name__ == "__main__":
    main()
