"""The drop-in itself: bridge/_build/dynearthsol3d.b200 is the reference's own object code
(input, init, tetgen meshing, markers, remeshing, output) with the step-loop call sites swapped
for the C-ABI (bridge/des_driver.cxx).  It and the unmodified reference binary
run the same .cfg on the same box; every frame they write is compared with the measure and the
threshold of benchmarks-cores/compare.py (max|new-old|/max|old| < 1e-8 per field)."""
import os
import subprocess

import pytest

import desframe
from conftest import ROOT, HERE

pytestmark = pytest.mark.gpu

REF = os.path.join(ROOT, "oracle", "_ref", "dynearthsol3d.ref")
B200 = os.path.join(ROOT, "bridge", "_build", "dynearthsol3d.b200")


def run(binary, cfg, wd):
    os.makedirs(wd, exist_ok=True)
    env = dict(os.environ, OMP_NUM_THREADS=str(min(16, os.cpu_count() or 1)))
    r = subprocess.run([binary, os.path.join(HERE, "cfg", cfg)], cwd=wd, env=env, capture_output=True,
                       text=True, timeout=1500)
    assert r.returncode == 0, (binary, r.stdout[-1500:], r.stderr[-1500:])
    return r.stdout


@pytest.mark.parametrize("cfg,model,frames", [
    ("t3d.cfg", "t3d", 6),                  # test-3d.cfg: 519 steps, tetgen mesh, Winkler, surface diffusion
    ("t3d-remesh.cfg", "t3dr", 5),          # test-3d-remesh.cfg: host remesh at step 300 + re-upload
    ("t3d-iso.cfg", "t3di", 2),             # + isostasy spin-up before the step loop (51 twin-call iterations)
    ("orift.cfg", "orift", 4),              # oblique-rift-3d.cfg (300 steps): EVP, 2 materials, PREM, averaged fields
])
def test_driver_matches_reference_frames(cfg, model, frames, tmp_path):
    if not (os.path.exists(REF) and os.path.exists(B200)):
        pytest.skip("reference / bridge binaries not built (need /root/reference at build time)")
    if not os.path.exists(os.path.join(HERE, "cfg", cfg)):
        pytest.skip("no such cfg")
    run(REF, cfg, str(tmp_path / "ref"))
    out = run(B200, cfg, str(tmp_path / "gpu"))
    worst_all = 0.0
    for k in range(frames):
        fn = "%s.save.%06d" % (model, k)
        a = desframe.read_frame(str(tmp_path / "ref" / fn))
        b = desframe.read_frame(str(tmp_path / "gpu" / fn))
        worst, report, bad = desframe.compare_frames(a, b)
        assert not bad, "frame %d: %s   [all: %s]" % (k, bad, " ".join(report))
        worst_all = max(worst_all, worst)
    print("%s: %d frames, worst relative difference %.2e" % (cfg, frames, worst_all))
