"""BASELINE.json configs[0]: benchmarks-cores/test-tiny.cfg via dynearthsol2d -- the reference's
own CPU-runnable plumbing case (2D triangles, OpenMP).  The 2D element kernels are outside the
hot-path scope, so nothing of ours computes here: the test pins
that the reference builds from its sources with oracle/Makefile's recipe (no -DTHREED, no
tetgen), runs the named config, is bit-identical across OpenMP thread counts (what makes it a
usable oracle) and passes its own benchmarks-cores/compare.py against itself."""
import os
import shutil
import subprocess
import sys

import pytest

import desframe
from conftest import ROOT, HERE

REF2D = os.path.join(ROOT, "oracle", "_ref", "dynearthsol2d.ref")
COMPARE = "/root/reference/benchmarks-cores/compare.py"


def write_poly(path, xlength=150e3, zlength=100e3):
    """The planar straight-line graph test-tiny.cfg meshes (mesh.poly_filename = cube.poly): a
    rectangle of one region, its four sides flagged x0 / z0 / x1 / z1 (triangle's .poly layout:
    points, segments with boundary flags, no holes, one region with material 0 and relative
    size 1)."""
    pts = [(0.0, 0.0), (0.0, -zlength), (xlength, -zlength), (xlength, 0.0)]
    flags = [1, 16, 2, 32]      # BOUNDX0, BOUNDZ0, BOUNDX1, BOUNDZ1 (constants.hpp:27-37)
    with open(path, "w") as f:
        f.write("4 2 0 0\n")
        for i, (x, z) in enumerate(pts):
            f.write("%d %r %r\n" % (i, x, z))
        f.write("4 1\n")
        for j in range(4):
            f.write("%d %d %d %d\n" % (j, j, (j + 1) % 4, flags[j]))
        f.write("0\n1\n0 %r %r 0 1.0\n" % (xlength / 2, -zlength / 2))


def run(wd, threads):
    os.makedirs(wd, exist_ok=True)
    shutil.copy(os.path.join(HERE, "cfg", "tiny2d.cfg"), wd)
    write_poly(os.path.join(wd, "cube.poly"))
    env = dict(os.environ, OMP_NUM_THREADS=str(threads))
    r = subprocess.run([REF2D, "tiny2d.cfg"], cwd=wd, env=env, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, (r.stdout[-1000:], r.stderr[-1000:])
    return r.stdout


def test_test_tiny_cfg_runs_on_the_2d_reference(tmp_path):
    if not os.path.exists(REF2D):
        pytest.skip("oracle/_ref/dynearthsol2d.ref not built (needs /root/reference at build time)")
    out = run(str(tmp_path / "t1"), 1)
    assert "Output # 4, step = 4" in out and "Ending simulation." in out
    run(str(tmp_path / "t4"), 4)
    for k in range(5):
        a = desframe.read_frame(str(tmp_path / "t1" / ("tiny2d.save.%06d" % k)))
        b = desframe.read_frame(str(tmp_path / "t4" / ("tiny2d.save.%06d" % k)))
        assert a["connectivity"].size % 3 == 0 and a["coordinate"].size == 2 * a["temperature"].size   # 2D
        worst, report, bad = desframe.compare_frames(a, b)
        assert not bad and worst == 0.0, (k, bad, report)
    if os.path.exists(COMPARE):
        env = dict(os.environ, PYTHONPATH="/root/reference")
        r = subprocess.run([sys.executable, COMPARE, str(tmp_path / "t1" / "tiny2d"), str(tmp_path / "t4" / "tiny2d"), "4"],
                           cwd=str(tmp_path), env=env, capture_output=True, text=True)
        assert r.returncode == 0, (r.stdout, r.stderr)
        assert "BIT-EXACT" in r.stdout, r.stdout
