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


def run(binary, cfg, wd, **extra_env):
    os.makedirs(wd, exist_ok=True)
    env = dict(os.environ, OMP_NUM_THREADS=str(min(16, os.cpu_count() or 1)), **extra_env)
    r = subprocess.run([binary, os.path.join(HERE, "cfg", cfg)], cwd=wd, env=env, capture_output=True,
                       text=True, timeout=1500)
    assert r.returncode == 0, (binary, r.stdout[-1500:], r.stderr[-1500:])
    return r.stdout


@pytest.mark.parametrize("cfg,model,frames", [
    ("t3d.cfg", "t3d", 6),                  # test-3d.cfg: 519 steps, tetgen mesh, Winkler, surface diffusion
    ("t3d-remesh.cfg", "t3dr", 5),          # test-3d-remesh.cfg: host remesh at step 300 + re-upload
    ("t3d-iso.cfg", "t3di", 2),             # + isostasy spin-up before the step loop (51 twin-call iterations)
    ("orift.cfg", "orift", 4),              # oblique-rift-3d.cfg (300 steps): EVP, 2 materials, PREM, averaged fields
    ("orift-full.cfg", "oriftfull", 3),     # oblique-rift-3d.cfg at its configured length (0.1 Myr, ~38 000 steps)
    ("t3d-pt.cfg", "t3dpt", 4),             # control.has_PT: the step through the twins + 3 pseudo-transient iterations
    ("t3d-pt-bfa.cfg", "t3dbf", 4),         # + initial_body_force_adjustment, prescribed tractions switched on after it
    ("t3d-nomarkers.cfg", "t3dnm", 6),      # sim.has_marker_output = no: frames written on a background thread
    ("orift-phase.cfg", "oriftpc", 4),      # phase changes on device-resident markers + hydration (8 rock types)
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
        if cfg == "orift-full.cfg" and k > 0:
            bad = _beyond_conditioning(a, b, k)
        assert not bad, "frame %d: %s   [all: %s]" % (k, bad, " ".join(report))
        worst_all = max(worst_all, worst)
    print("%s: %d frames, worst relative difference %.2e" % (cfg, frames, worst_all))


# frame field -> field of tests/golden/orift_full_conditioning.json
_COND = {"coordinate": "coord", "velocity": "vel", "temperature": "temperature", "stress": "stress",
         "strain": "strain", "strain-rate": "strain_rate", "plastic strain": "plstrain", "viscosity": "viscosity"}


def _beyond_conditioning(a, b, frame):
    """oblique-rift-3d.cfg at its configured length is not a 1e-8 test for ANY implementation that
    is not bit-identical in libm: the reference itself, restarted from an initial stress moved
    by ONE ULP, differs from its own run by 1.7e-4 in velocity and 5e-4 in strain at the first
    output time (tests/orift_conditioning.py -> tests/golden/orift_full_conditioning.json; shear
    localisation amplifies rounding).  The EVP viscosity calls pow/exp and the Mohr-Coulomb return
    goes through atan2/cos/sin of 3x3-C (-ffast-math), so CUDA's libm differs in the last bit
    from step 1 (5e-16 after 300 steps, test above).  Required here: every field within 10x of
    the reference's own one-ulp sensitivity at the same output time, integer fields identical,
    nothing NaN."""
    import json
    import numpy as np
    cond = json.load(open(os.path.join(HERE, "golden", "orift_full_conditioning.json")))["steps"]
    c = cond["19000" if frame == 1 else "38000"]
    generic = 10 * max(c.values())
    bad = []
    for name, x in a.items():
        y = b[name]
        if x.shape != y.shape:
            bad.append("%s: size" % name); continue
        if name in desframe.INT_FIELDS:
            if not np.array_equal(x, y):
                bad.append(name + ": integer field differs")
            continue
        if not (np.isfinite(x).all() and np.isfinite(y).all()):
            bad.append(name + ": NaN/Inf"); continue
        scale = np.abs(x).max() if x.size else 0.0
        rel = (np.abs(x - y).max() / (scale if scale > 0 else 1.0)) if x.size else 0.0
        bound = max(1e-8, 10 * c[_COND[name]]) if name in _COND else max(1e-8, generic)
        if not rel <= bound:
            bad.append("%s: %.3e > %.3e" % (name, rel, bound))
    return bad


@pytest.mark.parametrize("cfg,model,frames,nparts", [
    ("t3d-remesh.cfg", "t3dr", 5, 2),       # download (gather by owner) -> host remesh -> re-upload + REPARTITION
    ("orift.cfg", "orift", 4, 3),           # averaged output fields gathered from three parts
    ("t3d-remesh.cfg", "t3dr", 5, -2),      # the same two parts, each storing its sub-mesh in Morton order
])
def test_driver_over_several_devices_matches_reference_frames(cfg, model, frames, nparts, tmp_path):
    """The same binary with DES_BRIDGE_DEVICES: one handle over several devices (parts share
    device 0 on a one-GPU box).  Frames must equal the reference's as the one-device run's do."""
    import torch
    if not (os.path.exists(REF) and os.path.exists(B200)):
        pytest.skip("reference / bridge binaries not built (need /root/reference at build time)")
    ndev = max(1, torch.cuda.device_count())
    extra = {"DES_REORDER": "morton"} if nparts < 0 else {}
    nparts = abs(nparts)
    devices = ",".join(str(i % ndev) for i in range(nparts))
    run(REF, cfg, str(tmp_path / "ref"))
    out = run(B200, cfg, str(tmp_path / "gpu"), DES_BRIDGE_DEVICES=devices, **extra)
    assert "mesh partitioned over %d devices" % nparts in out
    if cfg == "t3d-remesh.cfg":
        assert out.count("mesh partitioned over") == 2 and "host remesh" in out     # initial + after the remesh
    for k in range(frames):
        fn = "%s.save.%06d" % (model, k)
        a = desframe.read_frame(str(tmp_path / "ref" / fn))
        b = desframe.read_frame(str(tmp_path / "gpu" / fn))
        worst, report, bad = desframe.compare_frames(a, b)
        assert not bad, "frame %d: %s   [all: %s]" % (k, bad, " ".join(report))


def _bytes_without_walltime(path):
    raw = bytearray(open(path, "rb").read())
    for line in bytes(raw[:desframe.HEADER]).splitlines()[1:]:
        if line.startswith(b"walltime_sec\t"):
            p = int(line.split(b"\t")[1])
            raw[p:p + 8] = b"\0" * 8
    return bytes(raw)


@pytest.mark.parametrize("cfg,model,frames", [
    ("t3d.cfg", "t3d", 6),                  # plain fields + marker sets (written synchronously)
    ("orift.cfg", "orift", 4),              # averaged fields, two materials, viscosity
    ("t3d-remesh.cfg", "t3dr", 5),          # the frame buffers follow the new mesh size
    ("t3d-nomarkers.cfg", "t3dnm", 6),      # background thread
])
def test_direct_frames_are_the_files_output_write_makes(cfg, model, frames, tmp_path):
    """Frames written from the device in the on-disk layout (des_gpu_download_frame +
    bridge/des_frame_writer.cxx, the default) against the same binary going through Variables and
    the reference's Output::write (DES_BRIDGE_DIRECT_OUTPUT=0): the same bytes, header included,
    except the wall-clock scalar; and the same .info rows except their wall-clock column."""
    if not os.path.exists(B200):
        pytest.skip("bridge binary not built")
    out = run(B200, cfg, str(tmp_path / "direct"), DES_BRIDGE_FRAME_TIMING="1")
    run(B200, cfg, str(tmp_path / "write"), DES_BRIDGE_DIRECT_OUTPUT="0")
    assert "[B200] frame from the device" in out
    if cfg == "t3d-nomarkers.cfg":
        assert "in the background" in out
    for k in range(frames):
        fn = "%s.save.%06d" % (model, k)
        a = _bytes_without_walltime(str(tmp_path / "write" / fn))
        b = _bytes_without_walltime(str(tmp_path / "direct" / fn))
        assert len(a) == len(b), (fn, len(a), len(b))
        assert a == b, fn
    rows = [[l.split() for l in open(str(tmp_path / w / (model + ".info")))] for w in ("write", "direct")]
    assert len(rows[0]) == len(rows[1]) == frames
    for ra, rb in zip(*rows):
        assert len(ra) == len(rb)
        diff = [i for i, (x, y) in enumerate(zip(ra, rb)) if x != y]
        assert len(diff) <= 1, (ra, rb)     # the run-time column


def test_driver_restart_matches_reference_frames(tmp_path):
    """sim.is_restarting: restart() (dynearthsol.cxx:231-437) rebuilds the host state from a
    checkpoint; the bridge uploads it like any other state.  The reference writes the
    checkpoints (t3d-ckpt.cfg), then both binaries restart from its frame 2 (t3d-restart.cfg)."""
    import shutil
    if not (os.path.exists(REF) and os.path.exists(B200)):
        pytest.skip("reference / bridge binaries not built (need /root/reference at build time)")
    run(REF, "t3d-ckpt.cfg", str(tmp_path / "ref"))
    os.makedirs(str(tmp_path / "gpu"))
    for f in os.listdir(str(tmp_path / "ref")):
        if f.startswith("t3dck."):
            shutil.copy(str(tmp_path / "ref" / f), str(tmp_path / "gpu"))
    run(REF, "t3d-restart.cfg", str(tmp_path / "ref"))
    out = run(B200, "t3d-restart.cfg", str(tmp_path / "gpu"))
    assert "Initializing mesh and field data from checkpoints" in out
    for k in (2, 3, 4, 5):
        fn = "t3drs.save.%06d" % k
        a = desframe.read_frame(str(tmp_path / "ref" / fn))
        b = desframe.read_frame(str(tmp_path / "gpu" / fn))
        worst, report, bad = desframe.compare_frames(a, b)
        assert not bad, "frame %d: %s   [all: %s]" % (k, bad, " ".join(report))


def test_bench_driver_mode(tmp_path):
    """bench.py --driver: the metric through the reference's own driver -- a generated
    meshing_option = 1 box .cfg, init() on the host with the wrapped topology builders
    (bridge/des_topology_wrap.cxx), the step loop on the GPU, the reference's "Compute" line."""
    import json
    import sys
    if not os.path.exists(B200):
        pytest.skip("bridge binary not built")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--driver", "--size", "100k", "--steps", "20"],
                       capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stderr[-800:]
    line = json.loads([l for l in r.stdout.splitlines() if l.startswith("{")][-1])
    assert line.get("impl") == "driver" and "unavailable" not in line, line
    assert line["value"] > 0 and line["steps"] >= 300
    assert "103680 tets" in line["config"]["workload"]
