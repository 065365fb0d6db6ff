"""Runs dynearthsol3d.ref (the unmodified reference, all host cores) and dynearthsol3d.b200 (the
same object code with the step loop on the GPU) on the named small configurations and records
the reference's own metric, "Compute ... s/step" (dynearthsol.cxx:990-1019), plus the printed
download / host remesh / re-upload seconds of the remeshing case.  The last frame of every run
is copied to <out>/frames so that benchmarks-cores/compare.py can be run on them where
/root/reference exists (tests/compare_frames.sh).

    python tests/bridge_perf.py <out_dir> [cfg ...]
"""
import json
import os
import re
import shutil
import subprocess
import sys
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
REF = os.path.join(ROOT, "oracle", "_ref", "dynearthsol3d.ref")
B200 = os.path.join(ROOT, "bridge", "_build", "dynearthsol3d.b200")
CASES = {"t3d.cfg": "t3d", "t3d-remesh.cfg": "t3dr", "orift-full.cfg": "oriftfull"}


def run(binary, cfg, wd, env=None):
    os.makedirs(wd, exist_ok=True)
    e = dict(os.environ, OMP_NUM_THREADS=str(os.cpu_count() or 1))
    e.update(env or {})
    r = subprocess.run([binary, os.path.join(HERE, "cfg", cfg)], cwd=wd, env=e, capture_output=True, text=True,
                       timeout=1200)
    assert r.returncode == 0, (binary, cfg, r.stdout[-800:], r.stderr[-800:])
    m = re.search(r"Compute\s*:\s*\S+.*?/\s*(\d+)\s*=\s*([0-9.eE+-]+)\s*s/step", r.stdout)
    out = {"steps": int(m.group(1)), "compute_s_per_step": float(m.group(2))}
    m = re.search(r"Remesh\s*:\s*\S+.*?/\s*(\d+)\s*=\s*([0-9.eE+-]+)\s*s/remesh", r.stdout)
    if m:       # the binary's own "Remesh" line (dynearthsol.cxx:1004-1011): seconds per remesh, 10 ms resolution
        out["remesh_s_per_remesh"] = float(m.group(2))
    rm = re.findall(r"\[B200\] download \+ host remesh \+ re-upload:.*?\(download ([0-9.e+-]+) s, host remesh ([0-9.e+-]+) s"
                    r"(?: of which field transfer on the device ([0-9.e+-]+) s)?, re-upload ([0-9.e+-]+) s\)", r.stdout)
    if rm:
        out["remesh_cycles"] = [{"download_s": float(a), "host_remesh_s": float(b),
                                 "device_field_transfer_s": float(t) if t else None, "reupload_s": float(c)}
                                for a, b, t, c in rm]
    pm = re.findall(r"\[B200\] mesh partitioned over (\d+) devices: partition ([0-9.]+) s, upload ([0-9.]+) s", r.stdout)
    if pm:
        out["partition"] = [{"parts": int(a), "partition_s": float(b), "upload_s": float(c)} for a, b, c in pm]
    return out


def main():
    outdir = sys.argv[1]
    cfgs = sys.argv[2:] or list(CASES)
    os.makedirs(os.path.join(outdir, "frames"), exist_ok=True)
    res = {"host_cores": os.cpu_count(), "cases": {}}
    for cfg in cfgs:
        model = CASES[cfg]
        wd = tempfile.mkdtemp(prefix="bridge_perf_")
        c = {"reference": run(REF, cfg, os.path.join(wd, "ref")), "b200": run(B200, cfg, os.path.join(wd, "gpu"))}
        if cfg == "t3d-remesh.cfg":
            c["b200_two_parts_on_one_device"] = run(B200, cfg, os.path.join(wd, "gpu2"), {"DES_BRIDGE_DEVICES": "0,0"})
        c["speedup_compute"] = c["reference"]["compute_s_per_step"] / c["b200"]["compute_s_per_step"]
        res["cases"][cfg] = c
        for side in ("ref", "gpu"):
            frames = sorted(f for f in os.listdir(os.path.join(wd, side)) if ".save." in f)
            d = os.path.join(outdir, "frames", side)
            os.makedirs(d, exist_ok=True)
            last = frames[-1]
            shutil.copy(os.path.join(wd, side, last), d)
            info = os.path.join(wd, side, model + ".info")
            if os.path.exists(info):
                shutil.copy(info, d)
            c.setdefault("last_frame", int(last.rsplit(".", 1)[1]))
        shutil.rmtree(wd, ignore_errors=True)
    json.dump(res, open(os.path.join(outdir, "bridge_perf.json"), "w"), indent=1)
    print(json.dumps(res, indent=1))


if __name__ == "__main__":
    main()
