"""Reader/differ for the reference's binary output frames (<model>.save.NNNNNN): a 4096-byte
ASCII header of `name<TAB>offset` lines, then raw little-endian arrays (binaryio.cxx:38-39,
106-176).  The measure is compare.py's: max|new - old| / max|old| per field, fail at 1e-8
(benchmarks-cores/compare.py:111-129,149-169) -- applied to every float field of the frame."""
import os
import numpy as np

INT_FIELDS = ("connectivity", "bcflag", "steps", "nnode", "nelem", "nseg", "markerset size", "markerset.elem",
              "markerset.mattype", "markerset.id", "markerset.genesis")
SKIP = ("walltime_sec",)
HEADER = 4096


def is_int_field(name):
    """integer arrays of a frame, for every marker set ("markerset", "hydrous-markerset", ...)"""
    return (name in INT_FIELDS or name.endswith("markerset size")
            or ("markerset." in name and name.rsplit(".", 1)[1] in ("elem", "mattype", "id", "genesis")))


def read_frame(path):
    with open(path, "rb") as f:
        raw = f.read()
    lines = raw[:HEADER].splitlines()
    assert lines[0].startswith(b"# DynEarthSol"), path
    pos = []
    for line in lines[1:]:
        if not line or line[0] == 0:
            break
        name, p = line.split(b"\t")
        pos.append((name.decode("ascii"), int(p)))
    out = {}
    ends = [p for _, p in pos[1:]] + [len(raw)]
    for (name, p), e in zip(pos, ends):
        if name in SKIP:
            continue
        dt = np.int32 if is_int_field(name) else np.float64
        out[name] = np.frombuffer(raw[p:e], dtype=dt)
    return out


def compare_frames(old, new, tol=1e-8):
    """Returns (worst relative difference over float fields, report lines, failures)."""
    worst, report, bad = 0.0, [], []
    for name, a in old.items():
        if name not in new:
            bad.append(name + ": missing"); continue
        b = new[name]
        if a.shape != b.shape:
            bad.append("%s: size %d vs %d" % (name, a.size, b.size)); continue
        if is_int_field(name):
            if not np.array_equal(a, b):
                bad.append(name + ": integer field differs")
            continue
        if not (np.isfinite(a).all() and np.isfinite(b).all()):
            bad.append(name + ": NaN/Inf"); continue
        scale = np.abs(a).max() if a.size else 0.0
        err = np.abs(a - b).max() if a.size else 0.0
        rel = 0.0 if err == 0 else err / (scale if scale > 0 else 1.0)
        report.append("%s:%s" % (name, "bit-exact" if err == 0 else "%.2e" % rel))
        worst = max(worst, rel)
        if not rel < tol:
            bad.append("%s: %.3e" % (name, rel))
    return worst, report, bad
