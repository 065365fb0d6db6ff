import os
import subprocess
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)
sys.path.insert(0, HERE)

REFLIB = os.path.join(ROOT, "oracle", "_ref", "libdesref.so")
_cache = {}


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def have_gpu():
    import ctypes
    try:
        cuda = ctypes.CDLL("libcudart.so.12")
    except OSError:
        try:
            cuda = ctypes.CDLL("/usr/local/cuda/lib64/libcudart.so")
        except OSError:
            return False
    n = ctypes.c_int(0)
    return cuda.cudaGetDeviceCount(ctypes.byref(n)) == 0 and n.value > 0


def cfg_variant(case, over):
    """tests/cfg/<case>.cfg with `section.key = value` entries replaced (or added to their
    section): the same file the reference would be given for that variant.  Returns its path."""
    import hashlib
    lines = open(os.path.join(HERE, "cfg", case + ".cfg")).read().splitlines()
    todo = {k: str(v) for k, v in over.items()}
    out, sec = [], None
    pending = {}                                    # section -> lines to add under its header
    for k, v in todo.items():
        pending.setdefault(k.split(".", 1)[0], {})[k.split(".", 1)[1]] = v
    for ln in lines:
        st = ln.strip()
        if st.startswith("[") and st.endswith("]"):
            sec = st[1:-1]
            out.append(ln)
            continue
        if sec in pending and "=" in st and not st.startswith("#"):
            key = st.split("=", 1)[0].strip()
            if key in pending[sec]:
                out.append("%s = %s" % (key, pending[sec].pop(key)))
                continue
        out.append(ln)
    final, sec = [], None
    for ln in out:
        final.append(ln)
        st = ln.strip()
        if st.startswith("[") and st.endswith("]") and st[1:-1] in pending:
            for key, v in pending.pop(st[1:-1]).items():
                final.append("%s = %s" % (key, v))
    for sec, kv in pending.items():
        if kv:
            final.append("[%s]" % sec)
            final += ["%s = %s" % (k, v) for k, v in kv.items()]
    text = "\n".join(final) + "\n"
    tag = hashlib.sha1(text.encode()).hexdigest()[:10]
    path = os.path.join("/tmp", "descfg_%s_%s.cfg" % (case, tag))
    with open(path, "w") as f:
        f.write(text)
    return path, tag


def ref_dump(case, steps, sets=(), over=None, perturb_at=None):
    """Reference inputs/outputs for tests/cfg/<case>.cfg: generated LIVE from the unmodified
    reference when oracle/_ref/libdesref.so is present (one subprocess per dump: the
    reference keeps process-wide state), else the committed golden fixture.
    over: {"section.key": value} cfg entries replaced before the reference reads the file;
    sets: (name, int) switches flipped after init() (refdump.py --set)."""
    key = (case, tuple(steps), tuple(sets), tuple(sorted((over or {}).items())), perturb_at)
    if key in _cache:
        return _cache[key]
    golden = os.path.join(HERE, "golden", case + ".npz")
    if os.path.exists(REFLIB):
        cfg, tag = (os.path.join(HERE, "cfg", case + ".cfg"), "") if not over else cfg_variant(case, over)
        out = os.path.join("/tmp", "desref_%s%s_%s_%s%s.npz" % (
            case, tag, "-".join(map(str, steps)), "-".join("%s%d" % s for s in sets),
            "" if perturb_at is None else "_ulp%d" % perturb_at))
        cmd = [sys.executable, os.path.join(HERE, "refdump.py"), cfg,
               out, "--steps", ",".join(map(str, steps))]
        for s in sets:
            cmd += ["--set", "%s=%d" % s]
        if perturb_at is not None:
            cmd += ["--perturb-at", str(perturb_at)]
        subprocess.run(cmd, check=True, stdout=subprocess.DEVNULL, stderr=subprocess.PIPE,
                       env=dict(os.environ, OMP_NUM_THREADS="4"))
        d = dict(np.load(out))
    elif os.path.exists(golden) and not sets and not over and perturb_at is None:
        d = dict(np.load(golden))
        missing = [k for k in steps if ("s%d_coord" % k) not in d]
        if missing:
            pytest.skip("golden fixture %s has no step(s) %s and libdesref.so is absent" % (case, missing))
    else:
        pytest.skip("no reference for case %s (libdesref.so absent, no golden fixture)" % case)
    _cache[key] = d
    return d


def ref_branches(suite, case="box-ep"):
    """Branch-level reference vectors (tests/refbranches.py), generated live from the unmodified
    reference in a subprocess; skipped where libdesref.so is absent."""
    key = ("branches", suite, case)
    if key in _cache:
        return _cache[key]
    if not os.path.exists(REFLIB):
        pytest.skip("oracle/_ref/libdesref.so absent: no reference to generate the %s vectors" % suite)
    out = os.path.join("/tmp", "desref_branches_%s_%s.npz" % (suite, case))
    subprocess.run([sys.executable, os.path.join(HERE, "refbranches.py"), suite,
                    os.path.join(HERE, "cfg", case + ".cfg"), out], check=True,
                   stdout=subprocess.DEVNULL, stderr=subprocess.PIPE, env=dict(os.environ, OMP_NUM_THREADS="4"))
    _cache[key] = dict(np.load(out))
    return _cache[key]


def ref_conditioning(case, steps, over=None, at=0):
    """{step: {field: relative change}} of the reference's OWN snapshots when every stress entry
    of its state at step `at` is moved by one ulp (tests/refdump.py --perturb-at).  Where plastic
    return mapping is active this is far above 1e-10: the Cardano eigenvalues lose ~1e5 of
    precision to cancellation (3x3-C/dsyevc3.c) and the return direction follows them, so a
    last-bit difference in libm (atan2/cos/sin: CUDA's vs glibc's) or in operation order
    (the reference's 3x3-C is built -O3 -ffast-math) moves the result by the same amount."""
    import refmodel
    a = ref_dump(case, steps, over=over)
    b = ref_dump(case, steps, over=over, perturb_at=at)
    out = {}
    for k in steps:
        if k <= at:
            continue
        out[k] = {f[len("s%d_" % k):]: refmodel.rel_err(b[f], a[f]) for f in a if f.startswith("s%d_" % k)
                  and a[f].ndim == 1 and a[f].size > 1}
    return out
