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


def ref_dump(case, steps, sets=()):
    """Reference inputs/outputs for tests/cfg/<case>.cfg: generated LIVE from the unmodified
    reference when oracle/_ref/libdesref.so is present (one subprocess per dump: the
    reference keeps process-wide state), else the committed golden fixture."""
    key = (case, tuple(steps), tuple(sets))
    if key in _cache:
        return _cache[key]
    golden = os.path.join(HERE, "golden", case + ".npz")
    if os.path.exists(REFLIB):
        out = os.path.join("/tmp", "desref_%s_%s_%s.npz" % (
            case, "-".join(map(str, steps)), "-".join("%s%d" % s for s in sets)))
        cmd = [sys.executable, os.path.join(HERE, "refdump.py"), os.path.join(HERE, "cfg", case + ".cfg"),
               out, "--steps", ",".join(map(str, steps))]
        for s in sets:
            cmd += ["--set", "%s=%d" % s]
        subprocess.run(cmd, check=True, stdout=subprocess.DEVNULL, stderr=subprocess.PIPE,
                       env=dict(os.environ, OMP_NUM_THREADS="4"))
        d = dict(np.load(out))
    elif os.path.exists(golden) and not sets:
        d = dict(np.load(golden))
        missing = [k for k in steps if ("s%d_coord" % k) not in d]
        if missing:
            pytest.skip("golden fixture %s has no step(s) %s and libdesref.so is absent" % (case, missing))
    else:
        pytest.skip("no reference for case %s (libdesref.so absent, no golden fixture)" % case)
    _cache[key] = d
    return d
