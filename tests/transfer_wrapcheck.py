"""One process: the reference stepped, then its own remesh(), then every field dumped to an .npz.
With --wrapcheck the library is oracle/_ref/libdesref_wrapcheck.so, where remesh()'s two transfer
calls go through bridge/des_transfer_wrap.cxx (see oracle/Makefile).  tests/test_transfer.py
compares the two dumps.  usage: python tests/transfer_wrapcheck.py CFG STEPS OUT.npz [--wrapcheck]"""
import os
import sys
import tempfile
import numpy as np

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import refharness  # noqa: E402

FIELDS = ["coord", "connectivity", "vel", "coord0", "temperature", "stress", "strain", "plstrain",
          "delta_plstrain", "volume", "volume_old", "radiogenic_source", "edvacc_surf", "mass", "tmass",
          "volume_n", "support_idx", "support_arr", "elemmarkers"]


def main():
    cfg, steps, out = sys.argv[1], int(sys.argv[2]), sys.argv[3]
    lib = None
    if "--wrapcheck" in sys.argv:
        lib = os.path.join(os.path.dirname(refharness.LIB), "libdesref_wrapcheck.so")
    work = tempfile.mkdtemp(prefix="deswc_")
    r = refharness.Ref(cfg, workdir=work, lib=lib)
    r.step(steps)
    cwd = os.getcwd(); os.chdir(work)
    try:
        assert r.lib.desref_remesh(1) == 0
    finally:
        os.chdir(cwd)
    r.step(5)                                   # the transferred state is stepped on
    d = {k: r.get(k) for k in FIELDS}
    d["dt"] = np.array([r.scalar("dt")])
    np.savez(out, **d)


if __name__ == "__main__":
    main()
