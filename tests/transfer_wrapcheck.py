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
    # barycentric_node_interpolation_forT onto the synthetic points of transfer_capture.py, with a
    # non-trivial "input" temperature
    import ctypes as C
    from transfer_capture import synthetic_points
    pts, _ = synthetic_points(r)
    n = r.int("nnode")
    x = r.get("coord").reshape(3, n)
    T0 = r.get("temperature")
    r.set("temperature", 300.0 + 1e-2 * np.abs(x[2]) + 1e-4 * x[0] + 3e-4 * x[1])
    forT = np.empty(pts.size // 3)
    r.lib.desref_forT_case.argtypes = [C.c_int, C.c_void_p, C.c_void_p]
    assert r.lib.desref_forT_case(forT.size, pts.ctypes.data, forT.ctypes.data) == 0
    r.set("temperature", T0)
    cwd = os.getcwd(); os.chdir(work)
    try:
        assert r.lib.desref_remesh(1) == 0
    finally:
        os.chdir(cwd)
    r.step(5)                                   # the transferred state is stepped on
    d = {k: r.get(k) for k in FIELDS}
    d["dt"] = np.array([r.scalar("dt")])
    d["forT"] = forT
    np.savez(out, **d)


if __name__ == "__main__":
    main()
