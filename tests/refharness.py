"""ctypes wrapper over oracle/_ref/libdesref.so (the UNMODIFIED reference, test infra only)."""
import ctypes as C
import os
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(HERE, "..", "oracle", "_ref", "libdesref.so")


def available():
    return os.path.exists(LIB)


class Ref:
    """One reference model per process (the reference keeps static state)."""

    def __init__(self, cfg_path, workdir=None, fast_init=False):
        # fast_init: bench.py's CPU arm on meshes the reference's O(nseg * nelem) boundary-facet
        # search cannot initialise (oracle/ref_harness.cxx: desref_fast_init); never for fixtures
        self.lib = C.CDLL(LIB)
        L = self.lib
        if fast_init:
            L.desref_fast_init(1)
        L.desref_init.argtypes = [C.c_char_p]
        L.desref_nbytes.argtypes = [C.c_char_p]; L.desref_nbytes.restype = C.c_long
        L.desref_get.argtypes = [C.c_char_p, C.c_void_p, C.c_long]
        L.desref_set.argtypes = [C.c_char_p, C.c_void_p, C.c_long]
        L.desref_get_scalar.argtypes = [C.c_char_p, C.POINTER(C.c_double)]
        L.desref_set_scalar.argtypes = [C.c_char_p, C.c_double]
        L.desref_get_int.argtypes = [C.c_char_p, C.POINTER(C.c_int)]
        L.desref_set_int.argtypes = [C.c_char_p, C.c_int]
        L.desref_get_matvec.argtypes = [C.c_char_p, C.c_void_p, C.c_int]
        L.desref_call.argtypes = [C.c_char_p]
        L.desref_step.argtypes = [C.c_int]; L.desref_step.restype = C.c_double
        L.desref_l2_residual.restype = C.c_double
        L.desref_bad_mesh_quality.argtypes = [C.c_double, C.c_double, C.c_double, C.c_int,
                                              C.POINTER(C.c_int), C.POINTER(C.c_double)]
        L.desref_get_mesh_param.argtypes = [C.c_char_p, C.POINTER(C.c_double)]
        cwd = os.getcwd()
        if workdir:
            os.chdir(workdir)
        try:
            rc = L.desref_init(os.path.abspath(os.path.join(cwd, cfg_path)).encode())
        finally:
            os.chdir(cwd)
        assert rc == 0, rc

    _INT = {"connectivity", "bcflag", "support_arr", "support_idx", "support_lidx", "top_elems",
            "top_nodes", "support_surf_arr", "support_surf_idx", "elem_and_nodes",
            "connectivity_surface", "edge_slot", "vbc_types", "stress_bc_types", "elemmarkers",
            "segment", "segflag"}

    def get(self, name):
        n = self.lib.desref_nbytes(name.encode())
        assert n >= 0, name
        isint = name in self._INT or name.startswith("bnodes") or name.startswith("bfacets")
        dt = np.uint32 if name == "bcflag" else (np.int32 if isint else np.float64)
        a = np.empty(n // np.dtype(dt).itemsize, dtype=dt)
        rc = self.lib.desref_get(name.encode(), a.ctypes.data, n)
        assert rc == 0, name
        return a

    def set(self, name, arr):
        arr = np.ascontiguousarray(arr)
        rc = self.lib.desref_set(name.encode(), arr.ctypes.data, arr.nbytes)
        assert rc == 0, name

    def scalar(self, name):
        v = C.c_double()
        assert self.lib.desref_get_scalar(name.encode(), C.byref(v)) == 0, name
        return v.value

    def set_scalar(self, name, v):
        assert self.lib.desref_set_scalar(name.encode(), float(v)) == 0, name

    def int(self, name):
        v = C.c_int()
        assert self.lib.desref_get_int(name.encode(), C.byref(v)) == 0, name
        return v.value

    def set_int(self, name, v):
        assert self.lib.desref_set_int(name.encode(), int(v)) == 0, name

    def matvec(self, name):
        n = self.int("nmat")
        a = np.empty(n)
        assert self.lib.desref_get_matvec(name.encode(), a.ctypes.data, n) == 0, name
        return a

    def call(self, fn):
        assert self.lib.desref_call(fn.encode()) == 0, fn

    def step(self, n=1):
        return self.lib.desref_step(n)

    def mesh_param(self, name):
        v = C.c_double()
        assert self.lib.desref_get_mesh_param(name.encode(), C.byref(v)) == 0, name
        return v.value

    def bad_mesh_quality(self, smallest_size, max_boundary_distortion, min_quality, remeshing_option):
        """bad_mesh_quality (remeshing.cxx:2752) with those Param.mesh knobs: (code, index, min_quality);
        index is -12345 and min_quality NaN where the reference leaves them unwritten."""
        idx = C.c_int(-12345)
        mq = C.c_double(float("nan"))
        code = self.lib.desref_bad_mesh_quality(smallest_size, max_boundary_distortion, min_quality,
                                                remeshing_option, C.byref(idx), C.byref(mq))
        return code, idx.value, mq.value
