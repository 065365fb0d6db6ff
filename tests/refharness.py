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

    def __init__(self, cfg_path, workdir=None, fast_init=False, lib=None):
        # fast_init: bench.py's CPU arm on meshes the reference's O(nseg * nelem) boundary-facet
        # search cannot initialise (oracle/ref_harness.cxx: desref_fast_init); never for fixtures
        self.lib = C.CDLL(lib or LIB)
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

    # ---- remesh field transfer ----
    _CAPTURED_INT = ("_conn", "is_changed", "bcflag", "support_arr", "support_idx")

    def remesh_captured(self, bad_quality=1):
        """The reference's own remesh() (remeshing.cxx:2869) on the current state, with the inputs
        and outputs of its nearest_neighbor_interpolation / barycentric_node_interpolation calls
        recorded by the harness: a dict name -> flat array (nn0_ elements, nn1_ top surface,
        brc_ nodes; old_* / new_* meshes, src_* / dst_* fields in the reference's own layout)."""
        L = self.lib
        L.desref_capture_transfers(1)
        assert L.desref_remesh(int(bad_quality)) == 0
        return self._captured()

    def transfer_nodes_captured(self, new_coord, new_bcflag):
        """The reference's barycentric_node_interpolation from the current mesh and nodal fields onto
        arbitrary points new_coord ([3][n] flat), recorded like remesh_captured: the brc_ entries."""
        L = self.lib
        new_coord = np.ascontiguousarray(new_coord, dtype=np.float64)
        new_bcflag = np.ascontiguousarray(new_bcflag, dtype=np.uint32)
        L.desref_transfer_nodes_case.argtypes = [C.c_int, C.c_void_p, C.c_void_p]
        L.desref_capture_transfers(1)
        assert L.desref_transfer_nodes_case(new_bcflag.size, new_coord.ctypes.data, new_bcflag.ctypes.data) == 0
        return {k: v for k, v in self._captured().items() if k.startswith("brc_")}

    def transfer_cells_captured(self, new_coord, new_conn):
        """The reference's nearest_neighbor_interpolation of the element fields from the current mesh
        onto (new_coord [3][n], new_conn [4][e]), recorded: (the nn0_ entries, the call's seconds)."""
        L = self.lib
        new_coord = np.ascontiguousarray(new_coord, dtype=np.float64).reshape(-1)
        new_conn = np.ascontiguousarray(new_conn, dtype=np.int32).reshape(-1)
        L.desref_transfer_cells_case.argtypes = [C.c_int, C.c_void_p, C.c_int, C.c_void_p]
        L.desref_transfer_cells_case.restype = C.c_double
        L.desref_capture_transfers(1)
        sec = L.desref_transfer_cells_case(new_coord.size // 3, new_coord.ctypes.data, new_conn.size // 4,
                                           new_conn.ctypes.data)
        return {k: v for k, v in self._captured().items() if k.startswith("nn0_")}, sec

    def _captured(self):
        L = self.lib
        L.desref_captured_nbytes.argtypes = [C.c_char_p]; L.desref_captured_nbytes.restype = C.c_long
        L.desref_captured_get.argtypes = [C.c_char_p, C.c_void_p, C.c_long]
        out = {}
        for pre, names in (("nn0_", ("old_coord", "old_conn", "new_coord", "new_conn", "old_size", "is_changed")
                            + tuple(d + f for d in ("src_", "dst_") for f in (
                                "plstrain", "delta_plstrain", "strain", "radiogenic_source", "dyn_fric_coeff",
                                "state_variable", "volume_old", "stress", "stressyy", "spr_blend_weight",
                                "spr_p_ref_old"))),
                           ("nn1_", ("old_coord", "old_conn", "new_coord", "new_conn", "old_size",
                                     "src_edvacc_surf", "dst_edvacc_surf")),
                           ("brc_", ("old_coord", "old_conn", "new_coord", "old_size", "new_bcflag",
                                     "old_support_arr", "old_support_idx")
                            + tuple(d + f for d in ("src_", "dst_") for f in (
                                "temperature", "ppressure", "dppressure", "vel", "coord0", "stress_n")))):
            for nm in names:
                key = pre + nm
                n = L.desref_captured_nbytes(key.encode())
                if n < 0:
                    continue
                isint = any(nm.endswith(t) for t in self._CAPTURED_INT)
                dt = np.uint32 if nm.endswith("bcflag") else (np.int32 if isint else np.float64)
                a = np.empty(n // np.dtype(dt).itemsize, dtype=dt)
                if n:
                    assert L.desref_captured_get(key.encode(), a.ctypes.data, n) == 0, key
                out[key] = a
        L.desref_capture_transfers(0)
        return out

    # ---- the material marker set (markersets[0]) and phase_changes ----
    def markers(self):
        """(elem [n], eta [4][n], mattype [n], tmp [n]) through MarkerSet's own accessors."""
        L = self.lib
        n = L.desref_marker_count()
        elem = np.empty(n, np.int32); mt = np.empty(n, np.int32); eta = np.empty((4, n)); tmp = np.empty(n)
        L.desref_markers_get.argtypes = [C.c_void_p] * 4
        assert L.desref_markers_get(elem.ctypes.data, eta.ctypes.data, mt.ctypes.data, tmp.ctypes.data) == n
        return elem, eta, mt, tmp

    def set_marker_mattype(self, mattype):
        mattype = np.ascontiguousarray(mattype, dtype=np.int32)
        self.lib.desref_markers_set_mattype.argtypes = [C.c_void_p]
        assert self.lib.desref_markers_set_mattype(mattype.ctypes.data) == 0

    def hydrous_count(self, new=None):
        """(*var.hydrous_elemmarkers)[e][0]; None when the model has no hydrous marker set."""
        a = np.zeros(self.int("nelem"), np.int32) if new is None else np.ascontiguousarray(new, dtype=np.int32)
        self.lib.desref_hydrous_count.argtypes = [C.c_void_p, C.c_int]
        rc = self.lib.desref_hydrous_count(a.ctypes.data, 0 if new is None else 1)
        return None if rc else a

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
