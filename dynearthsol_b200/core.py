"""ctypes binding of the C-ABI in include/des_gpu.h (libdes_b200.so) plus a thin host-side
mirror of the reference's step-loop call sites (dynearthsol.cxx:768-894).

There is no CPU fallback: if the CUDA library is missing this module raises on load, and if
no device is present `DesGpu(...)` raises.  Nothing here imports `oracle/`.
"""
import ctypes as C
import os
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("DES_LIB") or os.path.join(_HERE, "csrc", "libdes_b200.so")   # DES_LIB: A/B builds

DES_MAX_MAT = 16
DES_NBDRY = 10
DES_NKCLASS = 16

RH_ELASTIC, RH_VISCOUS, RH_MAXWELL, RH_EP, RH_EVP = 1, 2, 3, 5, 7

ERRORS = {0: "OK", 1: "NODEVICE", 2: "CUDA", 3: "ARG", 4: "UNSUPPORTED", 5: "STATE", 11: "NAN"}


class DesError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("des_gpu error %d (%s): %s" % (code, ERRORS.get(code, "?"), msg))
        self.code = code


_MATV = C.c_double * DES_MAX_MAT


class DesParam(C.Structure):
    _fields_ = [
        ("gravity", C.c_double), ("inertial_scaling", C.c_double), ("damping_factor", C.c_double),
        ("surface_diffusivity", C.c_double), ("surf_base_level", C.c_double),
        ("dt_fraction", C.c_double), ("fixed_dt", C.c_double), ("characteristic_speed", C.c_double),
        ("damping_option", C.c_int), ("is_quasi_static", C.c_int), ("has_thermal_diffusion", C.c_int),
        ("is_using_mixed_stress", C.c_int), ("surface_process_option", C.c_int),
        ("ref_pressure_option", C.c_int), ("has_moving_mesh", C.c_int),
        ("has_body_force_adjustment", C.c_int),
        ("has_winkler_foundation", C.c_int), ("has_water_loading", C.c_int),
        ("has_elastic_foundation", C.c_int),
        ("winkler_delta_rho", C.c_double), ("elastic_foundation_constant", C.c_double),
        ("sea_water_density", C.c_double), ("surface_temperature", C.c_double),
        ("vbc_types", C.c_int * DES_NBDRY), ("vbc_values", C.c_double * DES_NBDRY),
        ("vbc_val_lateral", C.c_double * 4), ("vbc_val_z1_loading_period", C.c_double),
        ("stress_bc_types", C.c_int * 6), ("max_vbc_val", C.c_double),
        ("zlength", C.c_double),
        ("rheol_type", C.c_int), ("nmat", C.c_int), ("mattype_ref", C.c_int),
        ("visc_min", C.c_double), ("visc_max", C.c_double), ("tension_max", C.c_double),
        ("therm_diff_max", C.c_double),
        ("rho0", _MATV), ("alpha", _MATV), ("bulk_modulus", _MATV), ("shear_modulus", _MATV),
        ("visc_exponent", _MATV), ("visc_coefficient", _MATV),
        ("visc_activation_energy", _MATV), ("visc_activation_volume", _MATV),
        ("heat_capacity", _MATV), ("therm_cond", _MATV),
        ("pls0", _MATV), ("pls1", _MATV), ("cohesion0", _MATV), ("cohesion1", _MATV),
        ("friction_angle0", _MATV), ("friction_angle1", _MATV),
        ("dilation_angle0", _MATV), ("dilation_angle1", _MATV), ("porosity", _MATV),
        ("stress_bc_values", C.c_double * 6),
    ]


MAT_ARRAYS = ["rho0", "alpha", "bulk_modulus", "shear_modulus", "visc_exponent", "visc_coefficient",
              "visc_activation_energy", "visc_activation_volume", "heat_capacity", "therm_cond",
              "pls0", "pls1", "cohesion0", "cohesion1", "friction_angle0", "friction_angle1",
              "dilation_angle0", "dilation_angle1", "porosity"]

_IP = C.POINTER(C.c_int)
_DP = C.POINTER(C.c_double)


class DesMesh(C.Structure):
    _fields_ = [
        ("nnode", C.c_int), ("nelem", C.c_int),
        ("connectivity", _IP), ("bcflag", C.POINTER(C.c_uint)),
        ("support_idx", _IP), ("support_arr", _IP), ("support_lidx", _IP),
        ("nbfacets", C.c_int * DES_NBDRY), ("bfacets", _IP * DES_NBDRY),
        ("nbnodes", C.c_int * DES_NBDRY), ("bnodes", _IP * DES_NBDRY),
        ("bnormals", _DP), ("n_edge_vec", C.c_int), ("edge_vec", _DP), ("edge_slot", _IP),
        ("ntop_elems", C.c_int), ("top_elems", _IP),
        ("ntop", C.c_int), ("top_nodes", _IP), ("etop", C.c_int),
        ("support_surf_idx", _IP), ("support_surf_arr", _IP),
        ("elem_and_nodes", _IP), ("connectivity_surface", _IP), ("coord", _DP),
    ]


STATE_FIELDS = ["coord", "vel", "coord0", "temperature", "stress", "strain", "strain_rate",
                "plstrain", "delta_plstrain", "volume", "volume_old", "viscosity", "dpressure",
                "edvoldt", "radiogenic_source", "force", "volume_n", "mass", "tmass", "ntmp",
                "dhacc", "edvacc_surf"]


class DesState(C.Structure):
    _fields_ = [(f, _DP) for f in STATE_FIELDS]


class DesDtReductions(C.Structure):
    _fields_ = [("minl", C.c_double), ("dt_maxwell", C.c_double), ("dt_diffusion", C.c_double),
                ("global_dt_min", C.c_double), ("global_max_vem", C.c_double),
                ("max_surf_vel", C.c_double)]


FRAME_NODE3 = ("coordinate", "velocity", "velocity_averaged", "force")
FRAME_NODE1 = ("temperature",)
FRAME_ELEM6 = ("strain_rate", "strain", "stress", "stress_averaged")
FRAME_ELEM1 = ("plstrain", "plstrain_rate", "viscosity", "density", "mesh_quality", "material")


class DesFrame(C.Structure):
    """The arrays of one output frame in the on-disk layout (include/des_gpu.h: DesFrame)."""
    _fields_ = ([("averaged", C.c_int), ("inv_dt", C.c_double), ("stress_avg_scale", C.c_double)] +
                [(k, C.POINTER(C.c_double)) for k in (
                    "coordinate", "velocity", "velocity_averaged", "temperature", "plstrain",
                    "plstrain_rate", "strain_rate", "strain", "stress", "viscosity", "stress_averaged",
                    "density", "mesh_quality", "material", "force")])


class DesTransferMeshes(C.Structure):
    _fields_ = [("old_nnode", C.c_int), ("old_coord", _DP), ("old_ncell", C.c_int), ("old_conn", _IP),
                ("old_size", _DP), ("new_nnode", C.c_int), ("new_coord", _DP), ("new_ncell", C.c_int),
                ("new_conn", _IP)]


class DesTransferField(C.Structure):
    _fields_ = [("src", _DP), ("dst", _DP), ("ncomp", C.c_int)]


class DesTransferStats(C.Structure):
    _fields_ = [("nchanged", C.c_int), ("nsample", C.c_int), ("n_on_face", C.c_int), ("n_outside", C.c_int),
                ("search_seconds", C.c_double), ("acm_seconds", C.c_double), ("field_seconds", C.c_double)]


class DesQualityParam(C.Structure):
    """Param.mesh members bad_mesh_quality reads (remeshing.cxx:2752-2866)."""
    _fields_ = [("smallest_size", C.c_double), ("resolution", C.c_double),
                ("max_boundary_distortion", C.c_double), ("min_quality", C.c_double),
                ("xlength", C.c_double), ("ylength", C.c_double), ("zlength", C.c_double),
                ("remeshing_option", C.c_int)]


class DesQuality(C.Structure):
    _fields_ = [("first_small_elem", C.c_int), ("first_bottom_node", C.c_int),
                ("first_side_node", C.c_int), ("worst_elem", C.c_int),
                ("worst_quality", C.c_double), ("bottom_value", C.c_double),
                ("side_value", C.c_double)]


# every symbol include/des_gpu.h declares (tests check the library exports them all)
ABI_SYMBOLS = [
    "des_gpu_create", "des_gpu_create_multi", "des_gpu_partition_info", "des_host_partition", "des_gpu_set_reorder",
    "des_gpu_destroy", "des_gpu_last_error", "des_gpu_upload_mesh",
    "des_gpu_upload_state", "des_gpu_download_state", "des_gpu_upload_elemmarkers",
    "des_gpu_set_compensation_pressure", "des_gpu_step", "des_gpu_compute_dt", "des_gpu_compute_dt_static",
    "des_dt_from_reductions", "des_gpu_sync",
    "des_gpu_update_temperature", "des_gpu_update_strain_rate", "des_gpu_compute_dvoldt",
    "des_gpu_compute_edvoldt", "des_gpu_update_stress", "des_gpu_nmd_stress",
    "des_gpu_update_force", "des_gpu_update_velocity", "des_gpu_apply_vbcs",
    "des_gpu_update_mesh", "des_gpu_update_mesh_part", "des_gpu_set_pt_jump",
    "des_gpu_set_body_force_adjustment", "des_gpu_compute_volume", "des_gpu_compute_mass",
    "des_gpu_rotate_stress", "des_gpu_residual_force", "des_gpu_isostasy_constrain_velocity",
    "des_gpu_launch_count", "des_gpu_enable_kernel_timing", "des_gpu_kernel_time",
    "des_gpu_device_ptr", "des_host_build_support", "des_abi_sizeof",
    "des_gpu_event_record", "des_gpu_event_elapsed_ms",
    "des_gpu_step_segment", "des_gpu_set_stream", "des_gpu_set_halo", "des_gpu_halo_buffers",
    "des_gpu_halo_width", "des_gpu_halo_pack", "des_gpu_halo_unpack",
    "des_gpu_set_interior", "des_gpu_step_segment_part",
    "des_gpu_mesh_quality", "des_quality_verdict", "des_gpu_average_fields",
    "des_gpu_download_averages", "des_gpu_check_nan",
    "des_gpu_download_frame", "des_host_alloc", "des_host_free",
    "des_gpu_build_support", "des_gpu_find_boundary_facets", "des_host_boundary_flags",
    "des_host_boundary_nodes", "des_host_bucket_facets", "des_host_surface_topology",
    "des_remesh_transfer_cells", "des_remesh_transfer_nodes",
    "des_gpu_upload_markers", "des_gpu_phase_changes", "des_gpu_download_elemmarkers",
]

_lib = None


def load_library():
    """Load libdes_b200.so (built in-tree by __graft_entry__.build()).  Fails loudly."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError("%s is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                          "(no CPU fallback exists)" % LIB_PATH)
    L = C.CDLL(LIB_PATH)
    H = C.c_void_p
    L.des_gpu_create.argtypes = [C.POINTER(DesParam), C.c_int, C.POINTER(H)]
    L.des_gpu_create_multi.argtypes = [C.POINTER(DesParam), C.c_int, _IP, C.POINTER(H)]
    L.des_gpu_partition_info.argtypes = [H, _IP, _DP, _DP, _IP, _IP, _IP]
    L.des_gpu_set_reorder.argtypes = [H, C.c_int]
    L.des_host_partition.argtypes = [C.c_int, C.c_int, _IP, _DP, C.c_int, _IP, _IP, _IP, _IP, _IP]
    L.des_gpu_destroy.argtypes = [H]
    L.des_gpu_last_error.argtypes = [H]
    L.des_gpu_last_error.restype = C.c_char_p
    L.des_gpu_upload_mesh.argtypes = [H, C.POINTER(DesMesh)]
    L.des_gpu_upload_state.argtypes = [H, C.POINTER(DesState)]
    L.des_gpu_download_state.argtypes = [H, C.POINTER(DesState)]
    L.des_gpu_upload_elemmarkers.argtypes = [H, _IP]
    L.des_gpu_set_compensation_pressure.argtypes = [H, C.c_double]
    L.des_gpu_step.argtypes = [H, C.c_double, C.c_double, C.c_int]
    L.des_gpu_compute_dt.argtypes = [H, C.POINTER(DesDtReductions)]
    L.des_gpu_compute_dt_static.argtypes = [H, C.POINTER(DesDtReductions)]
    L.des_dt_from_reductions.argtypes = [C.POINTER(DesParam), C.POINTER(DesDtReductions), _DP]
    L.des_dt_from_reductions.restype = C.c_double
    L.des_gpu_sync.argtypes = [H]
    for f in ["update_temperature", "update_stress", "update_velocity", "apply_vbcs", "update_mesh",
              "rotate_stress"]:
        getattr(L, "des_gpu_" + f).argtypes = [H, C.c_double]
    for f in ["update_strain_rate", "compute_dvoldt", "compute_edvoldt", "nmd_stress", "update_force",
              "compute_volume", "compute_mass", "isostasy_constrain_velocity"]:
        getattr(L, "des_gpu_" + f).argtypes = [H]
    L.des_gpu_residual_force.argtypes = [H, _DP]
    L.des_gpu_update_mesh_part.argtypes = [H, C.c_double, C.c_int]
    L.des_gpu_set_pt_jump.argtypes = [H, C.c_int]
    L.des_gpu_set_body_force_adjustment.argtypes = [H, C.c_int]
    L.des_gpu_launch_count.argtypes = [H]
    L.des_gpu_launch_count.restype = C.c_longlong
    L.des_gpu_enable_kernel_timing.argtypes = [H, C.c_int]
    L.des_gpu_kernel_time.argtypes = [H, C.c_int, _DP, C.POINTER(C.c_longlong), C.POINTER(C.c_char_p)]
    L.des_gpu_device_ptr.argtypes = [H, C.c_char_p]
    L.des_gpu_device_ptr.restype = C.c_void_p
    L.des_gpu_event_record.argtypes = [H, C.c_int]
    L.des_gpu_event_elapsed_ms.argtypes = [H, C.c_int, C.c_int, _DP]
    L.des_gpu_step_segment.argtypes = [H, C.c_int, C.c_double, C.c_double]
    L.des_gpu_set_stream.argtypes = [H, C.c_void_p, C.c_int]
    L.des_gpu_set_interior.argtypes = [H, C.c_int]
    L.des_gpu_step_segment_part.argtypes = [H, C.c_int, C.c_int, C.c_double, C.c_double, C.c_void_p]
    L.des_gpu_set_halo.argtypes = [H, C.c_int, _IP, C.c_int, _IP]
    L.des_gpu_halo_buffers.argtypes = [H, C.POINTER(C.c_void_p), C.POINTER(C.c_void_p)]
    L.des_gpu_halo_width.argtypes = [C.c_int]
    L.des_gpu_halo_pack.argtypes = [H, C.c_int]
    L.des_gpu_halo_unpack.argtypes = [H, C.c_int]
    L.des_host_build_support.argtypes = [C.c_int, C.c_int, _IP, _IP, _IP, _IP]
    L.des_abi_sizeof.argtypes = [C.c_int]
    L.des_gpu_mesh_quality.argtypes = [H, C.POINTER(DesQualityParam), C.POINTER(DesQuality)]
    L.des_quality_verdict.argtypes = [C.POINTER(DesQualityParam), C.POINTER(DesQuality),
                                      C.POINTER(C.c_int), _DP]
    _UP = C.POINTER(C.c_uint)
    L.des_gpu_build_support.argtypes = [C.c_int, C.c_int, C.c_int, _IP, _IP, _IP, _IP]
    L.des_gpu_find_boundary_facets.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, _IP, _IP, _IP, _UP,
                                               _IP, _IP, _IP, _IP]
    L.des_host_boundary_flags.argtypes = [C.c_int, C.c_int, _IP, _IP, _UP]
    L.des_host_boundary_nodes.argtypes = [C.c_int, _UP, C.POINTER(C.c_int * DES_NBDRY),
                                          C.POINTER(_IP * DES_NBDRY)]
    L.des_host_bucket_facets.argtypes = [C.c_int, C.c_int, _IP, _IP, _IP, _IP, C.POINTER(C.c_int * DES_NBDRY),
                                         C.POINTER(C.c_int * (DES_NBDRY + 1)), _IP, _IP]
    L.des_host_surface_topology.argtypes = [C.c_int, C.c_int, _IP, _DP, _IP, _IP, C.c_int, _IP, _IP,
                                            C.c_int, _IP, _IP, _IP, _IP, _IP]
    L.des_remesh_transfer_cells.argtypes = [C.c_int, C.POINTER(DesTransferMeshes), C.c_int, C.c_int,
                                            C.POINTER(DesTransferField), _IP, C.POINTER(DesTransferStats)]
    L.des_remesh_transfer_nodes.argtypes = [C.c_int, C.POINTER(DesTransferMeshes), _IP, _IP, C.c_int,
                                            C.POINTER(DesTransferField), _IP, _DP, C.POINTER(DesTransferStats)]
    L.des_gpu_upload_markers.argtypes = [H, C.c_int, C.c_int, _IP, _DP, _IP]
    L.des_gpu_phase_changes.argtypes = [H, C.c_int, _IP, C.c_int, _IP, _IP, _IP, _IP]
    L.des_gpu_download_elemmarkers.argtypes = [H, _IP]
    L.des_gpu_check_nan.argtypes = [H, C.POINTER(C.c_int)]
    L.des_gpu_average_fields.argtypes = [H, C.c_int]
    L.des_gpu_download_averages.argtypes = [H, _DP, _DP, _DP, _DP]
    L.des_gpu_download_frame.argtypes = [H, C.POINTER(DesFrame)]
    L.des_host_alloc.argtypes = [C.c_size_t, C.POINTER(C.c_void_p)]
    L.des_host_free.argtypes = [C.c_void_p]
    for i, st in enumerate((DesParam, DesMesh, DesState, DesDtReductions, DesQualityParam, DesQuality,
                            DesFrame, DesTransferMeshes, DesTransferField, DesTransferStats)):
        if L.des_abi_sizeof(i) != C.sizeof(st):
            raise ImportError("ABI mismatch: %s is %d bytes in the library, %d in core.py"
                              % (st.__name__, L.des_abi_sizeof(i), C.sizeof(st)))
    _lib = L
    return L


def host_build_support(nnode, conn):
    """create_support (mesh.cxx:3287-3329) through the library's host helper: conn is the
    SoA [4, nelem] int32 connectivity; returns (idx, arr, lidx)."""
    L = load_library()
    conn = np.ascontiguousarray(conn, dtype=np.int32)
    E = conn.size // 4
    idx = np.empty(nnode + 1, dtype=np.int32)
    arr = np.empty(4 * E, dtype=np.int32)
    lidx = np.empty(4 * E, dtype=np.int32)
    rc = L.des_host_build_support(nnode, E, _ptr(conn, C.c_int), _ptr(idx, C.c_int),
                                  _ptr(arr, C.c_int), _ptr(lidx, C.c_int))
    if rc:
        raise DesError(rc, "des_host_build_support")
    return idx, arr, lidx


def host_partition(coord, conn, nparts):
    """The multi-device handle's partitioner on the host (des_host_partition): returns
    (elem_part [E], node_owner [N], local_elems, local_nodes, ghost_nodes [nparts each])."""
    L = load_library()
    coord = np.ascontiguousarray(coord, dtype=np.float64)
    conn = np.ascontiguousarray(conn, dtype=np.int32)
    N, E = coord.shape[1], conn.shape[1]
    part = np.empty(E, dtype=np.int32); owner = np.empty(N, dtype=np.int32)
    le = np.empty(nparts, dtype=np.int32); ln = np.empty(nparts, dtype=np.int32); gh = np.empty(nparts, dtype=np.int32)
    rc = L.des_host_partition(N, E, _ptr(conn, C.c_int), _ptr(coord, C.c_double), nparts, _ptr(part, C.c_int),
                              _ptr(owner, C.c_int), _ptr(le, C.c_int), _ptr(ln, C.c_int), _ptr(gh, C.c_int))
    if rc:
        raise DesError(rc, "des_host_partition")
    return part, owner, le, ln, gh


def _ck_plain(rc, what):
    if rc:
        raise DesError(rc, what)


def host_boundary_flags(nnode, segment, segflag):
    """create_boundary_flags (mesh.cxx:2824-2851): segment is segment_t SoA [3*nseg]."""
    L = load_library()
    segment = np.ascontiguousarray(segment, dtype=np.int32).reshape(-1)
    segflag = np.ascontiguousarray(segflag, dtype=np.int32).reshape(-1)
    bc = np.zeros(nnode, dtype=np.uint32)
    _ck_plain(L.des_host_boundary_flags(nnode, segflag.size, _ptr(segment, C.c_int), _ptr(segflag, C.c_int),
                                        _ptr(bc, C.c_uint)), "des_host_boundary_flags")
    return bc


def host_boundary_nodes(bcflag):
    """create_boundary_nodes (mesh.cxx:2854-2880): list of ten int32 arrays."""
    L = load_library()
    bcflag = np.ascontiguousarray(bcflag, dtype=np.uint32)
    cnt = (C.c_int * DES_NBDRY)()
    _ck_plain(L.des_host_boundary_nodes(bcflag.size, _ptr(bcflag, C.c_uint), C.byref(cnt), None),
              "des_host_boundary_nodes")
    out = [np.empty(cnt[j], dtype=np.int32) for j in range(DES_NBDRY)]
    ptrs = (_IP * DES_NBDRY)(*[_ptr(a, C.c_int) for a in out])
    _ck_plain(L.des_host_boundary_nodes(bcflag.size, _ptr(bcflag, C.c_uint), C.byref(cnt), C.byref(ptrs)),
              "des_host_boundary_nodes")
    return out


def host_surface_topology(nnode, bnodes_z1, coord_x, support_idx, support_arr, conn_surface):
    """create_top_elems + create_surface_info integer tables (mesh.cxx:2882-3098)."""
    L = load_library()
    bz = np.ascontiguousarray(bnodes_z1, dtype=np.int32)
    cx = np.ascontiguousarray(coord_x, dtype=np.float64)
    cs = np.ascontiguousarray(conn_surface, dtype=np.int32).reshape(-1)
    etop = cs.size // 4
    ntop = bz.size
    top_nodes = np.empty(ntop, dtype=np.int32)
    cap = int(support_arr.size // 4)
    top_elems = np.empty(cap, dtype=np.int32)
    n_te = C.c_int()
    ean = np.empty(3 * etop, dtype=np.int32)
    sidx = np.empty(ntop + 1, dtype=np.int32)
    sarr = np.empty(3 * etop, dtype=np.int32)
    _ck_plain(L.des_host_surface_topology(nnode, ntop, _ptr(bz, C.c_int), _ptr(cx, C.c_double),
                                          _ptr(support_idx, C.c_int), _ptr(support_arr, C.c_int),
                                          etop, _ptr(cs, C.c_int), _ptr(top_nodes, C.c_int), cap,
                                          C.byref(n_te), _ptr(top_elems, C.c_int), _ptr(ean, C.c_int),
                                          _ptr(sidx, C.c_int), _ptr(sarr, C.c_int)), "des_host_surface_topology")
    return dict(top_nodes=top_nodes, top_elems=top_elems[:n_te.value].copy(), elem_and_nodes=ean,
                support_surf_idx=sidx, support_surf_arr=sarr)


def build_topology(coord, conn, segment, segflag, device=0):
    """All integer tables of DesMesh from connectivity + boundary segments (the reference's
    create_boundary_flags/nodes/facets, create_support, create_top_elems, create_surface_info:
    mesh.cxx:2824-3329), bit-identical to the reference's.  The two O(nelem) builders run on
    the GPU.  coord [3, N], conn [4, E], segment [3, nseg] (all SoA), segflag [nseg]."""
    L = load_library()
    coord = np.ascontiguousarray(coord, dtype=np.float64)
    conn = np.ascontiguousarray(conn, dtype=np.int32)
    segment = np.ascontiguousarray(segment, dtype=np.int32)
    segflag = np.ascontiguousarray(segflag, dtype=np.int32).reshape(-1)
    N, E, S = coord.shape[1], conn.shape[1], segflag.size
    m = {"nnode": N, "nelem": E, "connectivity": conn}
    m["bcflag"] = host_boundary_flags(N, segment, segflag)
    idx = np.empty(N + 1, dtype=np.int32); arr = np.empty(4 * E, dtype=np.int32); lidx = np.empty(4 * E, dtype=np.int32)
    _ck_plain(L.des_gpu_build_support(device, N, E, _ptr(conn, C.c_int), _ptr(idx, C.c_int), _ptr(arr, C.c_int),
                                      _ptr(lidx, C.c_int)), "des_gpu_build_support")
    m["support_idx"], m["support_arr"], m["support_lidx"] = idx, arr, lidx
    se = np.empty(S, dtype=np.int32); sf = np.empty(S, dtype=np.int32)
    _ck_plain(L.des_gpu_find_boundary_facets(device, N, E, S, _ptr(conn, C.c_int), _ptr(segment, C.c_int),
                                             _ptr(segflag, C.c_int), _ptr(m["bcflag"], C.c_uint),
                                             _ptr(idx, C.c_int), _ptr(arr, C.c_int), _ptr(se, C.c_int),
                                             _ptr(sf, C.c_int)), "des_gpu_find_boundary_facets")
    nb = (C.c_int * DES_NBDRY)(); off = (C.c_int * (DES_NBDRY + 1))()
    pairs = np.empty(2 * max(S, 1), dtype=np.int32)
    etop_max = int((segflag == 32).sum())
    cs = np.zeros(4 * etop_max, dtype=np.int32)
    _ck_plain(L.des_host_bucket_facets(E, S, _ptr(segflag, C.c_int), _ptr(se, C.c_int), _ptr(sf, C.c_int),
                                       _ptr(conn, C.c_int), C.byref(nb), C.byref(off), _ptr(pairs, C.c_int),
                                       _ptr(cs, C.c_int)), "des_host_bucket_facets")
    for k in range(DES_NBDRY):
        m["bfacets%d" % k] = pairs[2 * off[k]:2 * off[k + 1]].copy()
    m["connectivity_surface"] = cs.reshape(4, -1)
    bn = host_boundary_nodes(m["bcflag"])
    for k in range(DES_NBDRY):
        m["bnodes%d" % k] = bn[k]
    m.update(host_surface_topology(N, bn[5], coord[0], idx, arr, cs))
    m["seg_elem"], m["seg_facet"] = se, sf
    return m


def _transfer_args(old_coord, old_conn, old_size, new_coord, new_conn, fields, n_old, n_new):
    """ctypes views of the two meshes and the field list; returns (meshes, field array, outputs, keep-alive)."""
    oc = np.ascontiguousarray(old_coord, dtype=np.float64).reshape(-1)
    ocn = np.ascontiguousarray(old_conn, dtype=np.int32).reshape(-1)
    osz = np.ascontiguousarray(old_size, dtype=np.float64).reshape(-1)
    nc = np.ascontiguousarray(new_coord, dtype=np.float64).reshape(-1)
    keep = [oc, ocn, osz, nc]
    m = DesTransferMeshes(oc.size // 3, _ptr(oc, C.c_double), ocn.size // 4, _ptr(ocn, C.c_int),
                          _ptr(osz, C.c_double), nc.size // 3, _ptr(nc, C.c_double), 0, None)
    if osz.size != m.old_ncell:
        raise ValueError("old_size has %d entries for %d old cells" % (osz.size, m.old_ncell))
    if new_conn is not None:
        ncn = np.ascontiguousarray(new_conn, dtype=np.int32).reshape(-1)
        keep.append(ncn)
        m.new_ncell, m.new_conn = ncn.size // 4, _ptr(ncn, C.c_int)
    no, nn = n_old(m), n_new(m)
    arr = (DesTransferField * max(len(fields), 1))()
    out = {}
    for i, (name, src) in enumerate(fields.items()):
        src = np.ascontiguousarray(src, dtype=np.float64).reshape(-1)
        if src.size % no:
            raise ValueError("field %s: %d values for %d old items" % (name, src.size, no))
        w = src.size // no
        dst = np.full(w * nn, np.nan)
        keep.append(src)
        out[name] = dst
        arr[i] = DesTransferField(_ptr(src, C.c_double), _ptr(dst, C.c_double), w)
    return m, arr, out, keep


def remesh_transfer_cells(old_coord, old_conn, old_size, new_coord, new_conn, fields, is_surface=False, device=0):
    """nearest_neighbor_interpolation (nn-interpolation.cxx:661-702) on the device: element fields
    (or, is_surface, top-facet fields) from the old mesh to the new one.  Arrays in the reference's
    SoA layouts; fields: {name: [ncomp][old cells]}.  Returns ({name: [ncomp][new cells]},
    is_changed, stats)."""
    L = load_library()
    m, arr, out, keep = _transfer_args(old_coord, old_conn, old_size, new_coord, new_conn, fields,
                                       lambda m: m.old_ncell, lambda m: m.new_ncell)
    changed = np.empty(m.new_ncell, dtype=np.int32)
    st = DesTransferStats()
    _ck_plain(L.des_remesh_transfer_cells(device, C.byref(m), int(bool(is_surface)), len(fields), arr,
                                          _ptr(changed, C.c_int), C.byref(st)), "des_remesh_transfer_cells")
    return out, changed, st


def remesh_transfer_nodes(old_coord, old_conn, old_size, old_support_idx, old_support_arr, new_coord, fields,
                          device=0):
    """barycentric_node_interpolation (brc-interpolation.cxx:345-423) on the device: nodal fields from
    the old mesh to the new nodes.  Returns ({name: [ncomp][new nodes]}, el, brc [n][4], stats)."""
    L = load_library()
    m, arr, out, keep = _transfer_args(old_coord, old_conn, old_size, new_coord, None, fields,
                                       lambda m: m.old_nnode, lambda m: m.new_nnode)
    sidx = np.ascontiguousarray(old_support_idx, dtype=np.int32)
    sarr = np.ascontiguousarray(old_support_arr, dtype=np.int32)
    if sidx.size != m.old_nnode + 1 or sarr.size != 4 * m.old_ncell:
        raise ValueError("support lists do not match the old mesh")
    el = np.empty(m.new_nnode, dtype=np.int32)
    brc = np.empty(4 * m.new_nnode)
    st = DesTransferStats()
    _ck_plain(L.des_remesh_transfer_nodes(device, C.byref(m), _ptr(sidx, C.c_int), _ptr(sarr, C.c_int),
                                          len(fields), arr, _ptr(el, C.c_int), _ptr(brc, C.c_double),
                                          C.byref(st)), "des_remesh_transfer_nodes")
    return out, el, brc.reshape(-1, 4), st


def make_param(**kw):
    """DesParam from keyword scalars / per-material sequences, with input.cxx's defaults."""
    d = dict(gravity=10.0, inertial_scaling=1e5, damping_factor=0.8, surface_diffusivity=1e-6,
             surf_base_level=0.0, dt_fraction=1.0, fixed_dt=0.0, characteristic_speed=0.0,
             damping_option=1, is_quasi_static=1, has_thermal_diffusion=1, is_using_mixed_stress=1,
             surface_process_option=0, ref_pressure_option=0, has_moving_mesh=1,
             has_body_force_adjustment=0, has_winkler_foundation=1, has_water_loading=1,
             has_elastic_foundation=0, winkler_delta_rho=0.0, elastic_foundation_constant=0.0,
             sea_water_density=1030.0, surface_temperature=273.0,
             vbc_val_z1_loading_period=1e300, max_vbc_val=1e-12, zlength=0.0,
             rheol_type=RH_EP, nmat=1, mattype_ref=0, visc_min=1e19, visc_max=1e24,
             tension_max=1e9, therm_diff_max=5e-6)
    p = DesParam()
    d.update({k: v for k, v in kw.items() if k in d})
    for k, v in d.items():
        setattr(p, k, v)
    nmat = d["nmat"]
    for name in MAT_ARRAYS:
        if name in kw:
            a = np.broadcast_to(np.asarray(kw[name], dtype=np.float64), (nmat,))
            arr = getattr(p, name)
            for m in range(nmat):
                arr[m] = float(a[m])
    for name, n in (("vbc_types", DES_NBDRY), ("vbc_values", DES_NBDRY), ("vbc_val_lateral", 4),
                    ("stress_bc_types", 6), ("stress_bc_values", 6)):
        if name in kw:
            arr = getattr(p, name)
            for i in range(n):
                arr[i] = kw[name][i]
    unknown = set(kw) - set(d) - set(MAT_ARRAYS) - {"vbc_types", "vbc_values", "vbc_val_lateral",
                                                    "stress_bc_types", "stress_bc_values"}
    if unknown:
        raise TypeError("unknown DesParam fields: %s" % sorted(unknown))
    return p


def _ptr(a, ct):
    return a.ctypes.data_as(C.POINTER(ct)) if a is not None and a.size else C.POINTER(ct)()


class DesGpu:
    """One device-resident model.  Host arrays use the reference's SoA layout
    (Array2D a[d*n + i], array2d.hpp:410-425)."""

    def __init__(self, param, device=0, devices=None):
        """devices: a list of device indices -> one handle over several devices (or several parts
        on one device when an index repeats): des_gpu_create_multi."""
        self.lib = load_library()
        self.param = param
        self.h = C.c_void_p()
        self.devices = list(devices) if devices is not None else None
        if self.devices is not None:
            arr = (C.c_int * len(self.devices))(*self.devices)
            rc = self.lib.des_gpu_create_multi(C.byref(param), len(self.devices), arr, C.byref(self.h))
        else:
            rc = self.lib.des_gpu_create(C.byref(param), device, C.byref(self.h))
        if rc:
            msg = self.lib.des_gpu_last_error(self.h).decode() if self.h else "no handle"
            if self.h:
                self.lib.des_gpu_destroy(self.h)
                self.h = C.c_void_p()
            raise DesError(rc, msg)
        self.nnode = self.nelem = self.etop = 0
        self._keep = []

    def _ck(self, rc):
        if rc:
            raise DesError(rc, self.lib.des_gpu_last_error(self.h).decode())

    def close(self):
        if self.h:
            self.lib.des_gpu_destroy(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- data movement ----------------------------------------------------------------
    def upload_mesh(self, m):
        """`m`: dict with the reference's topology arrays (see DesMesh in des_gpu.h)."""
        keep = {}

        def arr(name, dt):
            a = m.get(name)
            if a is None:
                return None
            a = np.ascontiguousarray(a, dtype=dt)
            keep[name] = a
            return a

        dm = DesMesh()
        dm.nnode, dm.nelem = int(m["nnode"]), int(m["nelem"])
        dm.connectivity = _ptr(arr("connectivity", np.int32), C.c_int)
        dm.bcflag = _ptr(arr("bcflag", np.uint32), C.c_uint)
        dm.support_idx = _ptr(arr("support_idx", np.int32), C.c_int)
        dm.support_arr = _ptr(arr("support_arr", np.int32), C.c_int)
        dm.support_lidx = _ptr(arr("support_lidx", np.int32), C.c_int)
        for i in range(DES_NBDRY):
            bf = arr("bfacets%d" % i, np.int32)
            bn = arr("bnodes%d" % i, np.int32)
            dm.nbfacets[i] = 0 if bf is None else bf.size // 2
            dm.bfacets[i] = _ptr(bf, C.c_int)
            dm.nbnodes[i] = 0 if bn is None else bn.size
            dm.bnodes[i] = _ptr(bn, C.c_int)
        dm.bnormals = _ptr(arr("bnormals", np.float64), C.c_double)
        ev = arr("edge_vec", np.float64)
        dm.n_edge_vec = 0 if ev is None else ev.size
        dm.edge_vec = _ptr(ev, C.c_double)
        dm.edge_slot = _ptr(arr("edge_slot", np.int32), C.c_int)
        te = arr("top_elems", np.int32)
        dm.ntop_elems = 0 if te is None else te.size
        dm.top_elems = _ptr(te, C.c_int)
        tn = arr("top_nodes", np.int32)
        dm.ntop = 0 if tn is None else tn.size
        dm.top_nodes = _ptr(tn, C.c_int)
        dm.etop = dm.nbfacets[5]
        dm.support_surf_idx = _ptr(arr("support_surf_idx", np.int32), C.c_int)
        dm.support_surf_arr = _ptr(arr("support_surf_arr", np.int32), C.c_int)
        dm.elem_and_nodes = _ptr(arr("elem_and_nodes", np.int32), C.c_int)
        dm.connectivity_surface = _ptr(arr("connectivity_surface", np.int32), C.c_int)
        dm.coord = _ptr(arr("coord", np.float64), C.c_double)      # only a multi-device handle reads it
        self._ck(self.lib.des_gpu_upload_mesh(self.h, C.byref(dm)))
        self.nnode, self.nelem, self.etop = dm.nnode, dm.nelem, dm.etop

    def _sizes(self):
        N, E = self.nnode, self.nelem
        s = {f: E for f in STATE_FIELDS}
        for f in ("coord", "vel", "coord0", "force"):
            s[f] = 3 * N
        for f in ("temperature", "volume_n", "mass", "tmass", "ntmp", "dhacc"):
            s[f] = N
        for f in ("stress", "strain", "strain_rate"):
            s[f] = 6 * E
        s["edvacc_surf"] = self.etop
        return s

    def upload_state(self, **fields):
        st = DesState()
        sizes = self._sizes()
        keep = []
        for k, v in fields.items():
            a = np.ascontiguousarray(v, dtype=np.float64).reshape(-1)
            if a.size != sizes[k]:
                raise ValueError("%s: expected %d doubles, got %d" % (k, sizes[k], a.size))
            keep.append(a)
            setattr(st, k, _ptr(a, C.c_double))
        self._ck(self.lib.des_gpu_upload_state(self.h, C.byref(st)))

    def download_state(self, *names):
        st = DesState()
        sizes = self._sizes()
        out = {}
        for k in names:
            out[k] = np.empty(sizes[k], dtype=np.float64)
            setattr(st, k, _ptr(out[k], C.c_double) if out[k].size else _DP())
        self._ck(self.lib.des_gpu_download_state(self.h, C.byref(st)))
        return out

    def download_into(self, **bufs):
        """download_state into caller-owned (e.g. pinned) float64 arrays."""
        st = DesState()
        sizes = self._sizes()
        for k, a in bufs.items():
            if a.dtype != np.float64 or not a.flags["C_CONTIGUOUS"] or a.size != sizes[k]:
                raise ValueError("%s: need a contiguous float64 array of %d" % (k, sizes[k]))
            setattr(st, k, _ptr(a, C.c_double))
        self._ck(self.lib.des_gpu_download_state(self.h, C.byref(st)))

    def upload_elemmarkers(self, em):
        a = np.ascontiguousarray(em, dtype=np.int32).reshape(-1)
        if a.size != self.nelem * self.param.nmat:
            raise ValueError("elemmarkers: expected nelem*nmat ints")
        self._ck(self.lib.des_gpu_upload_elemmarkers(self.h, _ptr(a, C.c_int)))

    def upload_markers(self, elem, eta, mattype):
        """markersets[0]: element, shape-function weights [4][n] and rock type of every marker."""
        elem = np.ascontiguousarray(elem, dtype=np.int32)
        mattype = np.ascontiguousarray(mattype, dtype=np.int32)
        eta = np.ascontiguousarray(eta, dtype=np.float64).reshape(4, -1)
        self._markers = elem.size
        self._ck(self.lib.des_gpu_upload_markers(self.h, elem.size, eta.shape[1], _ptr(elem, C.c_int),
                                               _ptr(eta, C.c_double), _ptr(mattype, C.c_int)))

    def phase_changes(self, option=1, hydrous_count=None, cap=None):
        """phase_changes (phasechanges.cxx:109) on the resident markers: (changed marker ids,
        their new rock types, hyd_inc of every marker)."""
        n = self._markers
        cap = n if cap is None else cap
        nch = C.c_int(0)
        cm = np.empty(max(cap, 1), dtype=np.int32); ct = np.empty(max(cap, 1), dtype=np.int32)
        hyd = np.empty(max(n, 1), dtype=np.int32)
        hc = None if hydrous_count is None else np.ascontiguousarray(hydrous_count, dtype=np.int32)
        self._ck(self.lib.des_gpu_phase_changes(self.h, option, None if hc is None else _ptr(hc, C.c_int), cap,
                                              C.byref(nch), _ptr(cm, C.c_int), _ptr(ct, C.c_int), _ptr(hyd, C.c_int)))
        k = min(nch.value, cap)
        return cm[:k].copy(), ct[:k].copy(), hyd[:n], nch.value

    def download_elemmarkers(self):
        out = np.empty(self.nelem * self.param.nmat, dtype=np.int32)
        self._ck(self.lib.des_gpu_download_elemmarkers(self.h, _ptr(out, C.c_int)))
        return out.reshape(self.nelem, -1)

    def set_reorder(self, mode):
        """0 caller's order, 1 elements by smallest node, 2 Morton (nodes + elements; the mesh
        dict then needs "coord"); effective at the next upload_mesh."""
        self._ck(self.lib.des_gpu_set_reorder(self.h, int(mode)))

    def partition_info(self):
        n = C.c_int(); ps = C.c_double(); us = C.c_double()
        self._ck(self.lib.des_gpu_partition_info(self.h, C.byref(n), C.byref(ps), C.byref(us), None, None, None))
        le = (C.c_int * n.value)(); ln = (C.c_int * n.value)(); gh = (C.c_int * n.value)()
        self._ck(self.lib.des_gpu_partition_info(self.h, C.byref(n), C.byref(ps), C.byref(us), le, ln, gh))
        return dict(nparts=n.value, partition_seconds=ps.value, upload_seconds=us.value,
                    local_elems=list(le), local_nodes=list(ln), ghost_nodes=list(gh))

    def set_compensation_pressure(self, p):
        self._ck(self.lib.des_gpu_set_compensation_pressure(self.h, float(p)))

    # ---- stepping -----------------------------------------------------------------------
    def step(self, dt, time=0.0, steps=0):
        self._ck(self.lib.des_gpu_step(self.h, float(dt), float(time), int(steps)))

    def step_segment(self, seg, dt, time=0.0):
        self._ck(self.lib.des_gpu_step_segment(self.h, int(seg), float(dt), float(time)))

    def set_interior(self, n_interior):
        self._ck(self.lib.des_gpu_set_interior(self.h, int(n_interior)))

    def step_segment_part(self, seg, part, dt, time=0.0, stream_ptr=None):
        self._ck(self.lib.des_gpu_step_segment_part(self.h, seg, part, dt, time, stream_ptr))

    def set_stream(self, stream_ptr):
        """Run on the caller's CUDA stream (an integer cudaStream_t; 0 = legacy default stream)."""
        self._ck(self.lib.des_gpu_set_stream(self.h, C.c_void_p(stream_ptr), 0))

    def set_halo(self, send_nodes, recv_nodes):
        s = np.ascontiguousarray(send_nodes, dtype=np.int32)
        r = np.ascontiguousarray(recv_nodes, dtype=np.int32)
        self._ck(self.lib.des_gpu_set_halo(self.h, s.size, _ptr(s, C.c_int), r.size, _ptr(r, C.c_int)))

    def halo_buffers(self):
        a, b = C.c_void_p(), C.c_void_p()
        self._ck(self.lib.des_gpu_halo_buffers(self.h, C.byref(a), C.byref(b)))
        return a.value, b.value

    def halo_pack(self, phase):
        self._ck(self.lib.des_gpu_halo_pack(self.h, int(phase)))

    def halo_unpack(self, phase):
        self._ck(self.lib.des_gpu_halo_unpack(self.h, int(phase)))

    def compute_dt_reductions(self):
        r = DesDtReductions()
        self._ck(self.lib.des_gpu_compute_dt(self.h, C.byref(r)))
        return r

    def compute_dt_reductions_static(self):
        """des_gpu_compute_dt_static: the same reductions out of the library's static device
        scratch, no allocation (no counterpart in the reference; include/des_gpu.h)."""
        r = DesDtReductions()
        self._ck(self.lib.des_gpu_compute_dt_static(self.h, C.byref(r)))
        return r

    def compute_dt(self):
        """compute_dt (geometry.cxx:1480-1647): device reductions + host scalar algebra.
        Returns (dt, max_global_vel_mag, reductions)."""
        r = self.compute_dt_reductions()
        vmag = C.c_double()
        dt = self.lib.des_dt_from_reductions(C.byref(self.param), C.byref(r), C.byref(vmag))
        if not dt > 0:
            raise DesError(11, "dt <= 0")
        return dt, vmag.value, r

    def mesh_quality(self, **mesh_param):
        """bad_mesh_quality (remeshing.cxx:2752) on the device: returns (code, index, min_quality,
        DesQuality); min_quality is None when the reference would not have written it."""
        qp = DesQualityParam(**mesh_param)
        q = DesQuality()
        self._ck(self.lib.des_gpu_mesh_quality(self.h, C.byref(qp), C.byref(q)))
        idx = C.c_int(-12345)
        mq = C.c_double(float("nan"))
        code = self.lib.des_quality_verdict(C.byref(qp), C.byref(q), C.byref(idx), C.byref(mq))
        mqv = None if mq.value != mq.value else mq.value
        return code, (None if idx.value == -12345 else idx.value), mqv, q

    def check_nan(self):
        """check_nan (utils.hpp:323): how many NaN entries the arrays the reference scans hold."""
        c = C.c_int()
        self._ck(self.lib.des_gpu_check_nan(self.h, C.byref(c)))
        return c.value

    def average_fields(self, first):
        """Output::average_fields (output.cxx:312) with device-resident accumulators."""
        self._ck(self.lib.des_gpu_average_fields(self.h, 1 if first else 0))

    def download_averages(self):
        """(coord0 [3N], strain0 [6E], stress_avg [6E], delta_plstrain_avg [E]) as the reference lays
        them out."""
        N, E = self.nnode, self.nelem
        out = [np.empty(3 * N), np.empty(6 * E), np.empty(6 * E), np.empty(E)]
        self._ck(self.lib.des_gpu_download_averages(self.h, *[_ptr(a, C.c_double) for a in out]))
        return out

    def download_frame(self, averaged=False, inv_dt=0.0, stress_avg_scale=0.0, names=None):
        """des_gpu_download_frame: Output::_write's arrays (output.cxx:95-245) in the on-disk
        layout, [n][components]; returns {name: array}."""
        f = DesFrame()
        f.averaged = int(bool(averaged)); f.inv_dt = float(inv_dt); f.stress_avg_scale = float(stress_avg_scale)
        out = {}
        for group, n, w in ((FRAME_NODE3, self.nnode, 3), (FRAME_NODE1, self.nnode, 1),
                            (FRAME_ELEM6, self.nelem, 6), (FRAME_ELEM1, self.nelem, 1)):
            for k in group:
                if names is not None and k not in names:
                    continue
                if names is None and not averaged and k in ("velocity_averaged", "stress_averaged"):
                    continue
                out[k] = np.empty((n, w) if w > 1 else n, dtype=np.float64)
                setattr(f, k, _ptr(out[k], C.c_double))
        self._ck(self.lib.des_gpu_download_frame(self.h, C.byref(f)))
        return out

    def sync(self):
        self._ck(self.lib.des_gpu_sync(self.h))

    def run(self, nsteps, dt, time=0.0, steps=0, dt_interval=10):
        """The reference's step loop (dynearthsol.cxx:768-894) for `nsteps` steps:
        step, and compute_dt every `dt_interval` steps.  Returns (dt, time, steps)."""
        for _ in range(nsteps):
            steps += 1
            time += dt
            self.step(dt, time, steps)
            if steps % dt_interval == 0:
                dt = self.compute_dt()[0]
        return dt, time, steps

    # ---- twins ----------------------------------------------------------------------------
    def call(self, fn, *args):
        f = getattr(self.lib, "des_gpu_" + fn)
        self._ck(f(self.h, *[float(a) for a in args]))

    def residual_force(self):
        v = C.c_double()
        self._ck(self.lib.des_gpu_residual_force(self.h, C.byref(v)))
        return v.value

    # ---- instrumentation --------------------------------------------------------------
    def launch_count(self):
        return int(self.lib.des_gpu_launch_count(self.h))

    def enable_kernel_timing(self, on=True):
        self._ck(self.lib.des_gpu_enable_kernel_timing(self.h, int(on)))

    def kernel_times(self):
        out = []
        for i in range(DES_NKCLASS):
            ms, n, nm = C.c_double(), C.c_longlong(), C.c_char_p()
            self._ck(self.lib.des_gpu_kernel_time(self.h, i, C.byref(ms), C.byref(n), C.byref(nm)))
            if n.value:
                out.append((nm.value.decode(), ms.value, n.value))
        return out

    def event_record(self, slot):
        self._ck(self.lib.des_gpu_event_record(self.h, slot))

    def event_elapsed_ms(self, a, b):
        ms = C.c_double()
        self._ck(self.lib.des_gpu_event_elapsed_ms(self.h, a, b, C.byref(ms)))
        return ms.value

    def device_ptr(self, name):
        return self.lib.des_gpu_device_ptr(self.h, name.encode())
