"""Build a dynearthsol_b200.DesGpu model from a reference dump (tests/refdump.py .npz)."""
import numpy as np
from dynearthsol_b200 import core

MESH_KEYS = ["connectivity", "bcflag", "support_idx", "support_arr", "support_lidx", "bnormals",
             "edge_vec", "edge_slot", "top_elems", "top_nodes", "support_surf_idx",
             "support_surf_arr", "elem_and_nodes", "connectivity_surface"] + \
            ["bfacets%d" % i for i in range(10)] + ["bnodes%d" % i for i in range(10)]

UPLOAD = ["coord", "vel", "coord0", "temperature", "stress", "strain", "strain_rate", "plstrain",
          "delta_plstrain", "volume", "volume_old", "viscosity", "dpressure", "edvoldt",
          "radiogenic_source", "force", "volume_n", "mass", "tmass", "ntmp", "dhacc", "edvacc_surf"]


def param_from_dump(d, **override):
    kw = {}
    for k in ["gravity", "inertial_scaling", "damping_factor", "surf_base_level", "dt_fraction",
              "fixed_dt", "characteristic_speed", "winkler_delta_rho", "elastic_foundation_constant",
              "sea_water_density", "surface_temperature", "vbc_val_z1_loading_period", "max_vbc_val",
              "zlength", "visc_min", "visc_max", "tension_max", "therm_diff_max"]:
        kw[k] = float(d["p_" + k])
    kw["surface_diffusivity"] = float(d["p_surf_diff"])
    for k in ["damping_option", "is_quasi_static", "has_thermal_diffusion", "is_using_mixed_stress",
              "surface_process_option", "ref_pressure_option", "has_moving_mesh",
              "has_body_force_adjustment", "has_winkler_foundation", "has_water_loading",
              "has_elastic_foundation", "rheol_type", "nmat", "mattype_ref"]:
        kw[k] = int(d["p_" + k])
    for k in core.MAT_ARRAYS:
        kw[k] = np.asarray(d["p_" + k])
    kw["vbc_types"] = [int(x) for x in d["mesh_vbc_types"]]
    kw["vbc_values"] = [float(x) for x in d["mesh_vbc_values"]]
    kw["vbc_val_lateral"] = [float(d["p_vbc_val_%s_l" % s]) for s in ("x0", "x1", "y0", "y1")]
    kw["stress_bc_types"] = [int(x) for x in d["mesh_stress_bc_types"]]
    if "mesh_stress_bc_values" in d:
        kw["stress_bc_values"] = [float(x) for x in d["mesh_stress_bc_values"]]
    kw.update(override)
    return core.make_param(**kw)


def mesh_from_dump(d):
    m = {"nnode": int(d["p_nnode"]), "nelem": int(d["p_nelem"])}
    for k in MESH_KEYS:
        m[k] = d["mesh_" + k]
    if "s0_coord" in d:
        m["coord"] = d["s0_coord"]      # read by a multi-device handle (RCB) and by the Morton storage order
    return m


def model_from_dump(d, prefix="s0_", device=0, mesh=None, reorder=None, **override):
    """Create the device model and upload mesh + the state snapshot `prefix`."""
    g = core.DesGpu(param_from_dump(d, **override), device)
    if reorder is not None:
        g.set_reorder(reorder)
    g.upload_mesh(mesh if mesh is not None else mesh_from_dump(d))
    g.upload_elemmarkers(d["elemmarkers"])
    g.set_compensation_pressure(float(d["p_compensation_pressure"]))
    g.upload_state(**{k: d[prefix + k] for k in UPLOAD})
    return g


def rel_err(a, b):
    """max |a-b| / max |b| (the measure benchmarks-cores/compare.py uses, compare.py:149-169)."""
    a = np.asarray(a, dtype=np.float64); b = np.asarray(b, dtype=np.float64)
    if a.size == 0:
        return 0.0
    scale = np.abs(b).max()
    err = np.abs(a - b).max()
    return 0.0 if err == 0 else err / (scale if scale > 0 else 1.0)
