"""Run the UNMODIFIED reference (oracle/_ref/libdesref.so) on a cfg and dump inputs, per-function
outputs of the first step and whole-step states to an .npz.  This is the script that generated
tests/golden/*.npz; on a box where libdesref.so is present the GPU parity tests call it live.

usage: python tests/refdump.py CFG OUT.npz [--steps 1,20] [--set name=int ...]
"""
import argparse
import os
import sys
import tempfile
import numpy as np

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from refharness import Ref  # noqa: E402

MESH_ARRAYS = ["connectivity", "bcflag", "support_idx", "support_arr", "support_lidx", "bnormals",
               "edge_vec", "edge_slot", "top_elems", "top_nodes", "support_surf_idx",
               "support_surf_arr", "elem_and_nodes", "connectivity_surface", "vbc_types",
               "vbc_values", "stress_bc_types", "stress_bc_values", "segment", "segflag"] + \
              ["bfacets%d" % i for i in range(10)] + ["bnodes%d" % i for i in range(10)]
STATE_ARRAYS = ["coord", "vel", "coord0", "temperature", "stress", "strain", "strain_rate",
                "plstrain", "delta_plstrain", "volume", "volume_old", "viscosity", "dpressure",
                "edvoldt", "radiogenic_source", "force", "volume_n", "mass", "tmass", "ntmp",
                "dhacc", "edvacc_surf"]
SCALARS = ["dt", "time", "compensation_pressure", "max_vbc_val", "gravity", "inertial_scaling",
           "damping_factor", "dt_fraction", "surface_temperature", "winkler_delta_rho", "zlength",
           "surf_base_level", "sea_water_density", "visc_min", "visc_max", "tension_max",
           "therm_diff_max", "elastic_foundation_constant", "characteristic_speed", "fixed_dt",
           "surf_diff", "vbc_val_x0_l", "vbc_val_x1_l", "vbc_val_y0_l", "vbc_val_y1_l",
           "vbc_val_z1_loading_period", "max_surf_vel"]
INTS = ["nnode", "nelem", "nmat", "rheol_type", "mattype_ref", "damping_option", "is_quasi_static",
        "has_thermal_diffusion", "is_using_mixed_stress", "surface_process_option",
        "ref_pressure_option", "has_moving_mesh", "has_winkler_foundation", "has_water_loading",
        "has_elastic_foundation", "has_body_force_adjustment", "ntop", "etop", "ntop_elems"]
MATV = ["rho0", "alpha", "bulk_modulus", "shear_modulus", "visc_exponent", "visc_coefficient",
        "visc_activation_energy", "visc_activation_volume", "heat_capacity", "therm_cond", "pls0",
        "pls1", "cohesion0", "cohesion1", "friction_angle0", "friction_angle1", "dilation_angle0",
        "dilation_angle1", "porosity"]

# the reference's step, one call at a time (dynearthsol.cxx:776-876), and what each call writes
STEP_CALLS = [
    ("update_temperature", ["temperature"]),
    ("update_strain_rate", ["strain_rate"]),
    ("compute_dvoldt", ["ntmp"]),
    ("compute_edvoldt", ["edvoldt"]),
    ("update_stress", ["stress", "strain", "strain_rate", "plstrain", "delta_plstrain",
                       "dpressure", "viscosity"]),
    ("NMD_stress", ["stress", "ntmp"]),
    ("update_force", ["force"]),
    ("update_velocity", ["vel"]),
    ("apply_vbcs", ["vel"]),
    ("update_mesh", ["coord", "volume", "volume_old", "volume_n", "mass", "tmass", "stress",
                     "strain", "strain_rate", "plstrain", "dhacc", "edvacc_surf"]),
    ("rotate_stress", ["stress", "strain"]),
]


MESH_PARAMS = ["smallest_size", "resolution", "max_boundary_distortion", "min_quality", "xlength",
               "ylength", "zlength", "remeshing_option"]
# (smallest_size, max_boundary_distortion, min_quality, remeshing_option) overrides for
# bad_mesh_quality at the last dumped state; None = the cfg's own value.  Chosen to reach every
# return of remeshing.cxx:2752-2866: 0, 1 (distortion), 2 (bottom; sides with option 13), 3 (tiny).
QUALITY_CASES = [(None, None, None, None), (None, None, 0.999, None), (None, 1e-12, None, 1),
                 (None, 1e-12, None, 13), (None, 1e9, None, 13), (None, 1e-12, None, 0),
                 (1e3, None, None, None), (None, None, 1e-9, 11)]


def quality_cases(r, out):
    base = [r.mesh_param(k) for k in ("smallest_size", "max_boundary_distortion", "min_quality",
                                      "remeshing_option")]
    rows, res = [], []
    for case in QUALITY_CASES:
        a = [b if c is None else c for b, c in zip(base, case)]
        code, idx, mq = r.bad_mesh_quality(a[0], a[1], a[2], int(a[3]))
        rows.append(a)
        res.append((code, idx, mq))
    out["q_params"] = np.asarray(rows, dtype=np.float64)
    out["q_code"] = np.asarray([x[0] for x in res], dtype=np.int32)
    out["q_index"] = np.asarray([x[1] for x in res], dtype=np.int32)
    out["q_minq"] = np.asarray([x[2] for x in res], dtype=np.float64)
    # the side-node branch alone (remeshing_option 13, :2795-2838): same state with the bottom
    # nodes put back on z = -zlength, so the bottom check passes and the else-if chain decides
    coord = r.get("coord")
    n = r.int("nnode")
    c2 = coord.copy()
    bottom = (r.get("bcflag") & 16) != 0
    c2[2 * n:][bottom] = -r.mesh_param("zlength")
    r.set("coord", c2)
    code, idx, mq = r.bad_mesh_quality(base[0], 1e-12, base[2], 13)
    r.set("coord", coord)
    out["q2_coord"] = c2
    out["q2_result"] = np.asarray([code, idx], dtype=np.int32)


def phase_case(r, out, seed=11):
    """phase_changes (phasechanges.cxx:109-158) on the last dumped mesh: every marker gets a random
    rock type, every node a random temperature, every element a random hydrous-marker count, so
    that each arm of simple_subduction (:10-84) is taken by many markers; inputs and outputs as pc_*."""
    if r.int("phase_change_option") == 0 or r.hydrous_count() is None:
        return
    rng = np.random.default_rng(seed)
    n, ne, nmat = r.int("nnode"), r.int("nelem"), r.int("nmat")
    elem, eta, mt, _ = r.markers()
    mt0 = rng.choice(np.arange(nmat), size=mt.size, p=[0.2, 0.2, 0.2, 0.05, 0.2, 0.05] + [0.1 / (nmat - 6)] * (nmat - 6)).astype(np.int32)
    r.set_marker_mattype(mt0)
    T = rng.uniform(300.0, 1500.0, n)
    r.set("temperature", T)
    hyd = (rng.random(ne) < 0.5).astype(np.int32) * rng.integers(1, 4, ne).astype(np.int32)
    r.hydrous_count(hyd)
    out["pc_elem"], out["pc_eta"], out["pc_mattype0"] = elem, eta, mt0
    out["pc_temperature"], out["pc_coord"], out["pc_hydrous"] = T, r.get("coord"), hyd
    out["pc_elemmarkers0"] = r.get("elemmarkers")
    out["pc_option"] = np.int32(r.int("phase_change_option"))
    r.call("phase_changes")
    _, _, mt1, tmp = r.markers()
    out["pc_mattype1"], out["pc_tmp"] = mt1, tmp
    out["pc_elemmarkers1"] = r.get("elemmarkers")


def snapshot(r, out, prefix):
    for a in STATE_ARRAYS:
        out[prefix + a] = r.get(a)
    out[prefix + "dt"] = np.float64(r.scalar("dt"))
    out[prefix + "time"] = np.float64(r.scalar("time"))
    out[prefix + "max_surf_vel"] = np.float64(r.scalar("max_surf_vel"))


def dump(cfg, out_path, steps=(1,), sets=(), by_function=True, perturb_at=None):
    with tempfile.TemporaryDirectory() as wd:
        r = Ref(cfg, workdir=wd)
        for name, val in sets:
            r.set_int(name, val)
        out = {}
        for a in MESH_ARRAYS:
            out["mesh_" + a] = r.get(a)
        out["elemmarkers"] = r.get("elemmarkers")
        for s in SCALARS:
            out["p_" + s] = np.float64(r.scalar(s))
        for s in INTS:
            out["p_" + s] = np.int32(r.int(s))
        for s in MATV:
            out["p_" + s] = r.matvec(s)
        snapshot(r, out, "s0_")
        done = 0
        if perturb_at is not None:
            # conditioning probe: the same run with every stress entry moved by ONE ulp at step
            # `perturb_at`; the distance between its later snapshots and the unperturbed run's is
            # what the reference itself does with a last-bit difference in its input
            by_function = False
            if perturb_at > 0:
                r.step(perturb_at)
                done = perturb_at
            r.set("stress", np.nextafter(r.get("stress"), np.inf))
        if by_function:
            # step 1, call by call (same order as desref_step / the driver)
            r.set_int("steps", 1)
            r.set_scalar("time", r.scalar("time") + r.scalar("dt"))
            thermal = r.int("has_thermal_diffusion")
            mixed = r.int("is_using_mixed_stress")
            for fn, writes in STEP_CALLS:
                if fn == "update_temperature" and not thermal:
                    continue
                if fn == "NMD_stress" and not mixed:
                    continue
                if fn == "rotate_stress" and not (r.int("rheol_type") & 1):
                    continue
                if fn == "update_mesh" and not r.int("has_moving_mesh"):
                    continue
                r.call(fn)
                for a in writes:
                    out["f_%s_%s" % (fn, a)] = r.get(a)
            done = 1
            if 1 in steps:
                snapshot(r, out, "s1_")
        for k in sorted(steps):
            if k <= done:
                continue
            r.step(k - done)
            done = k
            snapshot(r, out, "s%d_" % k)
        for s in MESH_PARAMS:
            out["m_" + s] = np.float64(r.mesh_param(s))
        quality_cases(r, out)
        phase_case(r, out)
        out["steps"] = np.asarray(sorted(steps), dtype=np.int32)
        np.savez_compressed(out_path, **out)


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("cfg")
    ap.add_argument("out")
    ap.add_argument("--steps", default="1")
    ap.add_argument("--set", action="append", default=[])
    ap.add_argument("--perturb-at", type=int, default=None,
                    help="move the stress by one ulp at this step (0 = initial state) before going on")
    a = ap.parse_args()
    dump(a.cfg, a.out, steps=[int(s) for s in a.steps.split(",")],
         sets=[(s.split("=")[0], int(s.split("=")[1])) for s in a.set], perturb_at=a.perturb_at)
