"""Branch-level reference vectors: drive single hot functions of the UNMODIFIED reference
(oracle/_ref/libdesref.so) on hand-made inputs that reach the branches no whole-step config
reaches, and dump inputs + outputs to an .npz.  tests/test_branches_gpu.py feeds the same
inputs to the CUDA twins through the C-ABI and compares.

usage: python tests/refbranches.py SUITE CFG OUT.npz
  vbc       apply_vbcs (bc.cxx:227-659): every lateral type 0..7, z0 types 0..3, z1 types 0..4
            with and without an expired loading period, and the slant block n0..n3 (types
            1/3/11/13, plain and intersecting nodes, edge projection) on synthetic flags
  ql        update_stress (rheology.cxx:703) on stress states whose principal stresses are
            degenerate, so that dsyevh3 falls back to the QL solver (3x3-C/dsyevh3.c:102,117)
  residual  update_force + calculate_residual_force (fields.cxx:609-722)
"""
import os
import sys
import tempfile

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import refdump  # noqa: E402
from refharness import Ref  # noqa: E402


def base_dump(r, out):
    for a in refdump.MESH_ARRAYS:
        out["mesh_" + a] = r.get(a)
    out["elemmarkers"] = r.get("elemmarkers")
    for s in refdump.SCALARS:
        out["p_" + s] = np.float64(r.scalar(s))
    for s in refdump.INTS:
        out["p_" + s] = np.int32(r.int(s))
    for s in refdump.MATV:
        out["p_" + s] = r.matvec(s)
    refdump.snapshot(r, out, "s0_")


# ---- apply_vbcs ---------------------------------------------------------------------------
def vbc_cases(rng):
    """(types[10], values[10], lateral[4], z1_loading_period, time) per case."""
    cases = []
    vals = lambda: rng.uniform(-1e-9, 1e-9, size=10)
    lat = lambda: rng.uniform(-1e-9, 1e-9, size=4)
    # every lateral type on every face (latin-square style), z free
    for shift in range(8):
        t = [(shift + i) % 8 for i in range(4)] + [0, 0, 1, 1, 1, 1]
        cases.append((t, vals(), lat(), 1e300, 1.0))
    # z faces: every (z0, z1) pair, the loading period still running / expired
    for z0 in range(4):
        for z1 in range(5):
            for expired in (False, True):
                t = [1, 6, 0, 3, z0, z1, 1, 1, 1, 1]
                cases.append((t, vals(), lat(), 0.5 if expired else 2.0, 1.0))
    # slant block: each type on each slant boundary, with lateral types 0 and 1 (the two
    # intersection sub-branches, bc.cxx:509-527) and others (neither branch)
    for nt in (1, 3, 11, 13):
        for latt in (0, 1, 3):
            t = [latt, 1 - min(latt, 1), latt, 1, 0, 0, nt, nt, nt, nt]
            cases.append((t, vals(), lat(), 1e300, 1.0))
    cases.append(([0, 1, 1, 0, 2, 1, 1, 11, 3, 13], vals(), lat(), 1e300, 1.0))
    cases.append(([1, 0, 0, 1, 1, 3, 13, 3, 11, 1], vals(), lat(), 1e300, 1.0))
    return cases


def suite_vbc(r, out):
    rng = np.random.default_rng(20261105)
    n = r.int("nnode")
    flag0 = r.get("bcflag")
    # synthetic slant boundaries: bits 6..9 on interior nodes alone ("ordinary"), on nodes of
    # each box face (intersections), and several slant bits at once
    flag = flag0.copy()
    interior = np.nonzero(flag0 == 0)[0]
    for k, ib in enumerate((6, 7, 8, 9)):
        flag[interior[k::9][:12]] |= np.uint32(1 << ib)
        for face in range(6):
            on = np.nonzero(flag0 & (1 << face))[0]
            flag[on[(k + face)::7][:6]] |= np.uint32(1 << ib)
    flag[interior[5::11][:8]] |= np.uint32((1 << 6) | (1 << 8))
    flag[interior[7::13][:8]] |= np.uint32((1 << 7) | (1 << 9) | (1 << 6))
    r.set("bcflag", flag)
    # unit normals with a horizontal part (types 11/13 divide by sqrt(1 - nz^2))
    bn = np.zeros((3, 10))
    for ib in range(10):
        v = rng.normal(size=3)
        v[2] *= 0.5
        bn[:, ib] = v / np.sqrt((v * v).sum())
    r.set("bnormals", bn.reshape(-1))
    # edge table: a unit vector for about two thirds of the (ic < ib) pairs, -1 elsewhere
    slot = np.full(100, -1, dtype=np.int32)
    vecs = []
    for ic in range(6):
        for ib in range(6, 10):
            if (ic + ib) % 3 == 0:
                continue
            v = rng.normal(size=3)
            slot[ic * 10 + ib] = len(vecs)
            vecs.append(v / np.sqrt((v * v).sum()))
    r.set("edge_slot", slot)
    r.set("edge_vec", np.asarray(vecs).reshape(-1))
    out["vbc_bcflag"] = flag
    out["vbc_bnormals"] = bn.reshape(-1)
    out["vbc_edge_slot"] = slot
    out["vbc_edge_vec"] = np.asarray(vecs).reshape(-1)
    cases = vbc_cases(rng)
    out["vbc_n"] = np.int32(len(cases))
    for i, (types, values, lateral, period, time) in enumerate(cases):
        r.set("vbc_types", np.asarray(types, dtype=np.int32))
        r.set("vbc_values", np.asarray(values, dtype=np.float64))
        r.lib.desref_sync_vbc()
        for nm, v in zip(("vbc_val_x0_l", "vbc_val_x1_l", "vbc_val_y0_l", "vbc_val_y1_l"), lateral):
            r.set_scalar(nm, v)
        r.set_scalar("vbc_val_z1_loading_period", period)
        r.set_scalar("time", time)
        vel = rng.uniform(-1e-9, 1e-9, size=3 * n)
        r.set("vel", vel)
        r.call("apply_vbcs")
        out["vbc%d_types" % i] = np.asarray(types, dtype=np.int32)
        out["vbc%d_values" % i] = np.asarray(values, dtype=np.float64)
        out["vbc%d_lateral" % i] = np.asarray(lateral, dtype=np.float64)
        out["vbc%d_period" % i] = np.float64(period)
        out["vbc%d_time" % i] = np.float64(time)
        out["vbc%d_in" % i] = vel
        out["vbc%d_out" % i] = r.get("vel")


# ---- update_stress with degenerate principal stresses ---------------------------------------
def rot(rng):
    q, _ = np.linalg.qr(rng.normal(size=(3, 3)))
    return q


def suite_ql(r, out):
    """Yielding stress states whose eigen-decomposition leaves the Cardano/cross-product path:
    exactly diagonal tensors, two equal principal values (axis-aligned and rotated), nearly
    isotropic ones -- next to generic yielding states as a control."""
    rng = np.random.default_rng(777)
    E = r.int("nelem")
    kinds = np.arange(E) % 6
    s = np.zeros((6, E))
    for e in range(E):
        k = kinds[e]
        big = -rng.uniform(2e8, 6e8)
        if k == 0:      # diagonal, distinct
            p = np.asarray([1.0, rng.uniform(0.3, 0.7), rng.uniform(0.1, 0.25)]) * big
            t = np.diag(p[rng.permutation(3)])
        elif k == 1:    # diagonal, two equal
            a, b = rng.uniform(0.15, 0.3) * big, big
            t = np.diag(np.asarray([a, a, b])[rng.permutation(3)])
        elif k == 2:    # two equal principal values, rotated
            a, b = rng.uniform(0.15, 0.3) * big, big
            q = rot(rng)
            t = q @ np.diag([a, b, a]) @ q.T
        elif k == 3:    # one zero off-diagonal pair (block-diagonal)
            p = np.asarray([1.0, rng.uniform(0.3, 0.7), rng.uniform(0.1, 0.25)])[rng.permutation(3)] * big
            t = np.diag(p)
            t[0, 1] = t[1, 0] = rng.uniform(-0.2, 0.2) * big
        elif k == 4:    # tensile corner: nearly isotropic tension
            a = rng.uniform(1e8, 3e8)
            t = np.eye(3) * a + np.diag(rng.uniform(-1, 1, size=3)) * a * 1e-12
        else:           # generic yielding state
            q = rot(rng)
            t = q @ np.diag(np.asarray([1.0, rng.uniform(0.3, 0.7), rng.uniform(0.1, 0.25)]) * big) @ q.T
        s[:, e] = [t[0, 0], t[1, 1], t[2, 2], t[0, 1], t[0, 2], t[1, 2]]
    # zero strain increment on the shear components of the diagonal kinds keeps the trial diagonal
    sr = rng.uniform(-1e-15, 1e-15, size=(6, E))
    sr[3:, kinds <= 1] = 0.0
    sr[4:, kinds == 3] = 0.0
    r.set("stress", s.reshape(-1))
    r.set("strain_rate", sr.reshape(-1))
    r.set("edvoldt", (sr[0] + sr[1] + sr[2]).copy())
    out["ql_kinds"] = kinds.astype(np.int32)
    for a in ("stress", "strain", "strain_rate", "plstrain", "edvoldt", "volume", "volume_old", "temperature"):
        out["ql_in_" + a] = r.get(a)
    r.call("update_stress")
    for a in ("stress", "strain", "strain_rate", "plstrain", "delta_plstrain", "dpressure"):
        out["ql_out_" + a] = r.get(a)


# ---- residual force -------------------------------------------------------------------------
def suite_residual(r, out):
    r.step(3)
    refdump.snapshot(r, out, "res_")
    r.call("update_force")
    out["res_force"] = r.get("force")
    r.call("calculate_residual_force")
    out["res_l2"] = np.float64(r.lib.desref_l2_residual())


SUITES = {"vbc": suite_vbc, "ql": suite_ql, "residual": suite_residual}

if __name__ == "__main__":
    suite, cfg, path = sys.argv[1:4]
    with tempfile.TemporaryDirectory() as wd:
        r = Ref(cfg, workdir=wd)
        out = {}
        base_dump(r, out)
        SUITES[suite](r, out)
        np.savez_compressed(path, **out)
