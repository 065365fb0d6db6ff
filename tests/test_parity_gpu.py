"""GPU parity: the CUDA path, called through the C-ABI, against the UNMODIFIED reference on
the same inputs.  FP64 tolerance from BASELINE.json's north_star: 1e-10 relative after one
step (measured as max|a-b| / max|b| per field, the compare.py measure)."""
import numpy as np
import pytest

import refmodel
from conftest import ref_dump
from refdump import STEP_CALLS

pytestmark = pytest.mark.gpu

TOL = 1e-10
FIELDS = ["coord", "vel", "temperature", "stress", "strain", "strain_rate", "plstrain",
          "delta_plstrain", "volume", "volume_old", "force", "volume_n", "mass", "tmass",
          "dhacc", "edvacc_surf"]

# box-ep-100k: 103 680 tets -- many blocks per kernel, and more elements than the L2 prefetch
# distance of the element kernels (65 536), so that path runs too
CASES = ["box-ep-tiny", "box-ep", "box-evp", "box-maxwell", "box-elastic-nomix", "box-ep-100k"]


def compare(got, d, prefix, fields, tol=TOL, floor=None):
    """floor: {field: the reference's own change under a one-ulp input perturbation}
    (conftest.ref_conditioning); a field may differ by 8x that where it exceeds tol."""
    bad = []
    report = []
    for k in fields:
        if prefix + k not in d:
            continue
        err = refmodel.rel_err(got[k], d[prefix + k])
        exact = np.array_equal(got[k], d[prefix + k])
        report.append("%s:%s" % (k, "bit-exact" if exact else "%.2e" % err))
        lim = max(tol, 8 * floor.get(k, 0.0)) if floor else tol
        if not err < lim:
            bad.append((k, err, lim))
    assert not bad, "fields beyond %g: %s   [all: %s]" % (tol, bad, " ".join(report))
    return report


@pytest.mark.parametrize("case", CASES)
def test_one_step(case):
    d = ref_dump(case, [1, 20])
    g = refmodel.model_from_dump(d)
    dt = float(d["s0_dt"])
    g.step(dt, dt, 1)
    got = g.download_state(*FIELDS)
    compare(got, d, "s1_", FIELDS)
    g.close()


TWIN = {"update_temperature": ("update_temperature", "dt"), "update_strain_rate": ("update_strain_rate",),
        "compute_dvoldt": ("compute_dvoldt",), "compute_edvoldt": ("compute_edvoldt",),
        "update_stress": ("update_stress", "dt"), "NMD_stress": ("nmd_stress",),
        "update_force": ("update_force",), "update_velocity": ("update_velocity", "dt"),
        "apply_vbcs": ("apply_vbcs", "time"), "update_mesh": ("update_mesh", "dt"),
        "rotate_stress": ("rotate_stress", "dt")}


@pytest.mark.parametrize("case", ["box-ep-tiny", "box-evp"])
def test_by_function(case):
    """The fine-grained twins, in the reference's call order, each compared with what the
    reference function wrote (per-function rows of SURVEY.md section 8a)."""
    d = ref_dump(case, [1, 20])
    g = refmodel.model_from_dump(d)
    dt = float(d["s0_dt"])
    vals = {"dt": dt, "time": dt}
    for fn, writes in STEP_CALLS:
        if ("f_%s_%s" % (fn, writes[0])) not in d:
            continue
        t = TWIN[fn]
        g.call(t[0], *[vals[a] for a in t[1:]])
        names = [w for w in writes if w in FIELDS + ["ntmp", "edvoldt", "dpressure", "viscosity"]]
        got = g.download_state(*names)
        compare(got, d, "f_%s_" % fn, names)
    g.close()


@pytest.mark.parametrize("case", CASES)
def test_twenty_steps(case):
    """20 steps with compute_dt every 10 (dynearthsol.cxx:881-894) -- dt itself is compared."""
    d = ref_dump(case, [1, 20])
    g = refmodel.model_from_dump(d)
    dt, time, steps = g.run(20, float(d["s0_dt"]))
    assert abs(dt - float(d["s20_dt"])) <= 1e-12 * abs(dt), (dt, float(d["s20_dt"]))
    got = g.download_state(*FIELDS)
    compare(got, d, "s20_", FIELDS, tol=1e-9)
    g.close()


def test_plastic_yielding():
    """Start from the reference's state at step 250, where ~480 of 2880 elements are on the
    Mohr-Coulomb yield surface (eigen-decomposition, return mapping, rotation back), and
    compare one and twenty steps later."""
    d = ref_dump("box-ep", [250, 251, 270])
    assert (d["s251_delta_plstrain"] > 0).sum() > 100
    g = refmodel.model_from_dump(d, prefix="s250_")
    dt = float(d["s250_dt"])
    time = float(d["s250_time"])
    g.step(dt, time + dt, 251)
    got = g.download_state(*FIELDS)
    compare(got, d, "s251_", FIELDS)
    dt, time, steps = g.run(19, dt, time + dt, 251)
    got = g.download_state(*FIELDS)
    compare(got, d, "s270_", FIELDS, tol=1e-9)
    g.close()


def test_fused_equals_unfused():
    """The deferred-tail schedule (tail(n-1) fused into head(n)) against flushing after
    every step: bit-identical."""
    d = ref_dump("box-ep", [1, 20])
    dt = float(d["s0_dt"])
    a = refmodel.model_from_dump(d)
    b = refmodel.model_from_dump(d)
    t = 0.0
    for i in range(5):
        t += dt
        a.step(dt, t, i + 1)
        b.step(dt, t, i + 1)
        b.download_state("volume")      # forces the flush
    ga, gb = a.download_state(*FIELDS), b.download_state(*FIELDS)
    for k in FIELDS:
        assert np.array_equal(ga[k], gb[k]), k
    a.close(); b.close()


def test_determinism():
    d = ref_dump("box-ep", [1, 20])
    outs = []
    for _ in range(2):
        g = refmodel.model_from_dump(d)
        g.run(12, float(d["s0_dt"]))
        outs.append(g.download_state(*FIELDS))
        g.close()
    for k in FIELDS:
        assert np.array_equal(outs[0][k], outs[1][k]), k


def _quality_kwargs(d, row):
    return dict(smallest_size=float(row[0]), resolution=float(d["m_resolution"]),
                max_boundary_distortion=float(row[1]), min_quality=float(row[2]),
                xlength=float(d["m_xlength"]), ylength=float(d["m_ylength"]),
                zlength=float(d["m_zlength"]), remeshing_option=int(row[3]))


@pytest.mark.parametrize("case", ["box-ep-tiny", "box-ep"])
def test_mesh_quality(case):
    """bad_mesh_quality (remeshing.cxx:2752-2866) with its loops on the device: for every knob
    combination refdump.QUALITY_CASES records -- each return code 0..3 -- the code, the index
    (integer: exact) and min_quality (bit-exact: same arithmetic, host cube root) are the
    reference's, from the state 20 steps in.  The state is uploaded, and in a second pass
    reached by stepping on the device (so the deferred compute_volume is exercised)."""
    d = ref_dump(case, [1, 20])
    for mode in ("uploaded", "stepped"):
        if mode == "uploaded":
            g = refmodel.model_from_dump(d, prefix="s20_")
        else:
            g = refmodel.model_from_dump(d)
            g.run(20, float(d["s0_dt"]))
        for row, code, index, minq in zip(d["q_params"], d["q_code"], d["q_index"], d["q_minq"]):
            got, idx, mq, q = g.mesh_quality(**_quality_kwargs(d, row))
            assert got == int(code), (mode, row, got, code)
            assert (idx if idx is not None else -12345) == int(index), (mode, row, idx, index)
            if np.isnan(minq):
                assert mq is None, (mode, row, mq)
            elif mode == "uploaded":
                assert mq == float(minq), (mode, row, mq, minq)
            else:
                assert abs(mq - float(minq)) <= 1e-10 * float(minq), (mode, row, mq, minq)
        if mode == "uploaded":
            # the side-node branch alone: bottom nodes back on z = -zlength
            g.upload_state(coord=d["q2_coord"])
            row = d["q_params"][0].copy(); row[1] = 1e-12; row[3] = 13
            got, idx, mq, q = g.mesh_quality(**_quality_kwargs(d, row))
            assert (got, idx) == (int(d["q2_result"][0]), int(d["q2_result"][1])), (got, idx, d["q2_result"])
            assert q.first_bottom_node == -1 and q.first_side_node == idx
        g.close()


def test_average_fields():
    """Output::average_fields (output.cxx:312-352) with device accumulators: snapshot at the first
    step of the interval, running sums afterwards.  The host loop `s_avg[i] += s[i]` is restated
    here on the per-step downloads; the device accumulators must match it bit for bit (and
    tests/test_bridge_gpu.py compares the frames Output::write makes from them with the
    reference binary's)."""
    d = ref_dump("box-ep", [1, 20])
    g = refmodel.model_from_dump(d)
    dt = float(d["s0_dt"])
    t = 0.0
    want = None
    for i in range(6):
        t += dt
        g.step(dt, t, i + 1)
        s = g.download_state("coord", "strain", "stress", "delta_plstrain")
        g.average_fields(first=(i == 0))
        if i == 0:
            want = {k: v.copy() for k, v in s.items()}
        else:
            want["stress"] += s["stress"]
            want["delta_plstrain"] += s["delta_plstrain"]
    coord0, strain0, stress_avg, dpl_avg = g.download_averages()
    assert np.array_equal(coord0, want["coord"])
    assert np.array_equal(strain0, want["strain"])
    assert np.array_equal(stress_avg, want["stress"])
    assert np.array_equal(dpl_avg, want["delta_plstrain"])
    # and without the per-step downloads (the deferred tail is flushed by average_fields itself)
    h = refmodel.model_from_dump(d)
    t = 0.0
    for i in range(6):
        t += dt
        h.step(dt, t, i + 1)
        h.average_fields(first=(i == 0))
    for a, b in zip(h.download_averages(), (coord0, strain0, stress_avg, dpl_avg)):
        assert np.array_equal(a, b)
    g.close(); h.close()


def test_mesh_quality_reductions_at_scale():
    """Size-independent check of the device reductions of bad_mesh_quality on a 1 M-tet box with
    seeded node jitter (thousands of blocks, ties impossible to arrange by hand): the worst
    element, its quality, the first tiny element and the first displaced bottom node equal a numpy
    restatement of geometry.cxx:1873-1927 / remeshing.cxx:2764-2793 on the same arrays."""
    import sys
    from conftest import ROOT
    sys.path.insert(0, ROOT)
    import bench
    from dynearthsol_b200 import boxmesh
    prob = bench.build_problem("1M")
    N, E = prob["nnode"], prob["nelem"]
    rng = np.random.default_rng(20261104)
    coord = prob["state"]["coord"].reshape(3, N).copy()
    res = bench.SIZES["1M"][3]
    coord += rng.uniform(-0.2, 0.2, size=coord.shape) * res
    conn = prob["mesh"]["connectivity"].reshape(4, E)
    vol = boxmesh.tet_volumes(coord, conn)
    g, dt = bench.make_model(prob, 0)
    g.upload_state(coord=coord, volume=vol)
    p = [coord[:, conn[i]] for i in range(4)]

    def area(a, b, c):
        ab, ac = b - a, c - a
        d0 = ab[1] * ac[2] - ab[2] * ac[1]
        d1 = ab[2] * ac[0] - ab[0] * ac[2]
        d2 = ab[0] * ac[1] - ab[1] * ac[0]
        return np.sqrt(d0 * d0 + d1 * d1 + d2 * d2) / 2

    asum = area(p[0], p[1], p[2]) + area(p[0], p[1], p[3]) + area(p[2], p[3], p[0]) + area(p[2], p[3], p[1])
    quality = 216 * np.sqrt(3.0) * vol * vol / (asum * asum * asum)
    xl, yl, zl, _ = bench.SIZES["1M"]
    small_size = float(np.median(vol)) / (0.118 * res ** 3)       # half the elements are "too small"
    code, idx, mq, q = g.mesh_quality(smallest_size=small_size, resolution=res, max_boundary_distortion=0.1,
                                      min_quality=0.4, xlength=xl, ylength=yl, zlength=zl, remeshing_option=11)
    smallest_vol = small_size * 0.118 * res ** 3
    assert q.first_small_elem == int(np.nonzero(vol < smallest_vol)[0][0])
    assert (code, idx) == (3, q.first_small_elem)
    w = int(np.argmin(quality))
    assert q.worst_elem == w
    assert abs(q.worst_quality - quality[w]) <= 1e-13 * quality[w]
    bottom = np.nonzero(((prob["mesh"]["bcflag"] & 16) != 0) & (np.abs(coord[2] + zl) > 0.1 * res))[0]
    assert q.first_bottom_node == (int(bottom[0]) if len(bottom) else -1)
    assert abs(q.bottom_value - coord[2, q.first_bottom_node]) == 0 if len(bottom) else True
    g.close()


@pytest.mark.parametrize("option", [0, 2, 3, 4])
def test_damping_options(option):
    """apply_damping (fields.cxx:483-579), the non-default branches: none, proportional, the
    asymmetric one, and Rayleigh (4), which needs compute_mass's ymass column
    (geometry.cxx:1833-1866).  One and twenty steps against the reference run with the same
    control.damping_option."""
    d = ref_dump("box-ep", [1, 20], sets=(("damping_option", option),))
    assert int(d["p_damping_option"]) == option
    g = refmodel.model_from_dump(d)
    dt = float(d["s0_dt"])
    g.step(dt, dt, 1)
    compare(g.download_state(*FIELDS), d, "s1_", FIELDS)
    g.run(19, dt, dt, 1)
    compare(g.download_state(*FIELDS), d, "s20_", FIELDS, tol=1e-9)
    g.close()


def test_check_nan():
    """check_nan (utils.hpp:323-385) as a device count: zero on a healthy state, and exactly the
    number of poisoned entries in the arrays the reference scans."""
    d = ref_dump("box-ep", [1, 20])
    g = refmodel.model_from_dump(d)
    dt = float(d["s0_dt"])
    g.run(3, dt)
    assert g.check_nan() == 0
    s = g.download_state("stress", "temperature", "vel", "volume")
    s["stress"][[5, 2880 + 7, 5 * 2880 + 11]] = np.nan
    s["temperature"][3] = np.nan
    s["vel"][[0, 819 + 1]] = np.nan
    s["volume"][17] = np.nan
    g.upload_state(**s)
    assert g.check_nan() == 7
    g.close()
