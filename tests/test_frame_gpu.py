"""des_gpu_download_frame: the arrays of Output::_write
(output.cxx:95-245) straight from the device in the on-disk layout.  Checked here against the
fields des_gpu_download_state / des_gpu_download_averages return (the reference's in-memory SoA)
put through numpy restatements of _write's host loops -- IEEE double arithmetic in the same
operation order, so the comparison is bit-exact.  That the FILE matches the reference's is
tests/test_bridge_gpu.py's business (every frame of every driver test goes through this path)."""
import os

import numpy as np
import pytest

import refmodel
from dynearthsol_b200 import core

pytestmark = pytest.mark.gpu

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "box-ep-tiny.npz")


def aos(a, n, w):
    """Array2D::pack_to (array2d.hpp; binaryio.cxx:159): SoA [w][n] -> [n][w]."""
    return np.ascontiguousarray(np.asarray(a).reshape(w, n).T)


def tri_area(a, b, c):
    """triangle_area (geometry.cxx, 3D): half the norm of the cross product."""
    ab = b - a
    ac = c - a
    cx = ab[:, 1] * ac[:, 2] - ab[:, 2] * ac[:, 1]
    cy = ab[:, 2] * ac[:, 0] - ab[:, 0] * ac[:, 2]
    cz = ab[:, 0] * ac[:, 1] - ab[:, 1] * ac[:, 0]
    return np.sqrt(cx * cx + cy * cy + cz * cz) / 2


def check_plain(g, d):
    N, E = g.nnode, g.nelem
    f = g.download_frame()
    s = g.download_state("coord", "vel", "temperature", "force", "plstrain", "delta_plstrain", "strain_rate",
                         "strain", "stress", "viscosity", "volume")
    assert np.array_equal(f["coordinate"], aos(s["coord"], N, 3))
    assert np.array_equal(f["velocity"], aos(s["vel"], N, 3))
    assert np.array_equal(f["force"], aos(s["force"], N, 3))
    assert np.array_equal(f["temperature"], s["temperature"])
    assert np.array_equal(f["plstrain"], s["plstrain"])
    assert np.array_equal(f["plstrain_rate"], s["delta_plstrain"])
    assert np.array_equal(f["viscosity"], s["viscosity"])
    for k, src in (("strain_rate", "strain_rate"), ("strain", "strain"), ("stress", "stress")):
        assert np.array_equal(f[k], aos(s[src], E, 6)), k
    assert "velocity_averaged" not in f and "stress_averaged" not in f
    # density: MatProps::rho (matprops.cxx:642-664), one material
    conn = np.asarray(d["mesh_connectivity"]).reshape(4, E)
    T = s["temperature"]
    Te = (((T[conn[0]] + T[conn[1]]) + T[conn[2]]) + T[conn[3]]) / 4
    cnt = np.asarray(d["elemmarkers"]).reshape(E, -1)
    assert cnt.shape[1] == 1
    rho = float(np.atleast_1d(d["p_rho0"])[0]) * (1 - float(np.atleast_1d(d["p_alpha"])[0]) * (Te - 273)) * cnt[:, 0]
    assert np.array_equal(f["density"], rho / cnt[:, 0])
    assert np.array_equal(f["material"], np.zeros(E))
    # mesh quality: elem_quality (geometry.cxx:1873-1899); the areas go through a square root
    # numpy evaluates like the device, the cross products may differ in the last place
    x = aos(s["coord"], N, 3)
    a, b, c, dd = (x[conn[i]] for i in range(4))
    area = tri_area(a, b, c) + tri_area(a, b, dd) + tri_area(c, dd, a) + tri_area(c, dd, b)
    q = 216 * np.sqrt(3) * s["volume"] * s["volume"] / (area * area * area)
    assert np.allclose(f["mesh_quality"], q, rtol=1e-12, atol=0)
    assert (f["mesh_quality"] > 0).all() and (f["mesh_quality"] <= 1 + 1e-12).all()
    return f


@pytest.mark.parametrize("reorder", [0, 2])
def test_frame_matches_state_downloads(reorder):
    d = np.load(GOLDEN)
    keys = set(d.files)
    if not {"p_rho0", "p_alpha"} <= keys:
        pytest.skip("golden snapshot without material parameters")
    g = refmodel.model_from_dump(d, reorder=reorder)
    dt = float(d["s0_dt"])
    for i in range(3):
        g.step(dt, dt * (i + 1), i + 1)
    f1 = check_plain(g, d)
    f2 = g.download_frame()                 # the call changes nothing: the same frame again
    for k in f1:
        assert np.array_equal(f1[k], f2[k]), k
    g.close()


def test_frame_averaged_fields():
    """Output::_write with is_averaged (output.cxx:131-190): (coord - coord0) * inv_dt,
    delta_plstrain_avg * inv_dt, (strain - strain0) * inv_dt, stress_avg / (interval + 1)."""
    d = np.load(GOLDEN)
    g = refmodel.model_from_dump(d)
    N, E = g.nnode, g.nelem
    dt = float(d["s0_dt"])
    interval = 4
    for i in range(interval):
        g.step(dt, dt * (i + 1), i + 1)
        g.average_fields(i == 0)
    with_avg = dict(averaged=True, inv_dt=1.0 / (interval * dt), stress_avg_scale=1.0 / (interval + 1))
    f = g.download_frame(**with_avg)
    coord0, strain0, stress_sum, dpl_sum = g.download_averages()
    s = g.download_state("coord", "strain", "stress", "strain_rate")
    inv_dt = with_avg["inv_dt"]
    assert np.array_equal(f["velocity_averaged"], aos((s["coord"] - np.asarray(coord0).ravel()) * inv_dt, N, 3))
    assert np.array_equal(f["plstrain_rate"], np.asarray(dpl_sum).ravel() * inv_dt)
    assert np.array_equal(f["strain_rate"], aos((s["strain"] - np.asarray(strain0).ravel()) * inv_dt, E, 6))
    assert np.array_equal(f["stress_averaged"], aos(np.asarray(stress_sum).ravel() * with_avg["stress_avg_scale"], E, 6))
    assert np.array_equal(f["stress"], aos(s["stress"], E, 6))
    assert np.abs(f["velocity_averaged"]).max() > 0 and np.abs(f["stress_averaged"]).max() > 0
    # the accumulators are left as they were (the reference scales its own copies in place)
    again = g.download_frame(**with_avg)
    for k in f:
        assert np.array_equal(f[k], again[k]), k
    g.close()


def test_frame_refusals():
    d = np.load(GOLDEN)
    g = refmodel.model_from_dump(d)
    with pytest.raises(core.DesError) as ei:                 # no averaging interval was started
        g.download_frame(averaged=True, inv_dt=1.0, stress_avg_scale=0.5)
    assert ei.value.code == 5
    with pytest.raises(core.DesError) as ei:
        g.download_frame(names=("velocity_averaged",))
    assert ei.value.code == 3
    g.close()
    m = core.DesGpu(refmodel.param_from_dump(d), devices=[0, 0])
    m.upload_mesh(refmodel.mesh_from_dump(d))
    m.upload_elemmarkers(d["elemmarkers"])
    m.upload_state(**{k: d["s0_" + k] for k in refmodel.UPLOAD})
    with pytest.raises(core.DesError) as ei:
        m.download_frame()
    assert ei.value.code == 4
    m.close()
