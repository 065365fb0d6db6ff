"""des_gpu_compute_dt_static (include/des_gpu.h): compute_dt's reductions out of the library's
file-scope static device scratch, no allocation.  The entry has no counterpart in the reference
(it was added); the checker here is the existing entry
point des_gpu_compute_dt on the same handle, and both are min / max reductions of the same
per-element values, so the comparison is bit-exact.

This file is named so that pytest collects it last.
It needs nothing but the committed golden snapshot and passes alone in a fresh process as well
as after the other tests (the static array keeps whatever an earlier call left in it)."""
import os

import numpy as np
import pytest

import refmodel
from dynearthsol_b200 import boxmesh as bm, core

pytestmark = pytest.mark.gpu

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "box-ep-tiny.npz")
E_ARG, E_UNSUPPORTED, E_STATE = 3, 4, 5       # include/des_gpu.h (core.ERRORS)
FIELDS = ("minl", "dt_maxwell", "dt_diffusion", "global_dt_min", "global_max_vem", "max_surf_vel")


def as_tuple(r):
    return tuple(float(getattr(r, k)) for k in FIELDS)


def same(a, b):
    # bit-exact: identical doubles (np.array_equal would call two NaNs different, and there are none)
    return np.array_equal(np.array(as_tuple(a)).view(np.uint64), np.array(as_tuple(b)).view(np.uint64))


def test_static_scratch_matches_compute_dt_two_calls_in_a_row():
    d = np.load(GOLDEN)
    g = refmodel.model_from_dump(d)          # calls the new entry
    w = refmodel.model_from_dump(d)          # never does: the witness of an untouched step schedule
    dt = float(d["s0_dt"])

    # two calls in a row, before any step; then the existing entry on the same state
    a1 = g.compute_dt_reductions_static()
    a2 = g.compute_dt_reductions_static()
    ref = g.compute_dt_reductions()
    assert ref.minl > 0 and ref.global_dt_min > 0
    assert same(a1, ref), (as_tuple(a1), as_tuple(ref))
    assert same(a2, ref), (as_tuple(a2), as_tuple(ref))
    assert same(w.compute_dt_reductions(), ref)

    # mid-run, with the end-of-step work of the fused schedule still deferred, in both orders
    for i in range(3):
        g.step(dt, dt * (i + 1), i + 1)
        w.step(dt, dt * (i + 1), i + 1)
    b1 = g.compute_dt_reductions_static()       # before the existing entry (volume still pending)
    ref = g.compute_dt_reductions()
    b2 = g.compute_dt_reductions_static()       # after it (its element loop already ran)
    b3 = g.compute_dt_reductions_static()
    assert not same(ref, a1)                    # the state moved: a stale partial would show
    for b in (b1, b2, b3):
        assert same(b, ref), (as_tuple(b), as_tuple(ref))
    assert same(w.compute_dt_reductions(), ref)

    # the new entry left the handle's schedule alone: both handles keep stepping identically
    for i in range(3, 6):
        g.step(dt, dt * (i + 1), i + 1)
        w.step(dt, dt * (i + 1), i + 1)
    names = ("coord", "vel", "stress", "strain", "plstrain", "temperature", "volume")
    sg, sw = g.download_state(*names), w.download_state(*names)
    for k in names:
        assert np.array_equal(sg[k], sw[k]), k
    assert same(g.compute_dt_reductions_static(), w.compute_dt_reductions())
    g.close(); w.close()


def test_static_scratch_after_another_handle_used_it():
    """Every handle on the device shares the one array and it keeps the last call's contents: a
    second, different mesh right after must not see them."""
    d = np.load(GOLDEN)
    g = refmodel.model_from_dump(d)
    big = g.compute_dt_reductions_static()
    xl, yl, zl = 30e3, 20e3, 10e3
    coord, conn, bcflag = bm.regular_box(xl, yl, zl, 7, 5, 4)
    mesh = bm.boundary_topology(coord, conn, bcflag, support=core.host_build_support(coord.shape[1], conn))
    mesh["coord"] = coord
    phys = dict(bm.TEST3D_PHYSICS); phys["zlength"] = zl
    state, comp_p = bm.initial_fields(coord, conn, rho0=phys["rho0"][0], bulk_modulus=phys["bulk_modulus"][0],
                                      gravity=phys["gravity"], xlength=xl, ylength=yl, zlength=zl,
                                      resolution=5e3, weakzone=bm.TEST3D_WEAKZONE)
    s = core.DesGpu(core.make_param(**phys), 0)
    s.upload_mesh(mesh)
    s.upload_elemmarkers(np.full(conn.shape[1], 4, dtype=np.int32))
    s.set_compensation_pressure(comp_p)
    s.upload_state(**state)
    small = s.compute_dt_reductions_static()
    assert same(small, s.compute_dt_reductions())
    assert not same(small, big)
    assert same(g.compute_dt_reductions_static(), big)
    g.close(); s.close()


def test_static_scratch_refusals():
    """Sizes beyond the stated maximum, group handles and handles without a state are refused
    with a message, not served out of a too-small array."""
    d = np.load(GOLDEN)
    p = refmodel.param_from_dump(d)
    e = core.DesGpu(p, 0)
    with pytest.raises(core.DesError) as ei:
        e.compute_dt_reductions_static()
    assert ei.value.code == E_STATE
    e.close()

    m = core.DesGpu(p, devices=[0, 0])
    m.upload_mesh(refmodel.mesh_from_dump(d))
    m.upload_elemmarkers(d["elemmarkers"])
    m.upload_state(**{k: d["s0_" + k] for k in refmodel.UPLOAD})
    with pytest.raises(core.DesError) as ei:
        m.compute_dt_reductions_static()
    assert ei.value.code == E_UNSUPPORTED
    m.close()

    # 150^3 cells x 5 tetrahedra = 16 875 000 elements > 16 777 216
    limit = 16777216
    xl = yl = zl = 150e3
    coord, conn, bcflag = bm.regular_box(xl, yl, zl, 151, 151, 151)
    assert conn.shape[1] > limit
    mesh = bm.boundary_topology(coord, conn, bcflag, support=core.host_build_support(coord.shape[1], conn))
    mesh["coord"] = coord
    phys = dict(bm.TEST3D_PHYSICS); phys["zlength"] = zl
    state, comp_p = bm.initial_fields(coord, conn, rho0=phys["rho0"][0], bulk_modulus=phys["bulk_modulus"][0],
                                      gravity=phys["gravity"], xlength=xl, ylength=yl, zlength=zl,
                                      resolution=1e3, weakzone=bm.TEST3D_WEAKZONE)
    b = core.DesGpu(core.make_param(**phys), 0)
    b.upload_mesh(mesh)
    b.upload_elemmarkers(np.full(conn.shape[1], 4, dtype=np.int32))
    b.set_compensation_pressure(comp_p)
    b.upload_state(**state)
    with pytest.raises(core.DesError) as ei:
        b.compute_dt_reductions_static()
    assert ei.value.code == E_ARG
    assert str(limit) in str(ei.value) and "des_gpu_compute_dt" in str(ei.value)
    assert b.compute_dt_reductions().minl > 0      # the allocating entry still serves it
    b.close()
