"""One handle over several devices (des_gpu_create_multi, csrc/des_multi.cuh): the library
partitions the mesh, exchanges the ghost-node values between its parts and gathers / scatters
the host's global arrays, so the driver's call sites stay what they are.  Required: every field
BIT-IDENTICAL to the single-device handle (hence to the reference wherever that one is), run
after run, for 2, 3 and 4 parts -- on one device (parts share it) and, when the box has them,
on distinct devices."""
import numpy as np
import pytest

import refmodel
from conftest import ref_dump
from dynearthsol_b200 import core

pytestmark = pytest.mark.gpu

FIELDS = ["coord", "vel", "temperature", "stress", "strain", "strain_rate", "plstrain", "delta_plstrain",
          "volume", "volume_old", "force", "mass", "tmass", "volume_n", "dhacc", "edvacc_surf", "viscosity"]


def run(d, devices, nsteps=23, device=0, quality=False, reorder=0):
    mesh = refmodel.mesh_from_dump(d)
    mesh["coord"] = d["s0_coord"]
    g = core.DesGpu(refmodel.param_from_dump(d), device, devices=devices)
    g.set_reorder(reorder)
    g.upload_mesh(mesh)
    g.upload_elemmarkers(d["elemmarkers"])
    g.set_compensation_pressure(float(d["p_compensation_pressure"]))
    g.upload_state(**{k: d["s0_" + k] for k in refmodel.UPLOAD})
    dt = g.compute_dt()[0]
    snaps, t = {}, 0.0
    for s in range(1, nsteps + 1):
        t += dt
        if s == 7:      # the bridge's marker step: segments, a download in the middle, the swap
            for seg in range(5):
                g.step_segment(seg, dt, t)
            mid = g.download_state("coord", "volume", "dhacc", "edvacc_surf")
            g.upload_state(dhacc=mid["dhacc"], edvacc_surf=mid["edvacc_surf"])
            g.upload_elemmarkers(d["elemmarkers"])
            g.step_segment(5, dt, t)
            snaps["mid"] = mid
        else:
            g.step(dt, t, s)
        if s % 10 == 0:
            dt = g.compute_dt()[0]
            snaps["dt%d" % s] = {"dt": np.array([dt])}
        if s in (1, 12, nsteps):
            snaps[s] = g.download_state(*FIELDS)
    info = g.partition_info()
    q = None
    if quality:
        q = g.mesh_quality(smallest_size=0.01, resolution=float(d["p_zlength"]) / 4, max_boundary_distortion=1e-9,
                           min_quality=0.9, xlength=1.0, ylength=1.0, zlength=float(d["p_zlength"]), remeshing_option=11)
        q = q[:3] + (tuple(getattr(q[3], f) for f, _ in core.DesQuality._fields_),)
    g.close()
    return snaps, info, q


@pytest.mark.parametrize("case", ["box-ep", "box-evp"])
@pytest.mark.parametrize("nparts", [2, 3, 4])
def test_group_handle_equals_single_device(case, nparts):
    import torch
    d = ref_dump(case, [1])
    one, info1, q1 = run(d, None, quality=True)
    assert info1["nparts"] == 1
    ndev = torch.cuda.device_count()
    devices = [i % ndev for i in range(nparts)]
    many, info, q = run(d, devices, quality=True)
    assert info["nparts"] == nparts and min(info["local_elems"]) > 0
    assert sum(info["ghost_nodes"]) > 0 and info["partition_seconds"] > 0
    # every element lives on at least one part, most on exactly one
    assert sum(info["local_elems"]) >= int(d["p_nelem"])
    for k in one:
        for f in one[k]:
            assert np.array_equal(one[k][f], many[k][f]), (k, f, np.abs(one[k][f] - many[k][f]).max())
    assert q1 == q, (q1, q)
    again, _, _ = run(d, devices)
    for k in again:
        for f in again[k]:
            assert np.array_equal(again[k][f], many[k][f]), ("run to run", k, f)


@pytest.mark.parametrize("case", ["box-ep", "box-evp"])
def test_morton_storage_order_changes_no_bit(case):
    """des_gpu_set_reorder(h, 2): nodes and elements stored in Morton order inside the library.
    Every array crosses the C-ABI in the caller's order and every node's support keeps the
    caller's element order, so all fields, dt, the marker-step download in the middle of a step
    and the mesh-quality ids must equal the default order's bit for bit -- on one device, with the
    elements-by-smallest-node order, and with three parts that each reorder their own sub-mesh."""
    d = ref_dump(case, [1])
    one, _, q1 = run(d, None, quality=True)
    for reorder, devices in ((2, None), (1, None), (2, [0, 0, 0])):
        got, _, q = run(d, devices, quality=True, reorder=reorder)
        for k in one:
            for f in one[k]:
                assert np.array_equal(one[k][f], got[k][f]), (reorder, devices, k, f)
        assert q1 == q, (reorder, devices, q1, q)


def test_graph_launch_equals_kernel_by_kernel_launches(monkeypatch):
    """Meshes below DES_GRAPH elements (default 2M) run des_gpu_step as one CUDA Graph launch,
    dt reaching the kernels through device memory.  Same bits as launching kernel by kernel,
    through the marker step, the compute_dt steps and the steps right after them."""
    d = ref_dump("box-evp", [1])
    graphed, _, _ = run(d, None, nsteps=33)
    monkeypatch.setenv("DES_GRAPH", "0")
    plain, _, _ = run(d, None, nsteps=33)
    for k in plain:
        for f in plain[k]:
            assert np.array_equal(plain[k][f], graphed[k][f]), (k, f)


def test_group_handle_refuses_what_it_cannot_do():
    d = ref_dump("box-ep", [1])
    mesh = refmodel.mesh_from_dump(d)
    mesh.pop("coord", None)
    g = core.DesGpu(refmodel.param_from_dump(d), devices=[0, 0])
    with pytest.raises(core.DesError) as e:
        g.upload_mesh(mesh)             # no coordinates: nothing to bisect
    assert e.value.code == 3
    mesh["coord"] = d["s0_coord"]
    g.upload_mesh(mesh)
    g.upload_elemmarkers(d["elemmarkers"])
    g.upload_state(**{k: d["s0_" + k] for k in refmodel.UPLOAD})
    with pytest.raises(core.DesError) as e:
        g.call("update_force")          # the twins run on a single-device handle
    assert e.value.code == 4
    g.close()
    with pytest.raises(core.DesError):
        core.DesGpu(refmodel.param_from_dump(d), devices=[0, 99])


def test_non_positive_volume_is_reported():
    """the step reports NaN / negative volume through its status.  A node
    pushed through the opposite facet inverts its elements; the next sync point of the
    asynchronous step loop (des_gpu_compute_dt) returns DES_E_NAN instead of a dt."""
    d = ref_dump("box-ep", [1])
    g = refmodel.model_from_dump(d)
    dt = g.compute_dt()[0]
    g.step(dt, dt, 1)
    assert g.compute_dt()[0] > 0
    N = int(d["p_nnode"])
    coord = g.download_state("coord")["coord"].reshape(3, N).copy()
    inner = int(np.argmin(np.abs(coord[0] - coord[0].mean()) + np.abs(coord[1] - coord[1].mean())
                          + np.abs(coord[2] - coord[2].mean())))
    coord[2, inner] += 10 * (coord[2].max() - coord[2].min())      # far through its neighbours
    g.upload_state(coord=coord)
    g.step(dt, 2 * dt, 2)
    g.step(dt, 3 * dt, 3)       # the inverted volume is computed by this step's E-A (deferred compute_volume)
    with pytest.raises(core.DesError) as e:
        g.compute_dt()
    assert e.value.code == 11
    g.close()
