"""phase_changes on device-resident markers (des_gpu_upload_markers /
des_gpu_phase_changes) against the reference's own phase_changes (phasechanges.cxx:109-158) on
the oblique-rift mesh with eight rock types: every marker a random rock type, every node a random
temperature, every element a random hydrous-marker count (tests/refdump.py: phase_case).  Rock
types, the hyd_inc the reference leaves with set_tmp, and the elemmarkers counters: exact."""
import numpy as np
import pytest

import refmodel
from conftest import ref_dump
from dynearthsol_b200 import core

pytestmark = pytest.mark.gpu

OVER = {"mat.num_materials": 8, "mat.rho0": "[3300, 3280, 2850, 2700, 2400, 3280, 2850, 2700]",
        "mat.phase_change_option": 1, "control.has_hydration_processes": "yes",
        "mat.mattype_mantle": 0, "mat.mattype_crust": 2}


def case():
    d = ref_dump("orift", (1,), over=OVER)
    if "pc_mattype1" not in d:
        pytest.skip("no phase-change recording (oracle/_ref/libdesref.so absent)")
    return d


def model(d, reorder=None, ref_pressure_option=None):
    over = {} if ref_pressure_option is None else {"ref_pressure_option": ref_pressure_option}
    g = refmodel.model_from_dump(d, reorder=reorder, **over)
    g.upload_elemmarkers(d["pc_elemmarkers0"])
    g.upload_state(coord=d["pc_coord"], temperature=d["pc_temperature"])
    g.upload_markers(d["pc_elem"], d["pc_eta"], d["pc_mattype0"])
    return g


@pytest.mark.parametrize("reorder", [0, 2])
def test_phase_changes_match_reference(reorder):
    d = case()
    g = model(d, reorder=reorder)
    m0, m1 = d["pc_mattype0"], d["pc_mattype1"]
    cm, ct, hyd, n = g.phase_changes(option=int(d["pc_option"]), hydrous_count=d["pc_hydrous"])
    want = np.flatnonzero(m0 != m1)
    assert n == want.size and np.array_equal(cm, want)          # ascending marker order
    assert np.array_equal(ct, m1[want])
    assert np.array_equal(hyd, d["pc_tmp"].astype(np.int32))    # set_tmp(m, hyd_inc), phasechanges.cxx:131
    nmat = int(d["p_nmat"])
    assert np.array_equal(g.download_elemmarkers(), d["pc_elemmarkers1"].reshape(-1, nmat))
    # each arm of simple_subduction moved some markers
    for a, b in ((0, 1), (1, 0), (2, 3), (4, 5)):
        assert ((m0 == a) & (m1 == b)).any(), (a, b)
    # the device keeps its rock types current: called again it sees the NEW types, and with the
    # same temperature, pressure and hydrous counts nothing can move a second time
    cm2, ct2, _, n2 = g.phase_changes(option=1, hydrous_count=d["pc_hydrous"])
    assert n2 == 0 and cm2.size == 0
    g.close()


def test_phase_changes_small_cap_and_template_option():
    d = case()
    g = model(d)
    want = np.flatnonzero(d["pc_mattype0"] != d["pc_mattype1"])
    cm, ct, hyd, n = g.phase_changes(option=1, hydrous_count=d["pc_hydrous"], cap=10)
    assert n == want.size and np.array_equal(cm, want[:10])    # the count is complete, the list is cut
    cm, ct, hyd, n = g.phase_changes(option=101)                # custom_phase_change changes nothing
    assert n == 0 and not hyd.any()
    with pytest.raises(core.DesError) as ei:
        g.phase_changes(option=7)
    assert ei.value.code == 4
    g.close()


def test_phase_changes_without_hydrous_markers():
    """hydrous_count = NULL: no mantle marker serpentinises; everything else as before."""
    d = case()
    g = model(d)
    m0, m1 = d["pc_mattype0"], d["pc_mattype1"]
    cm, ct, hyd, n = g.phase_changes(option=1)
    want = np.flatnonzero((m0 != m1) & (m0 != 0))
    assert np.array_equal(cm, want) and np.array_equal(ct, m1[want])
    g.close()


def test_marker_set_can_grow_and_shrink_between_uploads():
    """Markers come and go on the host (surface processes, replenishment): a smaller set, then the
    whole one (the device buffers are re-made), then a smaller one again."""
    d = case()
    g = refmodel.model_from_dump(d)
    g.upload_elemmarkers(d["pc_elemmarkers0"])
    g.upload_state(coord=d["pc_coord"], temperature=d["pc_temperature"])
    m0, m1 = d["pc_mattype0"], d["pc_mattype1"]
    k = 500
    g.upload_markers(d["pc_elem"][:k], d["pc_eta"][:, :k], m0[:k])
    cm, ct, hyd, n = g.phase_changes(option=1, hydrous_count=d["pc_hydrous"])
    want = np.flatnonzero(m0[:k] != m1[:k])
    assert np.array_equal(cm, want) and np.array_equal(ct, m1[want]) and hyd.size == k
    g.upload_elemmarkers(d["pc_elemmarkers0"])      # the host's counters, as before the first call
    g.upload_markers(d["pc_elem"], d["pc_eta"], m0)
    cm, ct, hyd, n = g.phase_changes(option=1, hydrous_count=d["pc_hydrous"])
    want = np.flatnonzero(m0 != m1)
    assert np.array_equal(cm, want) and np.array_equal(ct, m1[want])
    assert np.array_equal(g.download_elemmarkers(), d["pc_elemmarkers1"].reshape(-1, int(d["p_nmat"])))
    g.upload_markers(d["pc_elem"][:k], d["pc_eta"][:, :k], m1[:k])
    cm, ct, hyd, n = g.phase_changes(option=1, hydrous_count=d["pc_hydrous"])
    assert n == 0
    g.close()
