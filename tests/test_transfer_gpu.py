"""Remesh field transfer on the device through the C-ABI
(des_remesh_transfer_cells / des_remesh_transfer_nodes) against the reference's own
nearest_neighbor_interpolation and barycentric_node_interpolation as recorded inside its remesh():
is_changed and every transferred field bit for bit."""
import numpy as np
import pytest

import transfer_case as tc
from dynearthsol_b200 import core

pytestmark = pytest.mark.gpu


def dev_cells(old_coord, old_conn, old_size, new_coord, new_conn, fields, is_surface):
    return core.remesh_transfer_cells(old_coord, old_conn, old_size, new_coord, new_conn, fields, is_surface)


def dev_nodes(old_coord, old_conn, old_size, sidx, sarr, new_coord, fields):
    return core.remesh_transfer_nodes(old_coord, old_conn, old_size, sidx, sarr, new_coord, fields)


def run_case(c, synthetic):
    changed, st = tc.check_cells(c, "nn0_", False, dev_cells)
    assert (changed == 0).any() and (changed == 1).any() and (changed == -1).any()
    assert st.nsample == 175
    _, st1 = tc.check_cells(c, "nn1_", True, dev_cells)
    assert st1.nsample == 210
    _, _, stn = tc.check_nodes(c, "brc_", dev_nodes)
    assert stn.n_outside > 0
    if synthetic:
        el, brc, sts = tc.check_nodes(c, "syn_", dev_nodes)
        assert sts.nchanged - sts.n_on_face - sts.n_outside > 0 and sts.n_on_face > 0 and sts.n_outside > 0
    return st, stn


def test_transfer_matches_recorded_reference_case():
    """The committed recording (1 314 -> 1 328 elements, every branch of the node search)."""
    run_case(tc.load_case(tc.GOLDEN), True)


def test_transfer_matches_reference_remesh_of_test_3d_remesh(tmp_path):
    """tests/cfg/t3d-remesh.cfg (benchmarks-cores/test-3d-remesh.cfg, BASELINE.json configs[4]) after
    300 steps, recorded now from the reference compiled under oracle/_ref: 13 850 -> 13 865 elements,
    4 741 of them changed (829 675 sample points located among their 32 nearest old elements)."""
    c = tc.live_case("t3d-remesh.cfg", 300, tmp_path)
    if c is None:
        pytest.skip("oracle/_ref/libdesref.so not built")
    st, stn = run_case(c, True)
    assert st.nchanged > 1000
    print("cells: search %.3f s, ratios %.3f s, fields %.3f s; nodes: %.3f / %.3f / %.3f s"
          % (st.search_seconds, st.acm_seconds, st.field_seconds,
             stn.search_seconds, stn.acm_seconds, stn.field_seconds))


@pytest.mark.parametrize("resolution", [6e3, 3e3])
def test_transfer_on_tiny_meshes(resolution, tmp_path):
    """box-ep-tiny at 6 km (10 elements, fewer than the 32 neighbours asked for) and 3 km (120),
    transferred onto its jittered self: every element changed; against the reference's own calls."""
    c = tc.tiny_box_case(resolution, tmp_path)
    if c is None:
        pytest.skip("oracle/_ref/libdesref.so not built")
    changed, st = tc.check_cells(c, "nn0_", False, dev_cells)
    assert st.nchanged == changed.size
    tc.check_nodes(c, "brc_", dev_nodes)


def test_transfer_in_many_small_passes(monkeypatch):
    """The sample points of the changed cells and the nodes of the wider search are processed in
    passes that bound the scratch (64 Mi sample points / 4 096 nodes); forced down to 37 cells /
    nodes per pass the recorded case still comes out bit-identical."""
    monkeypatch.setenv("DES_TRANSFER_BLOCK", "37")
    run_case(tc.load_case(tc.GOLDEN), True)


def test_transfer_runs_twice_with_the_same_bits():
    c = tc.load_case(tc.GOLDEN)
    f = tc.fields_of(c, "nn0_")
    a = dev_cells(c["nn0_old_coord"], c["nn0_old_conn"], c["nn0_old_size"], c["nn0_new_coord"], c["nn0_new_conn"], f, False)
    b = dev_cells(c["nn0_old_coord"], c["nn0_old_conn"], c["nn0_old_size"], c["nn0_new_coord"], c["nn0_new_conn"], f, False)
    assert np.array_equal(a[1], b[1])
    for k in a[0]:
        assert np.array_equal(a[0][k], b[0][k]), k


def test_transfer_onto_the_same_mesh_is_the_identity():
    """Old mesh == new mesh: every centroid distance is 0, nothing is changed, fields come back as
    they went in; every node takes the dist == 0 shortcut and gets weights (1, 0, 0, 0) permuted."""
    c = tc.load_case(tc.GOLDEN)
    f = tc.fields_of(c, "nn0_")
    out, changed, st = dev_cells(c["nn0_old_coord"], c["nn0_old_conn"], c["nn0_old_size"],
                                 c["nn0_old_coord"], c["nn0_old_conn"], f, False)
    assert not changed.any() and st.nchanged == 0
    for k in f:
        assert np.array_equal(out[k], np.asarray(f[k]).reshape(-1)), k
    nf = tc.fields_of(c, "brc_")
    out, el, brc, st = dev_nodes(c["brc_old_coord"], c["brc_old_conn"], c["brc_old_size"],
                                 c["brc_old_support_idx"], c["brc_old_support_arr"], c["brc_old_coord"], nf)
    assert st.nchanged == 0
    assert np.isin(brc, (0.0, 1.0)).all() and (brc.sum(axis=1) == 1).all()
    for k in nf:
        assert np.array_equal(out[k], np.asarray(nf[k]).reshape(-1)), k
