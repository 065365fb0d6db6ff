"""Remesh field transfer, the parts that need no GPU:
  * the recorded reference case is what the reference does now (when oracle/_ref is present);
  * the per-item routines of dynearthsol_b200/csrc/des_transfer.cuh -- the source the CUDA kernels
    are one-line loops over -- run in host loops (tests/hostcheck, test infrastructure) reproduce
    the reference's nearest_neighbor_interpolation and barycentric_node_interpolation bit for bit;
  * the cell-grid k-NN equals a brute-force search, ties included;
  * the C-ABI entry points reject bad arguments before touching a device."""
import ctypes as C
import os
import subprocess
import numpy as np
import pytest

import transfer_case as tc
from dynearthsol_b200 import core

HC_DIR = os.path.join(tc.HERE, "hostcheck")
HC_LIB = os.path.join(HC_DIR, "_build", "libtransfer_hostcheck.so")


def hostcheck():
    src = os.path.join(HC_DIR, "transfer_hostcheck.cpp")
    dep = os.path.join(tc.ROOT, "dynearthsol_b200", "csrc", "des_transfer.cuh")
    if not os.path.exists(HC_LIB) or os.path.getmtime(HC_LIB) < max(os.path.getmtime(src), os.path.getmtime(dep)):
        os.makedirs(os.path.dirname(HC_LIB), exist_ok=True)
        subprocess.check_call(["g++", "-O2", "-ffp-contract=off", "-fopenmp", "-fPIC", "-shared", "-std=c++17",
                               "-o", HC_LIB, src])
    L = C.CDLL(HC_LIB)
    V = C.c_void_p
    L.hostcheck_transfer_cells.argtypes = [V, C.c_int, C.c_int, V, V, V, V]
    L.hostcheck_transfer_nodes.argtypes = [V, V, V, C.c_int, V, V, V, V]
    L.hostcheck_knn.argtypes = [C.c_int, V, C.c_int, V, C.c_int, V, V]
    return L


def host_cells(L):
    def run(old_coord, old_conn, old_size, new_coord, new_conn, fields, is_surface):
        m, arr, out, keep = core._transfer_args(old_coord, old_conn, old_size, new_coord, new_conn, fields,
                                                lambda m: m.old_ncell, lambda m: m.new_ncell)
        changed = np.empty(m.new_ncell, dtype=np.int32)
        st = core.DesTransferStats()
        rc = L.hostcheck_transfer_cells(C.addressof(m), int(is_surface), len(fields), C.addressof(arr),
                                        changed.ctypes.data, C.addressof(st), None)
        assert rc == 0
        return out, changed, st
    return run


def host_nodes(L):
    def run(old_coord, old_conn, old_size, sidx, sarr, new_coord, fields):
        m, arr, out, keep = core._transfer_args(old_coord, old_conn, old_size, new_coord, None, fields,
                                                lambda m: m.old_nnode, lambda m: m.new_nnode)
        sidx = np.ascontiguousarray(sidx, dtype=np.int32); sarr = np.ascontiguousarray(sarr, dtype=np.int32)
        el = np.empty(m.new_nnode, dtype=np.int32); brc = np.empty(4 * m.new_nnode)
        st = core.DesTransferStats()
        rc = L.hostcheck_transfer_nodes(C.addressof(m), sidx.ctypes.data, sarr.ctypes.data, len(fields),
                                        C.addressof(arr), el.ctypes.data, brc.ctypes.data, C.addressof(st))
        assert rc == 0
        return out, el, brc.reshape(-1, 4), st
    return run


def test_golden_transfer_case_is_the_reference(tmp_path):
    """tests/golden/transfer-t3d-remesh-coarse.npz == what the reference records now."""
    live = tc.live_case("t3d-remesh-coarse.cfg", 300, tmp_path)
    if live is None:
        pytest.skip("oracle/_ref/libdesref.so not built")
    gold = tc.load_case(tc.GOLDEN)
    for k, v in gold.items():
        assert np.array_equal(v, live[k]), k


def test_per_item_routines_reproduce_the_reference_transfer():
    L = hostcheck()
    c = tc.load_case(tc.GOLDEN)
    changed, st = tc.check_cells(c, "nn0_", False, host_cells(L))
    # the three kinds of new element: same centroid, remapped by volume ratios, kept on the nearest
    assert (changed == 0).any() and (changed == 1).any() and (changed == -1).any()
    assert st.nsample == 175                    # 10 x 11 x 10 lattice points with eta_3 >= 1e-15
    _, st = tc.check_cells(c, "nn1_", True, host_cells(L))
    assert st.nsample == 210 and st.nchanged > 0
    _, _, st = tc.check_nodes(c, "brc_", host_nodes(L))
    assert st.n_outside > 0                     # boundary nodes of the new mesh off the old hull
    el, brc, st = tc.check_nodes(c, "syn_", host_nodes(L))
    # the synthetic points take every branch of prepare_interpolation (brc-interpolation.cxx:196-329)
    found_by_search = st.nchanged - st.n_on_face - st.n_outside
    assert found_by_search > 0 and st.n_on_face > 0 and st.n_outside > 0
    exact = np.isin(brc, (0.0, 1.0)).all(axis=1)
    assert exact.any() and not exact.all()      # the dist == 0 shortcut normalises the weights


@pytest.mark.parametrize("resolution", [6e3, 3e3])
def test_per_item_routines_on_tiny_meshes(resolution, tmp_path):
    """10 and 120 elements: fewer old elements than neighbours asked for, every element changed."""
    c = tc.tiny_box_case(resolution, tmp_path)
    if c is None:
        pytest.skip("oracle/_ref/libdesref.so not built")
    L = hostcheck()
    changed, st = tc.check_cells(c, "nn0_", False, host_cells(L))
    assert st.nchanged == changed.size
    tc.check_nodes(c, "brc_", host_nodes(L))


@pytest.mark.parametrize("kind", ["random", "lattice", "flat"])
def test_grid_knn_equals_brute_force(kind):
    L = hostcheck()
    rng = np.random.default_rng(3)
    if kind == "random":
        pts = rng.uniform(-1, 1, (3, 4000)) * np.array([[100e3], [10e3], [10e3]])
        q = rng.uniform(-1.2, 1.2, (3, 600)) * np.array([[100e3], [10e3], [10e3]])
    elif kind == "lattice":                     # exact ties everywhere: the lower index wins
        g = np.arange(12, dtype=np.float64)
        pts = np.stack(np.meshgrid(g, g, g, indexing="ij")).reshape(3, -1) * 250.0
        q = np.concatenate([pts[:, ::7] + 125.0, pts[:, ::11], rng.uniform(-300, 3300, (3, 200))], axis=1)
    else:                                       # the top-surface pass: z = 0 for every point
        pts = np.concatenate([rng.uniform(0, 1e5, (2, 900)), np.zeros((1, 900))])
        q = np.concatenate([rng.uniform(-1e3, 1.01e5, (2, 300)), rng.uniform(-500, 500, (1, 300))])
    pts = np.ascontiguousarray(pts); q = np.ascontiguousarray(q)
    for k in (1, 5, 6, 32):                  # 1 and 6 (ACM_K1) also through the fixed-size lists of the kernels
        a = np.empty((q.shape[1], k), dtype=np.int32); b = np.empty_like(a)
        assert L.hostcheck_knn(pts.shape[1], pts.ctypes.data, q.shape[1], q.ctypes.data, k,
                               a.ctypes.data, b.ctypes.data) == 0
        assert np.array_equal(a, b), (kind, k, int((a != b).any(axis=1).sum()))
        assert (a >= 0).all()


@pytest.mark.parametrize("kind", ["one-point", "coincident", "collinear", "two-cells-far-apart"])
def test_grid_knn_on_degenerate_point_sets(kind):
    """Point sets with no extent, or extent along one axis only: the cell grid degenerates to one
    cell or one row of cells and the search still equals the brute-force one."""
    L = hostcheck()
    rng = np.random.default_rng(1)
    if kind == "one-point":
        pts = np.array([[3.0], [4.0], [5.0]])
    elif kind == "coincident":
        pts = np.tile(np.array([[1.0], [2.0], [3.0]]), (1, 40))
    elif kind == "collinear":
        pts = np.stack([np.linspace(0, 1e5, 300), np.full(300, 7.0), np.full(300, -2.0)])
    else:
        pts = np.concatenate([rng.uniform(0, 1, (3, 50)), rng.uniform(0, 1, (3, 50)) + 1e6], axis=1)
    q = np.concatenate([pts[:, :20] + 0.25, rng.uniform(-10, 10, (3, 30)), rng.uniform(-1, 1, (3, 10)) * 2e6], axis=1)
    pts = np.ascontiguousarray(pts); q = np.ascontiguousarray(q)
    for k in (1, 6, 32):
        a = np.empty((q.shape[1], k), dtype=np.int32); b = np.empty_like(a)
        assert L.hostcheck_knn(pts.shape[1], pts.ctypes.data, q.shape[1], q.ctypes.data, k,
                               a.ctypes.data, b.ctypes.data) == 0
        assert np.array_equal(a, b), (kind, k)
        assert ((a >= 0).sum(axis=1) == min(k, pts.shape[1])).all()


def test_transfer_entry_points_check_their_arguments():
    """DES_E_ARG before any device is touched (this runs where there is none)."""
    L = core.load_library()
    c = tc.load_case(tc.GOLDEN)
    flds = {"plstrain": c["nn0_src_plstrain"]}

    def cells(**over):
        a = dict(old_coord=c["nn0_old_coord"], old_conn=c["nn0_old_conn"], old_size=c["nn0_old_size"],
                 new_coord=c["nn0_new_coord"], new_conn=c["nn0_new_conn"])
        a.update(over)
        m, arr, out, keep = core._transfer_args(a["old_coord"], a["old_conn"], a["old_size"], a["new_coord"],
                                                a["new_conn"], flds, lambda m: m.old_ncell, lambda m: m.new_ncell)
        return L.des_remesh_transfer_cells(0, C.byref(m), 0, 1, arr, None, None)

    bad = c["nn0_old_conn"].copy(); bad[5] = c["nn0_old_coord"].size // 3     # node id out of range
    assert cells(old_conn=bad) == 3
    bad = c["nn0_new_conn"].copy(); bad[0] = -1
    assert cells(new_conn=bad) == 3
    assert L.des_remesh_transfer_cells(0, None, 0, 0, None, None, None) == 3
    m, arr, out, keep = core._transfer_args(c["brc_old_coord"], c["brc_old_conn"], c["brc_old_size"],
                                            c["brc_new_coord"], None, {}, lambda m: m.old_nnode, lambda m: m.new_nnode)
    sidx = c["brc_old_support_idx"].copy(); sarr = c["brc_old_support_arr"].copy()
    IP = C.POINTER(C.c_int)
    assert L.des_remesh_transfer_nodes(0, C.byref(m), None, None, 0, None, None, None, None) == 3
    sarr[3] = c["brc_old_conn"].size // 4                                      # element id out of range
    assert L.des_remesh_transfer_nodes(0, C.byref(m), sidx.ctypes.data_as(IP), sarr.ctypes.data_as(IP),
                                       0, None, None, None, None) == 3


def test_bridge_wrap_file_through_the_reference_remesh(tmp_path):
    """bridge/des_transfer_wrap.cxx (which members of Variables each call replaces) inside the
    reference's own remesh(): oracle/_ref/libdesref_wrapcheck.so links that file with the transfer
    entry points answered by host loops over the library's per-item routines.  The state after the
    remesh and five more steps equals the unmodified reference's, every array bit for bit."""
    import sys
    wc = os.path.join(os.path.dirname(tc.REFLIB), "libdesref_wrapcheck.so")
    if not (os.path.exists(tc.REFLIB) and os.path.exists(wc)):
        pytest.skip("oracle/_ref not built")
    cfg = os.path.join(tc.HERE, "cfg", "t3d-remesh-coarse.cfg")
    script = os.path.join(tc.HERE, "transfer_wrapcheck.py")
    outs, procs = [], []
    for extra in ([], ["--wrapcheck"]):
        out = str(tmp_path / ("wc%d.npz" % len(outs)))
        outs.append(out)
        procs.append(subprocess.Popen([sys.executable, script, cfg, "300", out] + extra,
                                      stdout=subprocess.PIPE, stderr=subprocess.STDOUT,
                                      env=dict(os.environ, OMP_NUM_THREADS="4")))
    logs = [p.communicate(timeout=600)[0].decode() for p in procs]
    assert all(p.returncode == 0 for p in procs), logs
    assert "Element ACM" in logs[0] and "Element ACM" not in logs[1]      # the second run bypassed nn-interpolation.cxx
    a, b = np.load(outs[0]), np.load(outs[1])
    for k in a.files:
        assert np.array_equal(a[k], b[k]), k


def test_checker_code_is_not_in_the_product_binaries():
    """The host loops of tests/hostcheck exist only in the checker libraries: neither libdes_b200.so
    nor the bridge binary defines or imports a hostcheck symbol, and the library's transfer entry
    points come from its own CUDA translation unit."""
    lib = core.LIB_PATH
    out = subprocess.run(["nm", "-D", "--defined-only", lib], capture_output=True, text=True, check=True).stdout
    assert "des_remesh_transfer_cells" in out and "hostcheck" not in out
    b200 = os.path.join(tc.ROOT, "bridge", "_build", "dynearthsol3d.b200")
    if os.path.exists(b200):
        syms = subprocess.run(["nm", b200], capture_output=True, text=True, check=True).stdout
        assert "hostcheck" not in syms
        assert " U des_remesh_transfer_cells" in syms       # resolved by libdes_b200.so at load time


def test_transfer_has_no_cpu_path():
    """Valid arguments and no CUDA device: DES_E_NODEVICE, never a result computed on the host."""
    from conftest import have_gpu
    if have_gpu():
        pytest.skip("a GPU is present")
    c = tc.load_case(tc.GOLDEN)
    with pytest.raises(core.DesError) as ei:
        core.remesh_transfer_cells(c["nn0_old_coord"], c["nn0_old_conn"], c["nn0_old_size"], c["nn0_new_coord"],
                                   c["nn0_new_conn"], {"plstrain": c["nn0_src_plstrain"]})
    assert ei.value.code == 1
    with pytest.raises(core.DesError) as ei:
        core.remesh_transfer_nodes(c["brc_old_coord"], c["brc_old_conn"], c["brc_old_size"], c["brc_old_support_idx"],
                                   c["brc_old_support_arr"], c["brc_new_coord"], {"temperature": c["brc_src_temperature"]})
    assert ei.value.code == 1
