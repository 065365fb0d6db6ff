"""CPU checks of the drop-in boundary: libdes_b200.so loads, exports every symbol that
include/des_gpu.h declares, agrees with the ctypes mirror on struct layout, refuses to run
without a device (no CPU fallback), and its host-side integer/scalar helpers reproduce the
reference's arrays from the committed golden fixture (tests/golden, made by tests/refdump.py
from the unmodified reference)."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from conftest import ROOT, HERE, have_gpu
from dynearthsol_b200 import core, boxmesh

HEADER = os.path.join(ROOT, "include", "des_gpu.h")
GOLDEN = os.path.join(HERE, "golden", "box-ep-tiny.npz")


def header_symbols():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(des_[a-z0-9_]+)\s*\(", src)))


def test_header_symbols_exported():
    L = core.load_library()
    names = header_symbols()
    assert len(names) >= 30
    for n in names:
        assert hasattr(L, n), "libdes_b200.so does not export %s" % n
    # the python mirror lists the same set
    assert sorted(core.ABI_SYMBOLS) == names


def test_struct_sizes_match():
    L = core.load_library()
    for i, st in enumerate((core.DesParam, core.DesMesh, core.DesState, core.DesDtReductions,
                            core.DesQualityParam, core.DesQuality, core.DesFrame, core.DesTransferMeshes,
                            core.DesTransferField, core.DesTransferStats)):
        assert L.des_abi_sizeof(i) == C.sizeof(st)
    assert L.des_abi_sizeof(99) == -1


def test_no_cpu_fallback():
    """Without a CUDA device the product path must fail loudly (DES_E_NODEVICE)."""
    if have_gpu():
        pytest.skip("a GPU is present")
    with pytest.raises(core.DesError) as ei:
        core.DesGpu(core.make_param(), 0)
    assert ei.value.code == 1


def test_null_arguments_rejected():
    L = core.load_library()
    assert L.des_gpu_create(None, 0, None) == 3
    assert L.des_gpu_step(None, 1.0, 0.0, 0) == 3
    assert L.des_gpu_upload_mesh(None, None) == 3
    assert L.des_gpu_destroy(None) == 0
    assert L.des_host_build_support(0, 0, None, None, None, None) == 3
    # device-resident markers / phase changes: argument checks come before any device work
    assert L.des_gpu_upload_markers(None, 0, 0, None, None, None) == 3
    assert L.des_gpu_phase_changes(None, 1, None, 0, None, None, None, None) == 3
    assert L.des_gpu_download_elemmarkers(None, None) == 3


def test_make_param_rejects_unknown_field():
    with pytest.raises(TypeError):
        core.make_param(not_a_field=1)


def test_host_build_support_matches_reference():
    """create_support (mesh.cxx:3287-3329): bit-identical idx/arr/lidx."""
    d = np.load(GOLDEN)
    N, E = int(d["p_nnode"]), int(d["p_nelem"])
    idx, arr, lidx = core.host_build_support(N, d["mesh_connectivity"])
    assert np.array_equal(idx, d["mesh_support_idx"])
    assert np.array_equal(arr, d["mesh_support_arr"])
    assert np.array_equal(lidx, d["mesh_support_lidx"])
    # and the numpy builder used for the big synthetic boxes
    idx2, arr2, lidx2 = boxmesh.build_support(N, d["mesh_connectivity"].reshape(4, E))
    assert np.array_equal(idx2, idx) and np.array_equal(arr2, arr) and np.array_equal(lidx2, lidx)


def test_host_build_support_rejects_bad_connectivity():
    conn = np.array([[0], [1], [2], [7]], dtype=np.int32)      # node 7 out of range for nnode=4
    with pytest.raises(core.DesError):
        core.host_build_support(4, conn)


def _dt_reductions_numpy(d, prefix):
    """compute_dt's element loop (geometry.cxx:1511-1593) restated in numpy for the nmat=1 box."""
    N, E = int(d["p_nnode"]), int(d["p_nelem"])
    x = d[prefix + "coord"].reshape(3, N)
    v = d[prefix + "vel"].reshape(3, N)
    conn = d["mesh_connectivity"].reshape(4, E)
    p = [x[:, conn[i]] for i in range(4)]

    def area(a, b, c):
        ab, ac = b - a, c - a
        d0 = ab[1] * ac[2] - ab[2] * ac[1]
        d1 = ab[2] * ac[0] - ab[0] * ac[2]
        d2 = ab[0] * ac[1] - ab[1] * ac[0]
        return np.sqrt(d0 * d0 + d1 * d1 + d2 * d2) / 2

    maxa = np.maximum(np.maximum(area(p[0], p[1], p[2]), area(p[0], p[1], p[3])),
                      np.maximum(area(p[2], p[3], p[0]), area(p[2], p[3], p[1])))
    minh = 3 * d[prefix + "volume"] / maxa
    ve = np.zeros((3, E))
    for i in range(4):
        ve += v[:, conn[i]] * 0.25
    r = core.DesDtReductions()
    shearm = float(d["p_shear_modulus"][0])
    rho = float(d["p_rho0"][0])     # alpha = 0 in the fixture
    r.minl = float(minh.min())
    r.dt_maxwell = 0.5 * float(d["p_visc_min"]) / (1e-40 + shearm)
    r.dt_diffusion = float((0.5 * minh * minh / float(d["p_therm_diff_max"])).min())
    r.global_dt_min = float((minh / np.sqrt(shearm / rho) / 5.0).min())
    r.global_max_vem = float(np.sqrt((ve * ve).sum(axis=0)).max())
    r.max_surf_vel = float(d[prefix + "max_surf_vel"])
    return r


@pytest.mark.parametrize("prefix", ["s0_", "s10_"])
def test_dt_from_reductions_matches_reference(prefix):
    """Host scalar algebra of compute_dt (geometry.cxx:1597-1630) against the dt the reference
    computed for the same state (dt is recomputed every 10 steps, dynearthsol.cxx:893)."""
    import refmodel
    d = np.load(GOLDEN)
    L = core.load_library()
    par = refmodel.param_from_dump(d)
    r = _dt_reductions_numpy(d, prefix)
    vmag = C.c_double()
    dt = L.des_dt_from_reductions(C.byref(par), C.byref(r), C.byref(vmag))
    ref = float(d[prefix + "dt"])
    assert abs(dt - ref) <= 1e-14 * ref, (dt, ref)


def test_golden_fixture_is_the_reference():
    """When the reference library is present (this container; it also travels to the GPU box),
    regenerate the fixture live and require the committed file to be bit-identical to it."""
    from conftest import REFLIB, ref_dump
    if not os.path.exists(REFLIB):
        pytest.skip("oracle/_ref/libdesref.so absent")
    live = ref_dump("box-ep-tiny", [1, 10])
    gold = np.load(GOLDEN)
    for k in gold.files:
        assert k in live, k
        a, b = live[k], gold[k]
        if k == "mesh_bnormals":
            # rows n0..n3 have no facets on a box; the reference leaves them uninitialised
            a, b = a.reshape(3, 10)[:, :6], b.reshape(3, 10)[:, :6]
        if k == "mesh_connectivity_surface":
            # conn_t has 4 columns, the reference fills 3 (mesh.cxx:3249-3262)
            a, b = a.reshape(4, -1)[:3], b.reshape(4, -1)[:3]
        # q_minq holds NaN where bad_mesh_quality leaves min_quality unwritten
        same = np.array_equal(a, b, equal_nan=(a.dtype.kind == "f"))
        assert same, "golden fixture differs from the live reference: " + k


def test_quality_verdict_ladder():
    """des_quality_verdict is the decision ladder of bad_mesh_quality (remeshing.cxx:2766-2866):
    fed the reductions that the reference's loops would find, it returns the reference's code,
    index and min_quality for every case recorded in the golden fixture."""
    d = np.load(GOLDEN)
    L = core.load_library()
    N, E = int(d["p_nnode"]), int(d["p_nelem"])
    x = d["s10_coord"].reshape(3, N)
    vol = d["s10_volume"]
    conn = d["mesh_connectivity"].reshape(4, E)
    flag = d["mesh_bcflag"]
    p = [x[:, conn[i]] for i in range(4)]

    def area(a, b, c):
        ab, ac = b - a, c - a
        d0 = ab[1] * ac[2] - ab[2] * ac[1]
        d1 = ab[2] * ac[0] - ab[0] * ac[2]
        d2 = ab[0] * ac[1] - ab[1] * ac[0]
        return np.sqrt(d0 * d0 + d1 * d1 + d2 * d2) / 2

    asum = area(p[0], p[1], p[2]) + area(p[0], p[1], p[3]) + area(p[2], p[3], p[0]) + area(p[2], p[3], p[1])
    quality = 216 * np.sqrt(3.0) * vol * vol / (asum * asum * asum)
    res = float(d["m_resolution"])
    for row, code, index, minq in zip(d["q_params"], d["q_code"], d["q_index"], d["q_minq"]):
        qp = core.DesQualityParam(smallest_size=row[0], resolution=res, max_boundary_distortion=row[1],
                                  min_quality=row[2], xlength=float(d["m_xlength"]),
                                  ylength=float(d["m_ylength"]), zlength=float(d["m_zlength"]),
                                  remeshing_option=int(row[3]))
        q = core.DesQuality()
        small = np.nonzero(vol < row[0] * 0.118 * res ** 3)[0]
        q.first_small_elem = int(small[0]) if len(small) else -1
        dist = row[1] * res
        ro = int(row[3])
        q.first_bottom_node = q.first_side_node = -1
        if ro in (1, 2, 11, 13):
            bad = np.nonzero(((flag & 16) != 0) & (np.abs(x[2] + float(d["m_zlength"])) > dist))[0]
            q.first_bottom_node = int(bad[0]) if len(bad) else -1
        if ro == 13:
            off = np.where(flag & 1, np.abs(x[0]), np.where(flag & 2, np.abs(x[0] - float(d["m_xlength"])),
                  np.where(flag & 4, np.abs(x[1]), np.where(flag & 8, np.abs(x[1] - float(d["m_ylength"])), 0.0))))
            bad = np.nonzero(off > dist)[0]
            q.first_side_node = int(bad[0]) if len(bad) else -1
        w = int(np.argmin(quality)) if quality.min() < 1 else 0
        q.worst_elem, q.worst_quality = w, float(min(quality[w], 1.0))
        idx, mq = C.c_int(-12345), C.c_double(float("nan"))
        got = L.des_quality_verdict(C.byref(qp), C.byref(q), C.byref(idx), C.byref(mq))
        assert (got, idx.value) == (int(code), int(index)), (row, got, idx.value, code, index)
        assert (np.isnan(minq) and np.isnan(mq.value)) or mq.value == float(minq), (row, mq.value, minq)


def test_host_topology_builders_match_reference():
    """create_boundary_flags / create_boundary_nodes / create_top_elems / create_surface_info
    (mesh.cxx:2824-3098) through the library's host builders: bit-identical integer tables."""
    d = np.load(GOLDEN)
    N, E = int(d["p_nnode"]), int(d["p_nelem"])
    bc = core.host_boundary_flags(N, d["mesh_segment"], d["mesh_segflag"])
    assert np.array_equal(bc, d["mesh_bcflag"])
    bn = core.host_boundary_nodes(bc)
    for k in range(10):
        assert np.array_equal(bn[k], d["mesh_bnodes%d" % k]), k
    s = core.host_surface_topology(N, bn[5], d["s0_coord"][:N], d["mesh_support_idx"], d["mesh_support_arr"],
                                   d["mesh_connectivity_surface"])
    for k in ("top_nodes", "top_elems", "elem_and_nodes", "support_surf_idx", "support_surf_arr"):
        assert np.array_equal(s[k], d["mesh_" + k]), k


def test_new_entry_points_reject_bad_arguments_and_have_no_cpu_path():
    """The device-side builders / reductions fail loudly: NULL arguments -> DES_E_ARG, and on a
    box without a GPU the two device builders return DES_E_NODEVICE instead of computing on
    the host."""
    L = core.load_library()
    assert L.des_gpu_mesh_quality(None, None, None) == 3
    assert L.des_gpu_average_fields(None, 0) == 3
    assert L.des_gpu_download_averages(None, None, None, None, None) == 3
    assert L.des_gpu_set_interior(None, 0) == 3
    assert L.des_gpu_step_segment_part(None, 1, 0, 1.0, 0.0, None) == 3
    assert L.des_gpu_isostasy_constrain_velocity(None) == 3
    assert L.des_gpu_check_nan(None, None) == 3
    assert L.des_gpu_set_reorder(None, 2) == 3
    assert L.des_gpu_partition_info(None, None, None, None, None, None, None) == 3
    assert L.des_gpu_create_multi(None, 2, None, None) == 3
    p = core.make_param()
    hh = C.c_void_p()
    assert L.des_gpu_create_multi(C.byref(p), 0, None, C.byref(hh)) == 3          # no devices asked for
    if not have_gpu():
        assert L.des_gpu_create_multi(C.byref(p), 2, None, C.byref(hh)) == 1      # no CPU stand-in for a device group
    assert L.des_gpu_build_support(0, 4, 1, None, None, None, None) == 3
    assert L.des_host_boundary_flags(4, 1, None, None, None) == 3
    # out-of-range segment node
    seg = np.array([0, 1, 9], dtype=np.int32); flg = np.array([1], dtype=np.int32); bc = np.zeros(4, dtype=np.uint32)
    assert L.des_host_boundary_flags(4, 1, core._ptr(seg, C.c_int), core._ptr(flg, C.c_int), core._ptr(bc, C.c_uint)) == 3
    if not have_gpu():
        conn = np.array([[0], [1], [2], [3]], dtype=np.int32)
        idx = np.zeros(5, dtype=np.int32); arr = np.zeros(4, dtype=np.int32); lidx = np.zeros(4, dtype=np.int32)
        rc = L.des_gpu_build_support(0, 4, 1, core._ptr(conn, C.c_int), core._ptr(idx, C.c_int),
                                     core._ptr(arr, C.c_int), core._ptr(lidx, C.c_int))
        assert rc == 1


def test_reference_fast_init_equals_stock_init(tmp_path):
    """bench.py's CPU arm initialises large boxes with desref_fast_init(1): the boundary-facet
    search walks the support of the segment's first node instead of all elements
    (oracle/ref_harness.cxx).  Same arrays as the reference's own init(), bit for bit."""
    import subprocess
    import sys
    from refharness import available
    if not available():
        pytest.skip("oracle/_ref/libdesref.so not built")
    names = ["connectivity", "coord", "bcflag", "support_arr", "support_idx", "support_lidx", "top_elems",
             "top_nodes", "elem_and_nodes", "connectivity_surface", "support_surf_arr", "support_surf_idx",
             "stress", "plstrain", "temperature", "volume", "mass", "tmass", "volume_n", "vel"] \
        + ["bfacets%d" % i for i in range(10)] + ["bnodes%d" % i for i in range(10)]
    code = ("import sys; sys.path.insert(0, %r); import numpy as np\n"
            "from refharness import Ref\n"
            "r = Ref(%r, workdir=%r, fast_init=bool(int(sys.argv[1])))\n"
            "r.step(2)\n"
            "np.savez(sys.argv[2], dt=r.scalar('dt'), **{k: r.get(k) for k in %r})\n"
            % (HERE, os.path.join(HERE, "cfg", "box-ep.cfg"), str(tmp_path), names))
    out = []
    for fast in (0, 1):
        fn = str(tmp_path / ("init%d.npz" % fast))
        env = dict(os.environ, OMP_NUM_THREADS="1")     # the stock bucket order is thread-dependent
        subprocess.run([sys.executable, "-c", code, str(fast), fn], check=True, env=env,
                       capture_output=True, cwd=str(tmp_path))
        out.append(np.load(fn))
    a, b = out
    for k in names + ["dt"]:
        x, y = a[k], b[k]
        if k == "connectivity_surface":     # conn_t with 4 columns, the 4th is never written
            x, y = x[:3 * x.size // 4], y[:3 * y.size // 4]
        assert np.array_equal(x, y), k
