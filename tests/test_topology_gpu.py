"""the topology builders (device create_support and boundary-facet
search + host bucketing / surface tables) against the reference's own tables for the same
connectivity and boundary segments -- structured boxes and a tetgen mesh.  Integer work:
bit-exact."""
import numpy as np
import pytest

from conftest import ref_dump
from dynearthsol_b200 import core

pytestmark = pytest.mark.gpu

KEYS = ["bcflag", "support_idx", "support_arr", "support_lidx", "top_nodes", "top_elems",
        "elem_and_nodes", "support_surf_idx", "support_surf_arr"] + \
       ["bfacets%d" % i for i in range(10)] + ["bnodes%d" % i for i in range(10)]


@pytest.mark.parametrize("case", ["box-ep-tiny", "box-ep", "t3d", "orift"])
def test_build_topology_matches_reference(case):
    d = ref_dump(case, [1])
    N, E = int(d["p_nnode"]), int(d["p_nelem"])
    S = d["mesh_segflag"].size
    m = core.build_topology(d["s0_coord"].reshape(3, N), d["mesh_connectivity"].reshape(4, E),
                            d["mesh_segment"].reshape(3, S), d["mesh_segflag"])
    for k in KEYS:
        assert np.array_equal(np.asarray(m[k]).reshape(-1), d["mesh_" + k].reshape(-1)), (case, k)
    etop = m["bfacets5"].size // 2
    assert np.array_equal(m["connectivity_surface"].reshape(4, etop)[:3],
                          d["mesh_connectivity_surface"].reshape(4, etop)[:3])
    # every boundary segment was found, on the element the reference found
    assert (m["seg_elem"][(d["mesh_segflag"] & 1023) != 0] >= 0).all()


def test_build_support_device_equals_host_helper():
    """des_gpu_build_support == des_host_build_support on a 0.5 M-tet box (ragged patches of
    8..32 elements), and the boundary-facet search finds each box face's facets."""
    from dynearthsol_b200 import boxmesh
    coord, conn, bcflag = boxmesh.regular_box(150e3, 100e3, 100e3, 61, 41, 41)
    N, E = coord.shape[1], conn.shape[1]
    L = core.load_library()
    import ctypes as C
    idx = np.empty(N + 1, dtype=np.int32); arr = np.empty(4 * E, dtype=np.int32); lidx = np.empty(4 * E, dtype=np.int32)
    conn = np.ascontiguousarray(conn, dtype=np.int32)
    rc = L.des_gpu_build_support(0, N, E, core._ptr(conn, C.c_int), core._ptr(idx, C.c_int),
                                 core._ptr(arr, C.c_int), core._ptr(lidx, C.c_int))
    assert rc == 0
    hidx, harr, hlidx = core.host_build_support(N, conn)
    assert np.array_equal(idx, hidx) and np.array_equal(arr, harr) and np.array_equal(lidx, hlidx)
    # out-of-range connectivity is rejected, not dereferenced
    bad = conn.copy(); bad[2, 17] = N + 5
    assert L.des_gpu_build_support(0, N, E, core._ptr(bad, C.c_int), core._ptr(idx, C.c_int),
                                   core._ptr(arr, C.c_int), core._ptr(lidx, C.c_int)) == 3
