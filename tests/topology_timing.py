"""Wall time of the topology builders on a structured box: the device
create_support against the library's host helper and the numpy builder, and the boundary-facet
search through the support lists (the reference's own search is O(nseg * nelem))."""
import ctypes as C
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from dynearthsol_b200 import boxmesh, core

dims = {"1M": (78, 52, 52), "6M": (142, 95, 95), "50M": (286, 191, 191)}[sys.argv[1] if len(sys.argv) > 1 else "6M"]
coord, conn, bcflag = boxmesh.regular_box(150e3, 100e3, 100e3, *dims)
conn = np.ascontiguousarray(conn, dtype=np.int32)
N, E = coord.shape[1], conn.shape[1]
L = core.load_library()
out = {"nnode": N, "nelem": E}
idx = np.empty(N + 1, dtype=np.int32); arr = np.empty(4 * E, dtype=np.int32); lidx = np.empty(4 * E, dtype=np.int32)
args = (N, E, core._ptr(conn, C.c_int), core._ptr(idx, C.c_int), core._ptr(arr, C.c_int), core._ptr(lidx, C.c_int))
L.des_gpu_build_support(0, *args)          # context creation
t = time.time(); rc = L.des_gpu_build_support(0, *args); out["support_device_s"] = time.time() - t
assert rc == 0
t = time.time(); h = core.host_build_support(N, conn); out["support_host_helper_s"] = time.time() - t
assert np.array_equal(h[0], idx) and np.array_equal(h[1], arr) and np.array_equal(h[2], lidx)
t = time.time(); boxmesh.build_support(N, conn); out["support_numpy_s"] = time.time() - t
# boundary facets of the six box faces as segments: take the reference-order facets from the numpy builder
m = boxmesh.boundary_topology(coord, conn, bcflag, support=(idx, arr, lidx))
NOF = np.array([[1, 2, 3], [0, 3, 2], [0, 1, 3], [0, 2, 1]])
seg, flg = [], []
for k in range(6):
    p = m["bfacets%d" % k].reshape(-1, 2)
    seg.append(np.stack([conn[NOF[p[:, 1], j], p[:, 0]] for j in range(3)]))
    flg.append(np.full(p.shape[0], 1 << k, dtype=np.int32))
seg = np.ascontiguousarray(np.concatenate(seg, axis=1), dtype=np.int32); flg = np.concatenate(flg)
S = flg.size
se = np.empty(S, dtype=np.int32); sf = np.empty(S, dtype=np.int32)
t = time.time()
rc = L.des_gpu_find_boundary_facets(0, N, E, S, core._ptr(conn, C.c_int), core._ptr(seg, C.c_int), core._ptr(flg, C.c_int),
                                    core._ptr(bcflag.astype(np.uint32), C.c_uint), core._ptr(idx, C.c_int),
                                    core._ptr(arr, C.c_int), core._ptr(se, C.c_int), core._ptr(sf, C.c_int))
out["facet_search_device_s"] = time.time() - t
assert rc == 0 and (se >= 0).all()
out["nseg"] = int(S)
out["reference_search_pair_tests"] = float(S) * E / 2      # expected (segment, element) tests of mesh.cxx:3191
print(json.dumps(out))
