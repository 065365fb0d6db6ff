"""Remesh field transfer at scale, reference vs device, same inputs (run on the GPU box):
the regular box of bench.py at `size`, stepped a few steps so the fields are not trivial, then
transferred onto the same connectivity with every node displaced by up to `jitter` resolutions
(hull nodes kept on their faces) -- every element changes, every one goes through the volume-ratio
mapping (175 sample points each, located among the 32 nearest old elements).  The reference side is
its own nearest_neighbor_interpolation / barycentric_node_interpolation through the harness
(all host threads); the device side is the C-ABI on the recorded inputs; outputs are compared.

usage: python tests/transfer_timing.py [size=100k] [out.json]"""
import json
import os
import sys
import tempfile
import time
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, HERE); sys.path.insert(0, ROOT)
import bench  # noqa: E402
from refharness import Ref  # noqa: E402
from dynearthsol_b200 import core  # noqa: E402


def main():
    size = sys.argv[1] if len(sys.argv) > 1 else "100k"
    out_path = sys.argv[2] if len(sys.argv) > 2 else None
    xl, yl, zl, res = bench.SIZES[size]
    wd = tempfile.mkdtemp(prefix="destt_")
    cfg = os.path.join(wd, "sample.cfg")
    open(cfg, "w").write(bench.CFG_TEMPLATE % dict(xl=xl, yl=yl, zl=zl, res=res))
    t0 = time.time()
    r = Ref(cfg, workdir=wd, fast_init=True)
    r.step(3)
    init_s = time.time() - t0
    n, ne = r.int("nnode"), r.int("nelem")
    x = r.get("coord").reshape(3, n).copy()
    conn = r.get("connectivity").reshape(4, ne)
    flag = r.get("bcflag")
    rng = np.random.default_rng(5)
    d = rng.uniform(-0.2, 0.2, (3, n)) * res
    for axis, bits in ((0, 1 | 2), (1, 4 | 8), (2, 16 | 32)):      # BOUNDX0|X1, Y0|Y1, Z0|Z1 stay on their planes
        d[axis, (flag & bits) != 0] = 0
    xn = x + d
    device_only = os.environ.get("TT_DEVICE_ONLY") == "1"      # sizes the reference would take minutes for
    if device_only:
        names = ("plstrain", "delta_plstrain", "strain", "radiogenic_source", "volume_old", "stress")
        cells = {"nn0_old_coord": x.ravel(), "nn0_old_conn": conn.ravel(), "nn0_old_size": r.get("volume"),
                 "nn0_new_coord": xn.ravel(), "nn0_new_conn": conn.ravel()}
        cells.update({"nn0_src_" + k: r.get(k) for k in names})
        nodes = {"brc_old_coord": x.ravel(), "brc_old_conn": conn.ravel(), "brc_old_size": r.get("volume"),
                 "brc_old_support_idx": r.get("support_idx"), "brc_old_support_arr": r.get("support_arr"),
                 "brc_new_coord": xn.ravel()}
        nodes.update({"brc_src_" + k: r.get(k) for k in ("temperature", "vel", "coord0")})
        ref_cells_s = ref_nodes_s = None
    else:
        cells, ref_cells_s = r.transfer_cells_captured(xn.ravel(), conn.ravel())
        t0 = time.time()
        nodes = r.transfer_nodes_captured(xn.ravel(), flag)
        ref_nodes_s = time.time() - t0
    res_d = {"size": size, "nelem": ne, "nnode": n, "host_threads": os.cpu_count(), "reference_init_s": init_s,
             "reference_cells_s": ref_cells_s, "reference_nodes_s": ref_nodes_s}
    # device, twice (the first call pays the context)
    f = {k[8:]: v for k, v in cells.items() if k.startswith("nn0_src_")}
    for rep in range(2):
        t0 = time.time()
        out, changed, st = core.remesh_transfer_cells(cells["nn0_old_coord"], cells["nn0_old_conn"], cells["nn0_old_size"],
                                                      cells["nn0_new_coord"], cells["nn0_new_conn"], f, False)
        dev_cells_s = time.time() - t0
    res_d.update(device_cells_s=dev_cells_s, device_cells_search_s=st.search_seconds, device_cells_ratios_s=st.acm_seconds,
                 device_cells_fields_s=st.field_seconds, nchanged=st.nchanged, nsample=st.nsample,
                 sample_points=st.nchanged * st.nsample, nfields=len(f))
    if not device_only:
        res_d["cells_is_changed_equal"] = bool(np.array_equal(changed, cells["nn0_is_changed"]))
        res_d["cells_fields_differing_values"] = {k: int((out[k] != cells["nn0_dst_" + k]).sum()) for k in f}
    nf = {k[8:]: v for k, v in nodes.items() if k.startswith("brc_src_")}
    for rep in range(2):
        t0 = time.time()
        outn, el, brc, stn = core.remesh_transfer_nodes(nodes["brc_old_coord"], nodes["brc_old_conn"], nodes["brc_old_size"],
                                                        nodes["brc_old_support_idx"], nodes["brc_old_support_arr"],
                                                        nodes["brc_new_coord"], nf)
        dev_nodes_s = time.time() - t0
        if rep == 0:
            res_d["device_nodes_first_call_s"] = dev_nodes_s
    res_d.update(device_nodes_s=dev_nodes_s, device_nodes_locate_s=stn.search_seconds,
                 device_nodes_wider_search_s=stn.acm_seconds, device_nodes_fields_s=stn.field_seconds,
                 nodes_by_wider_search=stn.nchanged, nodes_outside=stn.n_outside)
    if not device_only:
        res_d["nodes_fields_differing_values"] = {k: int((outn[k] != nodes["brc_dst_" + k]).sum()) for k in nf}
        res_d["speedup_cells"] = ref_cells_s / dev_cells_s
        res_d["speedup_nodes"] = ref_nodes_s / res_d["device_nodes_s"]
    print(json.dumps(res_d))
    if out_path:
        json.dump(res_d, open(out_path, "w"), indent=1)


if __name__ == "__main__":
    main()
