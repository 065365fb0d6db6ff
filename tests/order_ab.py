"""Node + element ordering A/B on a general (tetgen) mesh -- row N1 of the round-1 verdict.

The reference hands the step loop a mesh renumbered along its longest axis
(renumbering_mesh, mesh.cxx:2696-2821).  Here the SAME mesh is stepped in three orders, every
node- and element-indexed array renumbered on the host before upload (so no library change is
needed to measure it):
    delivered   as the reference produced it
    morton      nodes by the Morton code of their coordinates, elements by the Morton code
                of their centroid
    random      a random permutation of both (what an unsorted mesh would cost)
Per order: ms/step and the per-kernel times (CUDA events around every launch).

    python tests/order_ab.py abtest/tetgen_1M.npz out.json [steps]
"""
import json
import sys
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from dynearthsol_b200 import boxmesh as bm, core


def morton3(x, y, z, bits=21):
    def spread(v):
        v = v.astype(np.uint64) & np.uint64((1 << bits) - 1)
        v = (v | (v << np.uint64(32))) & np.uint64(0x1f00000000ffff)
        v = (v | (v << np.uint64(16))) & np.uint64(0x1f0000ff0000ff)
        v = (v | (v << np.uint64(8))) & np.uint64(0x100f00f00f00f00f)
        v = (v | (v << np.uint64(4))) & np.uint64(0x10c30c30c30c30c3)
        v = (v | (v << np.uint64(2))) & np.uint64(0x1249249249249249)
        return v
    return spread(x) | (spread(y) << np.uint64(1)) | (spread(z) << np.uint64(2))


def quantise(c):
    lo, hi = c.min(axis=1, keepdims=True), c.max(axis=1, keepdims=True)
    return np.floor((c - lo) / np.maximum(hi - lo, 1e-300) * (2 ** 21 - 1)).astype(np.uint64)


def renumber(coord, conn, segment, nperm, eperm):
    """new node i = old node nperm[i]; new element j = old element eperm[j]"""
    inv = np.empty(nperm.size, dtype=np.int32)
    inv[nperm] = np.arange(nperm.size, dtype=np.int32)
    return (np.ascontiguousarray(coord[:, nperm]), np.ascontiguousarray(inv[conn][:, eperm]),
            np.ascontiguousarray(inv[segment]))


def run(coord, conn, segment, segflag, steps, reorder=0):
    N, E = coord.shape[1], conn.shape[1]
    mesh = core.build_topology(coord, conn, segment, segflag)
    xl, yl, zl = [float(coord[d].max() - coord[d].min()) for d in range(3)]
    phys = dict(bm.TEST3D_PHYSICS); phys["zlength"] = zl
    state, comp_p = bm.initial_fields(coord, conn, rho0=phys["rho0"][0], bulk_modulus=phys["bulk_modulus"][0],
                                      gravity=phys["gravity"], xlength=xl, ylength=yl, zlength=zl,
                                      resolution=(xl * yl * zl / E * 6) ** (1 / 3.0), weakzone=None)
    g = core.DesGpu(core.make_param(**phys), 0)
    g.set_reorder(reorder)
    mesh["coord"] = coord
    mesh["bnormals"] = np.zeros(30); mesh["edge_slot"] = np.full(100, -1, dtype=np.int32)
    g.upload_mesh(mesh)
    g.upload_elemmarkers(np.full(E, 4, dtype=np.int32))
    g.set_compensation_pressure(comp_p)
    g.upload_state(**state)
    g.call("apply_vbcs", 0.0); g.call("compute_mass")
    dt = g.compute_dt()[0]
    dt, t, s = g.run(5, dt)
    g.sync()
    g.event_record(0)
    dt, t, s = g.run(steps, dt, t, s)
    g.event_record(1)
    ms = g.event_elapsed_ms(0, 1) / steps
    g.enable_kernel_timing(True)
    g.run(10, dt, t, s)
    kt = {name.split("(")[0]: m / n for name, m, n in g.kernel_times() if n}
    g.enable_kernel_timing(False)
    chk = g.download_state("vel")["vel"]
    g.close()
    return {"ms_per_step": ms, "kernels_avg_ms": kt, "vel_abs_sum": float(np.abs(chk).sum())}


def main():
    d = np.load(sys.argv[1])
    out = sys.argv[2]
    steps = int(sys.argv[3]) if len(sys.argv) > 3 else 40
    coord, conn = d["coord"], d["conn"]
    N, E = coord.shape[1], conn.shape[1]
    segment = d["segment"].reshape(3, -1); segflag = d["segflag"].reshape(-1)
    res = {"mesh": "%s: %d nodes, %d tets, %d boundary segments" % (os.path.basename(sys.argv[1]), N, E, segflag.size),
           "steps": steps, "orders": {}}
    cent = (coord[:, conn[0]] + coord[:, conn[1]] + coord[:, conn[2]] + coord[:, conn[3]]) / 4
    qn, qe = quantise(coord), quantise(cent)
    rng = np.random.default_rng(1)
    orders = {
        "delivered": (np.arange(N), np.arange(E)),
        "morton": (np.argsort(morton3(qn[0], qn[1], qn[2]), kind="stable"), np.argsort(morton3(qe[0], qe[1], qe[2]), kind="stable")),
        "random": (rng.permutation(N), rng.permutation(E)),
    }
    for name, (npm, epm) in orders.items():
        c, cn, sg = renumber(coord, conn, segment, npm, epm)
        res["orders"][name] = run(c, cn, sg, segflag, steps)
        print(name, json.dumps(res["orders"][name]), flush=True)
    # the same, done by the library: the caller keeps the delivered order, des_gpu_set_reorder(h, 2)
    res["orders"]["delivered, stored in Morton order by the library"] = run(coord, conn, segment, segflag, steps, reorder=2)
    print("library", json.dumps(res["orders"]["delivered, stored in Morton order by the library"]), flush=True)
    json.dump(res, open(out, "w"), indent=1)


if __name__ == "__main__":
    main()
