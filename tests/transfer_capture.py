"""Record what the UNMODIFIED reference's remesh field transfer does (oracle/_ref/libdesref.so):
  * a real remesh() cycle of a cfg after `steps` steps -- the inputs and outputs of its
    nearest_neighbor_interpolation (elements, top surface) and barycentric_node_interpolation;
  * a synthetic node case on the stepped (pre-remesh) mesh: points displaced so that the capped
    breadth-first search, the on-a-face and the outside-the-domain branches of
    prepare_interpolation (brc-interpolation.cxx:226-329) are all taken ("syn_" entries).
This is the script that generated tests/golden/transfer-*.npz; where libdesref.so is present the
tests call it live (one subprocess per case: the reference keeps process-wide state).

usage: python tests/transfer_capture.py CFG STEPS OUT.npz [--slim]
"""
import os
import sys
import tempfile
import numpy as np

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from refharness import Ref  # noqa: E402

SLIM_DROP = ("delta_plstrain", "radiogenic_source", "dyn_fric_coeff", "state_variable", "stressyy",
             "spr_p_ref_old", "ppressure", "dppressure")


def synthetic_points(r, seed=7):
    """Points that take every branch of prepare_interpolation (brc-interpolation.cxx:183-329):
    old nodes left where they are (the dist == 0 shortcut); random points of the domain, among them
    the ones whose enclosing old element is NOT around their nearest old node (found only by the
    breadth-first search); top-surface nodes raised by 2e-11..3e-10 of the resolution (just outside
    every element: the on-a-face band) or by up to 0.2 resolutions (outside the old domain)."""
    rng = np.random.default_rng(seed)
    n, ne = r.int("nnode"), r.int("nelem")
    x = r.get("coord").reshape(3, n)
    conn = r.get("connectivity").reshape(4, ne)
    flag = r.get("bcflag")
    res = r.mesh_param("resolution")
    keep = x[:, rng.random(n) < 0.3]
    lo, hi = x.min(axis=1), x.max(axis=1)
    p = lo[:, None] + rng.random((3, 6000)) * (hi - lo)[:, None]
    # barycentric coordinates of every point in every old element (plain numpy, only to pick points)
    a = x[:, conn[0]]
    T = np.stack([x[:, conn[1]] - a, x[:, conn[2]] - a, x[:, conn[3]] - a], axis=-1).transpose(1, 0, 2)
    lam = np.einsum("eij,jep->eip", np.linalg.inv(T), p[:, None, :] - a[:, :, None])
    inside = (lam.min(axis=1) > 1e-6) & (lam.sum(axis=1) < 1 - 1e-6)        # [ne, npts]
    has = inside.any(axis=0)
    owner = inside.argmax(axis=0)
    nearest = ((p[:, None, :] - x[:, :, None]) ** 2).sum(axis=0).argmin(axis=0)
    in_ring = (conn[:, owner] == nearest[None, :]).any(axis=0)
    hard = np.flatnonzero(has & ~in_ring)[:80]
    easy = np.flatnonzero(has & in_ring)[:150]
    top = np.flatnonzero((flag & (1 << 5)) != 0)          # BOUNDZ1
    # points of top facets whose nearest old node is not a node of the facet's element, raised until
    # their lowest barycentric coordinate in that element is -6e-11 .. -9e-11: outside is_inside's
    # 5e-11, inside the 1e-10 the search accepts afterwards (brc-interpolation.cxx:301-305)
    pairs = r.get("bfacets5").reshape(-1, 2)
    etop = pairs.shape[0]
    cs = r.get("connectivity_surface").reshape(4, etop)[:3]
    w = rng.dirichlet(np.ones(3) * 0.7, size=(etop, 12))                       # [etop, 12, 3]
    fp = np.einsum("dkf,fsk->dfs", x[:, cs], w).reshape(3, -1)                 # [3, etop*12]
    fe = np.repeat(pairs[:, 0], 12)
    near = ((fp[:, None, :] - x[:, :, None]) ** 2).sum(axis=0).argmin(axis=0)
    pick = np.flatnonzero(~(conn[:, fe] == near[None, :]).any(axis=0))[:60]
    Ti = np.linalg.inv(T)
    up = np.array([0.0, 0.0, 1.0])
    slope = np.array([min((Ti[e] @ up).min(), -(Ti[e] @ up).sum()) for e in fe[pick]])   # d(min lambda)/dz < 0
    tiny = fp[:, pick].copy()
    tiny[2] += rng.uniform(6e-11, 9e-11, pick.size) / -slope
    far = x[:, top[::2]].copy()
    far[2] += res * rng.uniform(0.01, 0.2, far.shape[1])
    pts = np.concatenate([keep, p[:, hard], p[:, easy], tiny, far], axis=1)
    bc = np.zeros(pts.shape[1], dtype=np.uint32)
    bc[keep.shape[1] + hard.size + easy.size:] = 32
    return np.ascontiguousarray(pts).ravel(), bc


def jittered(r, amount, seed=2):
    """The current mesh with every node moved by up to `amount` resolutions (hull nodes stay on their
    planes): the same connectivity, every centroid displaced."""
    rng = np.random.default_rng(seed)
    n = r.int("nnode")
    x = r.get("coord").reshape(3, n).copy()
    flag = r.get("bcflag")
    d = rng.uniform(-amount, amount, (3, n)) * r.mesh_param("resolution")
    for axis, bits in ((0, 1 | 2), (1, 4 | 8), (2, 16 | 32)):
        d[axis, (flag & bits) != 0] = 0
    return (x + d).ravel(), flag


def main():
    cfg, steps, out = sys.argv[1], int(sys.argv[2]), sys.argv[3]
    slim = "--slim" in sys.argv
    work = tempfile.mkdtemp(prefix="destr_")
    r = Ref(cfg, workdir=work)
    r.step(steps)
    if "--jitter" in sys.argv:
        # no remesh: the reference's two transfer calls from the current mesh onto its jittered self
        xn, flag = jittered(r, float(sys.argv[sys.argv.index("--jitter") + 1]))
        cap, _ = r.transfer_cells_captured(xn, r.get("connectivity"))
        cap.update(r.transfer_nodes_captured(xn, flag))
        np.savez_compressed(out, **cap)
        return
    xs, fs = synthetic_points(r)
    syn = r.transfer_nodes_captured(xs, fs)
    cap = {"syn_" + k[4:]: v for k, v in syn.items()}
    cwd = os.getcwd(); os.chdir(work)
    try:
        cap.update(r.remesh_captured(1))
    finally:
        os.chdir(cwd)
    if slim:
        # one copy of the arrays the three calls share, a subset of the fields
        for pre in ("nn1_", "brc_"):
            for nm in ("old_coord", "new_coord"):
                assert np.array_equal(cap[pre + nm], cap["nn0_" + nm])
                del cap[pre + nm]
        assert np.array_equal(cap["brc_old_conn"], cap["nn0_old_conn"]); del cap["brc_old_conn"]
        assert np.array_equal(cap["brc_old_size"], cap["nn0_old_size"]); del cap["brc_old_size"]
        for k in list(cap):
            if any(k.endswith("src_" + d) or k.endswith("dst_" + d) for d in SLIM_DROP):
                del cap[k]
    np.savez_compressed(out, **cap)
    print("captured", {k: v.size for k, v in cap.items()})


if __name__ == "__main__":
    main()
