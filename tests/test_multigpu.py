"""Multi-GPU host logic: partition invariants (CPU), the ghost exchange over torch.distributed
with the gloo backend at world_size 2 (CPU), and -- on a GPU -- a two-part run emulated on one
device and a real 2-rank NCCL run, both required to be BIT-IDENTICAL to the single-GPU run."""
import os
import socket
import subprocess
import sys

import numpy as np
import pytest

from conftest import ROOT, HERE
from dynearthsol_b200 import boxmesh as bm, partition as pt

BOX = (24e3, 10e3, 8e3, 2e3)        # 12 x 5 x 4 cells, 1200 tets


def global_box():
    xl, yl, zl, res = BOX
    nx, ny, nz = bm.grid_dims(xl, yl, zl, res)
    coord, conn, bcflag = bm.regular_box(xl, yl, zl, nx, ny, nz)
    return (nx, ny, nz), coord, conn, bcflag


@pytest.mark.parametrize("nparts", [1, 2, 3, 4, 8])
def test_partition_invariants(nparts):
    (nx, ny, nz), coord, conn, bcflag = global_box()
    parts = pt.partition_mesh(coord, conn, bcflag, nparts)
    N, E = coord.shape[1], conn.shape[1]
    owned_count = np.zeros(N, dtype=int)
    elem_cover = np.zeros(E, dtype=int)
    for p in parts:
        assert np.all(np.diff(p["node_gid"]) > 0) and np.all(np.diff(p["elem_gid"]) > 0)   # global order kept
        owned_count[p["node_gid"][p["owned"]]] += 1
        elem_cover[p["elem_gid"]] += 1
        # an owned node sees its complete support locally
        gsup = np.bincount(conn.reshape(-1), minlength=N)
        lsup = np.bincount(p["conn"].reshape(-1), minlength=p["node_gid"].size)
        assert np.array_equal(lsup[p["owned"]], gsup[p["node_gid"][p["owned"]]])
        # local connectivity maps back to the global one
        assert np.array_equal(p["node_gid"][p["conn"]], conn[:, p["elem_gid"]])
    assert np.all(owned_count == 1)
    assert np.all(elem_cover >= 1)
    # message lists agree pairwise, in order, in global ids; every ghost is received once
    for a in parts:
        pa = a["plan"]
        ghosts = np.nonzero(~a["owned"])[0]
        assert sorted(pa.recv_nodes.tolist()) == sorted(ghosts.tolist())
        for q, rb in enumerate(pa.neighbours):
            b = parts[rb]
            pb = b["plan"]
            qb = pb.neighbours.index(a["rank"])
            sa = pa.send_nodes[pa.send_offsets[q]:pa.send_offsets[q + 1]]
            rbn = pb.recv_nodes[pb.recv_offsets[qb]:pb.recv_offsets[qb + 1]]
            assert np.array_equal(a["node_gid"][sa], b["node_gid"][rbn])
            assert np.all(a["owned"][sa])


def test_direct_box_part_is_the_restriction_of_the_global_box():
    """regular_box_part (what the 200M-tet runs use: no global mesh) == RCB restriction."""
    (nx, ny, nz), coord, conn, bcflag = global_box()
    xl, yl, zl, res = BOX
    ii = np.rint(coord[0] / (xl / (nx - 1))).astype(np.int64)
    jj = np.rint(coord[1] / (yl / (ny - 1))).astype(np.int64)
    kk = np.rint(coord[2] / (-zl / (nz - 1))).astype(np.int64)
    gkey = (ii * ny + jj) * nz + kk
    for nparts in (2, 4, 8):
        boxes = pt.grid_boxes(nx, ny, nz, (xl, yl, zl), nparts)
        gowner = pt.owner_of_grid_nodes(ii, jj, kk, boxes)
        assert (gowner >= 0).all()
        for r in range(nparts):
            c, e, b, key, own = bm.regular_box_part(xl, yl, zl, nx, ny, nz, r, boxes)
            eg, ng, cl = pt.local_subset(conn, gowner, r)
            assert np.array_equal(key, gkey[ng]) and np.array_equal(c, coord[:, ng])
            assert np.array_equal(b, bcflag[ng]) and np.array_equal(own, gowner[ng])
            assert np.array_equal(e, cl)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


GLOO_WORKER = r"""
import os, sys
sys.path.insert(0, %(root)r)
import numpy as np, torch, torch.distributed as dist
from dynearthsol_b200 import boxmesh as bm, partition as pt
from dynearthsol_b200.multigpu import HaloExchanger, HALO_WIDTH
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
dist.init_process_group("gloo", rank=rank, world_size=world)
xl, yl, zl, res = %(box)r
nx, ny, nz = bm.grid_dims(xl, yl, zl, res)
boxes = pt.grid_boxes(nx, ny, nz, (xl, yl, zl), world)
coord, conn, bcflag, key, owner = bm.regular_box_part(xl, yl, zl, nx, ny, nz, rank, boxes)
plan = pt.halo_plan(rank, conn, owner, key)
send = torch.zeros(plan.send_nodes.size * 7, dtype=torch.float64)
recv = torch.zeros(plan.recv_nodes.size * 7, dtype=torch.float64)
x = HaloExchanger(plan, send, recv)
owned = owner == rank
for phase, w in enumerate(HALO_WIDTH):
    # a field every rank can evaluate from the global key: value f of node `key` is w numbers
    def field(k):
        return (k[:, None] * 1.5 + np.arange(w)[None, :] * 0.25 + phase).astype(np.float64)
    vals = np.full((key.size, w), np.nan)
    vals[owned] = field(key[owned])
    send[:plan.send_nodes.size * w] = torch.from_numpy(vals[plan.send_nodes].reshape(-1))    # pack
    x.exchange(w)
    vals[plan.recv_nodes] = recv[:plan.recv_nodes.size * w].numpy().reshape(-1, w)           # unpack
    assert not np.isnan(vals).any(), "a ghost was not received"
    assert np.array_equal(vals, field(key)), "ghost values differ from the owner's"
# the dt reduction: min over ranks
t = torch.tensor([float(rank + 3)], dtype=torch.float64)
dist.all_reduce(t, op=dist.ReduceOp.MIN)
assert t.item() == 3.0
dist.barrier()
print("RANK_OK", rank, plan.neighbours, plan.send_counts, plan.recv_counts)
"""


@pytest.mark.parametrize("world", [2, 4])
def test_gloo_ghost_exchange(world, tmp_path):
    """world_size-2 (and 4) CPU run of the N>1 exchange path on the gloo backend."""
    script = tmp_path / "worker.py"
    script.write_text(GLOO_WORKER % dict(root=ROOT, box=BOX))
    port = _free_port()
    procs = []
    for r in range(world):
        env = dict(os.environ, RANK=str(r), WORLD_SIZE=str(world), MASTER_ADDR="127.0.0.1",
                   MASTER_PORT=str(port), OMP_NUM_THREADS="1")
        procs.append(subprocess.Popen([sys.executable, str(script)], env=env, stdout=subprocess.PIPE,
                                      stderr=subprocess.STDOUT, text=True))
    outs = [p.communicate(timeout=300)[0] for p in procs]
    for r, (p, o) in enumerate(zip(procs, outs)):
        assert p.returncode == 0 and "RANK_OK %d" % r in o, o[-2000:]


# ------------------------------------------------------------------------------------------
FIELDS = ["coord", "vel", "temperature", "stress", "strain", "strain_rate", "plstrain",
          "delta_plstrain", "volume", "volume_old", "force", "volume_n", "mass", "tmass", "dhacc"]


def _problem():
    from dynearthsol_b200 import core
    xl, yl, zl, res = BOX
    (nx, ny, nz), coord, conn, bcflag = global_box()
    phys = dict(bm.TEST3D_PHYSICS); phys["zlength"] = zl
    state, comp_p = bm.initial_fields(coord, conn, rho0=2700.0, bulk_modulus=50e9, gravity=10.0, xlength=xl,
                                      ylength=yl, zlength=zl, resolution=res, weakzone=bm.TEST3D_WEAKZONE)
    return coord, conn, bcflag, phys, state, comp_p


def _owned_merge(parts, outs, N, E):
    """Global arrays from the per-rank downloads: node fields from the owner, element fields
    from any rank holding the element (they must agree -- checked)."""
    merged = {}
    for k in FIELDS:
        per = outs[0][k].size
        p0 = parts[0]
        nl, el = p0["node_gid"].size, p0["elem_gid"].size
        if per in (nl, 3 * nl):
            c = per // nl
            g = np.full((c, N), np.nan)
            for p, o in zip(parts, outs):
                a = o[k].reshape(c, -1)
                g[:, p["node_gid"][p["owned"]]] = a[:, p["owned"]]
        else:
            c = per // el
            g = np.full((c, E), np.nan)
            for p, o in zip(parts, outs):
                a = o[k].reshape(c, -1)
                prev = g[:, p["elem_gid"]]
                seen = ~np.isnan(prev[0])
                assert np.array_equal(prev[:, seen], a[:, seen]), "redundant elements disagree: " + k
                g[:, p["elem_gid"]] = a
        assert not np.isnan(g).any(), k
        merged[k] = g.reshape(-1)
    return merged


@pytest.mark.gpu
@pytest.mark.parametrize("nparts,overlap", [(2, False), (3, False), (2, True), (3, True)])
def test_parts_on_one_gpu_bit_identical(nparts, overlap):
    """Ranks emulated sequentially on one device (no kernel waits on another): the segments,
    pack/unpack kernels and the exchange plan give the single-handle result bit for bit.
    overlap: every part stores its ghost-free elements first and runs E-B / E-C over them
    BEFORE the exchange they overlap with in the real run delivers the ghost values
    (des_gpu_set_interior, des_gpu_step_segment_part)."""
    import torch
    from dynearthsol_b200 import core, multigpu as mg
    coord, conn, bcflag, phys, state, comp_p = _problem()
    N, E = coord.shape[1], conn.shape[1]
    mesh = mg.local_mesh(coord, conn, bcflag)
    g1 = mg.init_model(mesh, phys, state, comp_p, 0)
    g1.call("apply_vbcs", 0.0); g1.call("compute_mass")
    dt0 = g1.compute_dt()[0]
    CHECK = (1, 5, 10, 12)
    refs, dts = {}, {}
    dt1, t1, s1 = dt0, 0.0, 0
    for s in CHECK:
        dt1, t1, s1 = g1.run(s - s1, dt1, t1, s1)
        refs[s] = g1.download_state(*FIELDS)
        dts[s] = dt1
    g1.close()

    parts = pt.partition_mesh(coord, conn, bcflag, nparts)
    dev = torch.device("cuda", 0)
    gs, bufs = [], []
    for p in parts:
        perm, nint = (pt.interior_first(p["conn"], p["owner"], p["rank"]) if overlap else (None, 0))
        if overlap:
            p["elem_gid"] = p["elem_gid"][perm]
        st = pt.restrict_state(state, p["elem_gid"], p["node_gid"], N, E)
        g = mg.init_model(mg.local_mesh(p["coord"], p["conn"], p["bcflag"], perm), phys, st, comp_p, 0)
        if overlap:
            assert 0 < nint < p["conn"].shape[1]
            g.set_interior(nint)
        g.set_halo(p["plan"].send_nodes, p["plan"].recv_nodes)
        sp, rp = g.halo_buffers()
        bufs.append((mg.device_tensor(sp, p["plan"].send_nodes.size * 7, dev),
                     mg.device_tensor(rp, p["plan"].recv_nodes.size * 7, dev)))
        g.call("apply_vbcs", 0.0); g.call("compute_mass")
        gs.append(g)

    def exchange(phase):
        w = mg.HALO_WIDTH[phase]
        for g in gs:
            g.halo_pack(phase); g.sync()
        for a, p in enumerate(parts):
            pa = p["plan"]
            for q, b in enumerate(pa.neighbours):
                pb = parts[b]["plan"]
                qb = pb.neighbours.index(a)
                so, sc = int(pa.send_offsets[q]) * w, pa.send_counts[q] * w
                ro = int(pb.recv_offsets[qb]) * w
                bufs[b][1][ro:ro + sc].copy_(bufs[a][0][so:so + sc])
        torch.cuda.synchronize()
        for g in gs:
            g.halo_unpack(phase); g.sync()

    def global_dt():
        import ctypes as C
        rs = [g.compute_dt_reductions() for g in gs]
        r = rs[0]
        for o in rs[1:]:
            r.minl = min(r.minl, o.minl); r.dt_maxwell = min(r.dt_maxwell, o.dt_maxwell)
            r.dt_diffusion = min(r.dt_diffusion, o.dt_diffusion); r.global_dt_min = min(r.global_dt_min, o.global_dt_min)
            r.global_max_vem = max(r.global_max_vem, o.global_max_vem); r.max_surf_vel = max(r.max_surf_vel, o.max_surf_vel)
        return gs[0].lib.des_dt_from_reductions(C.byref(gs[0].param), C.byref(r), None)

    dt = global_dt()
    assert dt == dt0
    t = 0.0
    for s in range(1, 13):
        t += dt
        for seg in range(6):
            if overlap and seg in (1, 2):
                # part 0 ran before exchange(seg - 1) below; the ghost-dependent rest follows it
                for g in gs:
                    g.step_segment_part(seg, 1, dt, t)
                    g.step_segment_part(seg, 2, dt, t)
            else:
                for g in gs:
                    g.step_segment(seg, dt, t)
            if overlap and seg in (0, 1):
                for g in gs:
                    g.step_segment_part(seg + 1, 0, dt, t)      # interior elements: ghosts still stale
            if seg < 4:
                exchange(seg)
        if s % 10 == 0:
            dt = global_dt()
        if s in CHECK:
            assert dt == dts[s], (s, dt, dts[s])
            outs = [g.download_state(*FIELDS) for g in gs]
            merged = _owned_merge(parts, outs, N, E)
            bad = [k for k in FIELDS if not np.array_equal(merged[k], refs[s][k])]
            assert not bad, "step %d: fields differ from the single-GPU run: %s" % (s, bad)
            # ghost copies too: a rank-local download (or a host surface-marker pass on a
            # partition) must see the owner's accumulated surface change
            for p, o in zip(parts, outs):
                assert np.array_equal(o["dhacc"], refs[s]["dhacc"][p["node_gid"]]), "ghost dhacc drifts"
    for g in gs:
        g.close()


@pytest.mark.gpu
@pytest.mark.parametrize("world,box", [
    (2, BOX),                                   # 1 200 tets
    (4, (150e3, 100e3, 100e3, 1923.1)),         # 1 054 560 tets: 3-D RCB, edge and corner ghosts
    (8, (150e3, 100e3, 100e3, 1923.1)),         # up to 7 neighbours per rank
])
def test_ranks_nccl_bit_identical(world, box, tmp_path):
    """A real one-process-per-GPU NCCL run (needs `world` GPUs) against the single-GPU run of the
    same box: every node and element field of every owner, and dt after the allreduce, bit for bit."""
    import torch
    if torch.cuda.device_count() < world:
        pytest.skip("needs %d GPUs" % world)
    out = tmp_path / "mg"
    out.mkdir()
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(world),
           "--master-addr", "127.0.0.1", "--master-port", str(_free_port()),
           os.path.join(HERE, "mgpu_worker.py"), str(out)] + [repr(float(x)) for x in box]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    from dynearthsol_b200 import multigpu as mg
    xl, yl, zl, res = box
    nx, ny, nz = bm.grid_dims(xl, yl, zl, res)
    coord, conn, bcflag = bm.regular_box(xl, yl, zl, nx, ny, nz)
    phys = dict(bm.TEST3D_PHYSICS); phys["zlength"] = zl
    state, comp_p = bm.initial_fields(coord, conn, rho0=2700.0, bulk_modulus=50e9, gravity=10.0, xlength=xl,
                                      ylength=yl, zlength=zl, resolution=res, weakzone=bm.TEST3D_WEAKZONE)
    N, E = coord.shape[1], conn.shape[1]
    g1 = mg.init_model(mg.local_mesh(coord, conn, bcflag), phys, state, comp_p, 0)
    g1.call("apply_vbcs", 0.0); g1.call("compute_mass")
    dt1, _, _ = g1.run(12, g1.compute_dt()[0])
    ref = g1.download_state("coord", "vel", "temperature", "force", "mass", "stress", "plstrain", "volume")
    g1.close()
    dx, dy, dz = xl / (nx - 1), yl / (ny - 1), zl / (nz - 1)
    ii = np.rint(coord[0] / dx).astype(np.int64); jj = np.rint(coord[1] / dy).astype(np.int64)
    kk = np.rint(coord[2] / -dz).astype(np.int64)
    gkey = (ii * ny + jj) * nz + kk
    inv = np.argsort(gkey)
    ek = np.sort(gkey[conn], axis=0)
    eorder = np.lexsort(ek[::-1])
    eks = ek[:, eorder]
    seen_n, seen_e = np.zeros(N, dtype=np.int32), np.zeros(E, dtype=np.int32)
    for rnk in range(world):
        d = np.load(out / ("rank%d.npz" % rnk))
        assert float(d["dt"][0]) == dt1
        pos = inv[np.searchsorted(gkey[inv], d["key"])]
        seen_n[pos] += 1
        for k, c in (("coord", 3), ("vel", 3), ("force", 3), ("temperature", 1), ("mass", 1)):
            assert np.array_equal(d[k].reshape(c, -1), ref[k].reshape(c, N)[:, pos]), (rnk, k)
        # elements by their sorted node keys (rows compared lexicographically through a structured view)
        def rows(a):
            return np.ascontiguousarray(a.T).view([("", a.dtype)] * 4).reshape(-1)
        epos = eorder[np.searchsorted(rows(eks), rows(d["ekey"]))]
        assert np.array_equal(ek[:, epos], d["ekey"])
        seen_e[epos] += 1
        assert np.array_equal(d["stress"], ref["stress"].reshape(6, E)[:, epos]), (rnk, "stress")
        for k in ("plstrain", "volume"):
            assert np.array_equal(d[k], ref[k][epos]), (rnk, k)
    assert (seen_n == 1).all() and (seen_e == 1).all()      # every node / element reported by exactly one rank


@pytest.mark.parametrize("world", [2, 8])
def test_interior_first_order_keeps_the_summation_order(world):
    """The device order of a rank's elements (ghost-free elements first, for the overlap of the
    exchanges with E-B / E-C) must not change any per-node sum: support and facet lists stay in
    the global element order."""
    from dynearthsol_b200 import multigpu as mg
    xl, yl, zl, res = BOX
    nx, ny, nz = bm.grid_dims(xl, yl, zl, res)
    boxes = pt.grid_boxes(nx, ny, nz, (xl, yl, zl), world)
    for rank in range(world):
        coord, conn, bcflag, key, owner = bm.regular_box_part(xl, yl, zl, nx, ny, nz, rank, boxes)
        perm, nint = pt.interior_first(conn, owner, rank)
        E = conn.shape[1]
        assert sorted(perm.tolist()) == list(range(E))
        ghost = owner != rank
        dep = ghost[conn].any(axis=0)
        assert not dep[perm[:nint]].any() and dep[perm[nint:]].all()
        assert np.all(np.diff(perm[:nint]) > 0) and np.all(np.diff(perm[nint:]) > 0)      # stable
        plain = mg.local_mesh(coord, conn, bcflag)
        re = mg.local_mesh(coord, conn, bcflag, perm)
        assert np.array_equal(re["connectivity"], conn[:, perm])
        # same support entries, in the same (global) order, renamed
        assert np.array_equal(re["support_idx"], plain["support_idx"])
        assert np.array_equal(perm[re["support_arr"]], plain["support_arr"])
        assert np.array_equal(re["support_lidx"], plain["support_lidx"])
        for b in range(10):
            a, p = re["bfacets%d" % b].reshape(-1, 2), plain["bfacets%d" % b].reshape(-1, 2)
            assert np.array_equal(perm[a[:, 0]], p[:, 0]) and np.array_equal(a[:, 1], p[:, 1])
        assert np.array_equal(re["connectivity_surface"], plain["connectivity_surface"])
        assert np.array_equal(re["elem_and_nodes"], plain["elem_and_nodes"])
        assert np.array_equal(re["support_surf_arr"], plain["support_surf_arr"])


@pytest.mark.parametrize("nparts", [2, 3, 8])
def test_library_partitioner_on_the_host(nparts):
    """The partitioner behind des_gpu_create_multi (csrc/des_multi.cuh), run without a device
    through des_host_partition: RCB parts balanced to one element, every node owned by the lowest
    part around it, and the local element / node / ghost counts equal to what partition.py's
    local_subset (the scheme the NCCL harness uses) builds from the same parts."""
    from dynearthsol_b200 import core
    (nx, ny, nz), coord, conn, bcflag = global_box()
    N, E = coord.shape[1], conn.shape[1]
    part, owner, le, ln, gh = core.host_partition(coord, conn, nparts)
    counts = np.bincount(part, minlength=nparts)
    assert counts.sum() == E and counts.max() - counts.min() <= nparts      # count-proportional bisection
    assert np.array_equal(owner, pt.node_owner_from_parts(conn, part, N))
    # each part is a coordinate box of centroids along the cut axes: no part's bounding boxes interleave badly
    cen = pt.element_centroids(coord, conn)
    for r in range(nparts):
        elem_gid, node_gid, conn_local = pt.local_subset(conn, owner, r)
        assert le[r] == elem_gid.size and ln[r] == node_gid.size
        assert gh[r] == int((owner[node_gid] != r).sum())
        assert np.isfinite(cen[:, part == r]).all()
    # the first cut is along x (the longest side of the 24 x 10 x 8 km box): parts 0..left-1 lie left of the rest
    left = nparts // 2
    assert cen[0, part < left].max() <= cen[0, part >= left].min()
