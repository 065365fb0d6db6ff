"""Element partitioning and ghost-node exchange plans for the multi-GPU explicit step.

The reference has no distributed mode; this is the host-side integer
work that shards its hot path.  Scheme:

  * elements are split by recursive coordinate bisection (RCB) of their centroids;
  * a node is OWNED by the lowest part among the elements around it;
  * a rank keeps every element that touches a node it owns (its RCB part plus a one-element
    layer, computed redundantly) and the nodes of those elements; the ones it does not own
    are GHOSTS;
  * local elements and nodes keep their global relative order, so an owned node sums its
    support in the reference's ascending element order (mesh.cxx:3287-3329) and the result
    is bit-identical to the single-GPU (and the reference's) result;
  * after each nodal barrier of a step the owner sends the new values of the nodes that are
    ghosts elsewhere -- a pure copy, no reduction, hence deterministic.

Everything here is numpy on the host (integer and index work); the exchange itself runs
over torch.distributed in dynearthsol_b200/multigpu.py.
"""
import numpy as np


def rcb(points, nparts):
    """Recursive coordinate bisection of `points` [3, n] into `nparts` parts (any nparts >= 1):
    split the longest extent at the count-proportional quantile, recurse.  Returns part [n]."""
    n = points.shape[1]
    part = np.zeros(n, dtype=np.int32)

    def split(ids, p0, np_):
        if np_ == 1:
            part[ids] = p0
            return
        pts = points[:, ids]
        ext = pts.max(axis=1) - pts.min(axis=1)
        ax = int(np.argmax(ext))
        left = np_ // 2
        k = int(round(ids.size * left / np_))
        order = np.argsort(pts[ax], kind="stable")
        split(ids[order[:k]], p0, left)
        split(ids[order[k:]], p0 + left, np_ - left)

    split(np.arange(n), 0, int(nparts))
    return part


def element_centroids(coord, conn):
    c = np.zeros((3, conn.shape[1]))
    for i in range(4):
        c += coord[:, conn[i]]
    return c / 4


def node_owner_from_parts(conn, part, nnode):
    """owner[n] = lowest part among the elements that contain node n"""
    owner = np.full(nnode, np.iinfo(np.int32).max, dtype=np.int32)
    for i in range(4):
        np.minimum.at(owner, conn[i], part)
    return owner


def local_subset(conn, owner, rank):
    """Elements touching a node owned by `rank` and the nodes of those elements, both in
    ascending global order.  Returns (elem_gid, node_gid, conn_local)."""
    mine = owner == rank
    touch = mine[conn[0]] | mine[conn[1]] | mine[conn[2]] | mine[conn[3]]
    elem_gid = np.nonzero(touch)[0]
    sub = conn[:, elem_gid]
    node_gid = np.unique(sub)
    conn_local = np.searchsorted(node_gid, sub).astype(np.int32)
    return elem_gid, node_gid, conn_local


def interior_first(conn_local, owner_local, rank):
    """Stable reordering of a rank's elements: those that touch no ghost node first.  Returns
    (perm, n_interior) with new element i = old element perm[i].  The ghost exchange that
    precedes an element kernel can then overlap with the kernel's run over [0, n_interior)."""
    ghost = owner_local != rank
    dep = ghost[conn_local[0]] | ghost[conn_local[1]] | ghost[conn_local[2]] | ghost[conn_local[3]]
    perm = np.concatenate([np.nonzero(~dep)[0], np.nonzero(dep)[0]]).astype(np.int64)
    return perm, int((~dep).sum())


class HaloPlan:
    """Who sends what to whom.  For neighbour ranks in ascending order:
    send_nodes / recv_nodes are local node indices concatenated over neighbours, each
    neighbour's block sorted by the nodes' global key, so that the sender's block for rank B
    lists the same nodes in the same order as B's receive block from the sender."""

    def __init__(self, rank, neighbours, send_nodes, send_counts, recv_nodes, recv_counts):
        self.rank = rank
        self.neighbours = [int(r) for r in neighbours]
        self.send_nodes = np.ascontiguousarray(send_nodes, dtype=np.int32)
        self.recv_nodes = np.ascontiguousarray(recv_nodes, dtype=np.int32)
        self.send_counts = [int(c) for c in send_counts]
        self.recv_counts = [int(c) for c in recv_counts]
        self.send_offsets = np.concatenate([[0], np.cumsum(self.send_counts)]).astype(np.int64)
        self.recv_offsets = np.concatenate([[0], np.cumsum(self.recv_counts)]).astype(np.int64)


def halo_plan(rank, conn_local, owner_local, node_key):
    """Exchange plan from purely local information.

    conn_local [4, E]   local connectivity of the rank's elements
    owner_local [N]     owning rank of every local node
    node_key [N]        a globally agreed integer key of every local node (global node id)

    Rank B needs my node n iff n lies in an element that touches a B-owned node; every element
    around n is local to me (I own n), so the test needs no communication."""
    own = owner_local[conn_local]                       # [4, E]
    dests, nodes = [], []
    for i in range(4):
        for j in range(4):
            if i == j:
                continue
            m = (own[i] == rank) & (own[j] != rank)
            if m.any():
                dests.append(own[j][m])
                nodes.append(conn_local[i][m])
    if dests:
        pairs = np.unique(np.stack([np.concatenate(dests).astype(np.int64),
                                    np.concatenate(nodes).astype(np.int64)], axis=1), axis=0)
    else:
        pairs = np.zeros((0, 2), dtype=np.int64)
    ghosts = np.nonzero(owner_local != rank)[0]
    neigh = sorted(set(pairs[:, 0].tolist()) | set(owner_local[ghosts].tolist()))
    send_nodes, send_counts, recv_nodes, recv_counts = [], [], [], []
    for r in neigh:
        s = pairs[pairs[:, 0] == r, 1]
        s = s[np.argsort(node_key[s], kind="stable")]
        send_nodes.append(s); send_counts.append(s.size)
        g = ghosts[owner_local[ghosts] == r]
        g = g[np.argsort(node_key[g], kind="stable")]
        recv_nodes.append(g); recv_counts.append(g.size)
    cat = lambda xs: np.concatenate(xs) if xs else np.zeros(0, dtype=np.int64)
    return HaloPlan(rank, neigh, cat(send_nodes), send_counts, cat(recv_nodes), recv_counts)


def partition_mesh(coord, conn, bcflag, nparts):
    """RCB partition of a global mesh.  Returns a list (one per rank) of dicts with
    elem_gid, node_gid, conn, coord, bcflag, owner (per local node), owned (mask), plan."""
    nnode = coord.shape[1]
    part = rcb(element_centroids(coord, conn), nparts)
    owner = node_owner_from_parts(conn, part, nnode)
    out = []
    for rank in range(nparts):
        elem_gid, node_gid, conn_local = local_subset(conn, owner, rank)
        owner_local = owner[node_gid]
        plan = halo_plan(rank, conn_local, owner_local, node_gid)
        out.append(dict(rank=rank, elem_gid=elem_gid, node_gid=node_gid, conn=conn_local,
                        coord=np.ascontiguousarray(coord[:, node_gid]), bcflag=bcflag[node_gid],
                        owner=owner_local, owned=owner_local == rank, plan=plan))
    return out


def restrict_state(state, elem_gid, node_gid, nnode, nelem):
    """Local copy of a dict of global SoA state arrays (DesState names)."""
    out = {}
    for k, v in state.items():
        a = np.asarray(v)
        flat = a.reshape(-1)
        if flat.size == nnode:
            out[k] = flat[node_gid]
        elif flat.size == 3 * nnode:
            out[k] = flat.reshape(3, nnode)[:, node_gid]
        elif flat.size == nelem:
            out[k] = flat[elem_gid]
        elif flat.size == 6 * nelem:
            out[k] = flat.reshape(6, nelem)[:, elem_gid]
        else:
            raise ValueError("cannot restrict %s (size %d)" % (k, flat.size))
        out[k] = np.ascontiguousarray(out[k])
    return out


# ---- structured boxes: the node-index space is bisected directly, no global mesh needed ----

def grid_boxes(nx, ny, nz, lengths, nparts):
    """RCB of the node index space [0,nx) x [0,ny) x [0,nz) into `nparts` boxes of owned nodes:
    always cut the physically longest side at the count-proportional index.
    Returns a list of ((ia, ib), (ja, jb), (ka, kb)), half-open."""
    spacing = [lengths[0] / (nx - 1), lengths[1] / (ny - 1), lengths[2] / (nz - 1)]
    boxes = []

    def split(box, np_):
        if np_ == 1:
            boxes.append(tuple(box))
            return
        ext = [(box[d][1] - box[d][0]) * spacing[d] for d in range(3)]
        ax = int(np.argmax(ext))
        left = np_ // 2
        a, b = box[ax]
        cut = a + int(round((b - a) * left / np_))
        lo = list(box); hi = list(box)
        lo[ax] = (a, cut); hi[ax] = (cut, b)
        split(lo, left)
        split(hi, np_ - left)

    split([(0, nx), (0, ny), (0, nz)], int(nparts))
    return boxes


def owner_of_grid_nodes(i, j, k, boxes):
    owner = np.full(i.shape, -1, dtype=np.int32)
    for r, ((ia, ib), (ja, jb), (ka, kb)) in enumerate(boxes):
        m = (i >= ia) & (i < ib) & (j >= ja) & (j < jb) & (k >= ka) & (k < kb)
        owner[m] = r
    return owner
