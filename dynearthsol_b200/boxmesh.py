"""Synthetic structured-box meshes and their topology, generated directly (no host init).

The reference cannot initialise the 50 M / 200 M-tet boxes (its create_boundary_facets is
O(nseg*nelem), mesh.cxx:3177-3216), so the benchmark mesh, its topology arrays and the
initial fields are produced here with the SAME semantics as the reference's regular
generator and builders:

  nodes / cells / 5-tet split   new_mesh_regular: mesh.cxx:147-347 (parity-alternating split)
  renumbering                   renumbering_mesh: mesh.cxx:2696-2821 (sort by x + 1e-3 y + 1e-6 z)
  bcflag / bnodes / bfacets     mesh.cxx:2824-2880, 3161-3246
  support CSR                   create_support: mesh.cxx:3287-3329 (ascending element ids)
  top_elems / surface info      mesh.cxx:2882-2931, 3031-3101
  boundary normals, edges       create_boundary_normals: bc.cxx:94-224
  lithostatic stress, weak zone ic.cxx:322-362, 29-69, 497-654

Integer outputs are compared bit-for-bit with the reference's own on small boxes
(tests/test_boxmesh.py).  Everything is numpy on the host; it is mesh *input* preparation,
not part of the timed path.
"""
import numpy as np

NODE_OF_FACET = np.array([[1, 2, 3], [0, 3, 2], [0, 1, 3], [0, 2, 1]], dtype=np.int32)

# divide_hexahedron_to_tetrahedra_index, mesh.cxx:190-262: local cell vertices of the 5 tets
TETS_EVEN = np.array([[0, 1, 2, 5], [0, 2, 3, 7], [0, 4, 5, 7], [2, 5, 6, 7], [0, 5, 2, 7]], dtype=np.int32)
TETS_ODD = np.array([[1, 2, 3, 6], [0, 1, 3, 4], [1, 4, 5, 6], [3, 4, 6, 7], [1, 3, 4, 6]], dtype=np.int32)


def grid_dims(xlength, ylength, zlength, resolution):
    """init_var, dynearthsol.cxx:127-140"""
    nx = int(round(xlength / resolution)) + 1
    ny = int(round(ylength / resolution)) + 1
    nz = int(round(zlength / resolution)) + 1
    return nx, ny, nz


def regular_box(xlength, ylength, zlength, nx, ny, nz, renumber=True, order="reference"):
    """Returns coord [3,N] float64, connectivity [4,E] int32 (both SoA like Array2D), bcflag [N].

    order="reference": the reference's numbering (renumbering_mesh when renumber=True).
    order="morton"   : nodes and cells reordered along a Morton curve over (i, j, k) blocks --
                       same mesh, better gather locality on the GPU.
    """
    N = nx * ny * nz
    dx = xlength / (nx - 1)
    dy = ylength / (ny - 1)
    dz = -zlength / (nz - 1)
    # create_rect_node, mesh.cxx:320-347: idx = k + j*nz + i*ny*nz
    ii, jj, kk = np.meshgrid(np.arange(nx, dtype=np.int32), np.arange(ny, dtype=np.int32),
                             np.arange(nz, dtype=np.int32), indexing="ij")
    coord = np.empty((3, N), dtype=np.float64)
    coord[0] = (ii * dx).reshape(-1)
    coord[1] = (jj * dy).reshape(-1)
    coord[2] = (kk * dz).reshape(-1)
    bcflag = np.zeros(N, dtype=np.uint32)
    bcflag[(ii == 0).reshape(-1)] |= 1
    bcflag[(ii == nx - 1).reshape(-1)] |= 2
    bcflag[(jj == 0).reshape(-1)] |= 4
    bcflag[(jj == ny - 1).reshape(-1)] |= 8
    bcflag[(kk == nz - 1).reshape(-1)] |= 16      # bottom: z = -zlength
    bcflag[(kk == 0).reshape(-1)] |= 32           # top: z = 0
    del ii, jj, kk

    # create_quadrilateral_cells + create_elem_from_cell, mesh.cxx:147-317
    ci, cj, ck = np.meshgrid(np.arange(nx - 1, dtype=np.int32), np.arange(ny - 1, dtype=np.int32),
                             np.arange(nz - 1, dtype=np.int32), indexing="ij")
    ci = ci.reshape(-1); cj = cj.reshape(-1); ck = ck.reshape(-1)
    ncell = ci.size
    idx0 = ci * (ny * nz) + cj * nz + ck
    verts = np.empty((8, ncell), dtype=np.int32)
    verts[0] = idx0
    verts[1] = idx0 + nz
    verts[2] = idx0 + nz + ny * nz
    verts[3] = idx0 + ny * nz
    verts[4:8] = verts[0:4] + 1
    odd = ((ci + cj + ck) % 2).astype(bool)
    cell_key = None
    if order == "morton":
        cell_key = _morton3(ci, cj, ck)
    del ci, cj, ck, idx0
    conn = np.empty((4, ncell, 5), dtype=np.int32)
    cells = np.arange(ncell)
    for n in range(5):
        for v in range(4):
            conn[v, :, n] = np.where(odd, verts[TETS_ODD[n, v], cells], verts[TETS_EVEN[n, v], cells])
    del verts, odd, cells
    if order == "morton":
        perm = np.argsort(cell_key, kind="stable")
        conn = conn[:, perm, :]
        del perm, cell_key
    conn = conn.reshape(4, ncell * 5)

    if order == "morton":
        # nodes in first-touch order of the Morton-ordered elements
        first = np.full(N, np.iinfo(np.int64).max, dtype=np.int64)
        flat = conn.T.reshape(-1)
        np.minimum.at(first, flat, np.arange(flat.size, dtype=np.int64))
        nd_idx = np.argsort(first, kind="stable").astype(np.int32)
        del first, flat
        nd_inv = np.empty(N, dtype=np.int32)
        nd_inv[nd_idx] = np.arange(N, dtype=np.int32)
        coord = coord[:, nd_idx]
        bcflag = bcflag[nd_idx]
        conn = nd_inv[conn]
    elif renumber:
        # renumbering_mesh with meshing_elem_shape=1: dmax=x, dmid=y, dmin=z; f=1e-3
        f = 1e-3
        wn = coord[0] + f * coord[1] + f * f * coord[2]
        we = wn[conn[0]] + wn[conn[1]] + wn[conn[2]] + wn[conn[3]]
        nd_idx = np.argsort(wn, kind="stable").astype(np.int32)
        el_idx = np.argsort(we, kind="stable")
        nd_inv = np.empty(N, dtype=np.int32)
        nd_inv[nd_idx] = np.arange(N, dtype=np.int32)
        coord = coord[:, nd_idx]
        bcflag = bcflag[nd_idx]
        conn = nd_inv[conn[:, el_idx]]
    return np.ascontiguousarray(coord), np.ascontiguousarray(conn.astype(np.int32)), bcflag


def regular_box_part(xlength, ylength, zlength, nx, ny, nz, rank, boxes):
    """The part of the regular box that rank `rank` holds when the node index space is split
    into `boxes` (partition.grid_boxes): the elements touching a node the rank owns, and their
    nodes, generated directly -- the global mesh is never built.  Node coordinates, the
    parity-alternating 5-tet split and the relative order of nodes and elements are those of
    regular_box(), so the local mesh is the restriction of the global one.

    Returns coord [3,N], conn [4,E] (local indices), bcflag [N], node_key [N] (global
    lexicographic node id, int64) and owner [N] (owning rank of every local node)."""
    from . import partition
    (ia, ib), (ja, jb), (ka, kb) = boxes[rank]
    i0, i1 = max(ia - 1, 0), min(ib, nx - 1)        # local node index ranges, inclusive
    j0, j1 = max(ja - 1, 0), min(jb, ny - 1)
    k0, k1 = max(ka - 1, 0), min(kb, nz - 1)
    nxl, nyl, nzl = i1 - i0 + 1, j1 - j0 + 1, k1 - k0 + 1
    dx = xlength / (nx - 1)
    dy = ylength / (ny - 1)
    dz = -zlength / (nz - 1)
    ii, jj, kk = np.meshgrid(np.arange(i0, i1 + 1, dtype=np.int32), np.arange(j0, j1 + 1, dtype=np.int32),
                             np.arange(k0, k1 + 1, dtype=np.int32), indexing="ij")
    ii = ii.reshape(-1); jj = jj.reshape(-1); kk = kk.reshape(-1)
    coord = np.empty((3, ii.size), dtype=np.float64)
    coord[0] = ii * dx
    coord[1] = jj * dy
    coord[2] = kk * dz
    bcflag = np.zeros(ii.size, dtype=np.uint32)
    bcflag[ii == 0] |= 1
    bcflag[ii == nx - 1] |= 2
    bcflag[jj == 0] |= 4
    bcflag[jj == ny - 1] |= 8
    bcflag[kk == nz - 1] |= 16
    bcflag[kk == 0] |= 32
    node_key = (ii.astype(np.int64) * ny + jj) * nz + kk
    owner = partition.owner_of_grid_nodes(ii, jj, kk, boxes)
    del ii, jj, kk

    ci, cj, ck = np.meshgrid(np.arange(i0, i1, dtype=np.int32), np.arange(j0, j1, dtype=np.int32),
                             np.arange(k0, k1, dtype=np.int32), indexing="ij")
    ci = ci.reshape(-1); cj = cj.reshape(-1); ck = ck.reshape(-1)
    ncell = ci.size
    idx0 = (ci - i0) * (nyl * nzl) + (cj - j0) * nzl + (ck - k0)
    verts = np.empty((8, ncell), dtype=np.int32)
    verts[0] = idx0
    verts[1] = idx0 + nzl
    verts[2] = idx0 + nzl + nyl * nzl
    verts[3] = idx0 + nyl * nzl
    verts[4:8] = verts[0:4] + 1
    odd = ((ci + cj + ck) % 2).astype(bool)
    del ci, cj, ck, idx0
    conn = np.empty((4, ncell, 5), dtype=np.int32)
    cells = np.arange(ncell)
    for n in range(5):
        for v in range(4):
            conn[v, :, n] = np.where(odd, verts[TETS_ODD[n, v], cells], verts[TETS_EVEN[n, v], cells])
    del verts, odd, cells
    conn = conn.reshape(4, ncell * 5)
    # keep the elements that touch an owned node, and the nodes they use
    mine = owner == rank
    keep = mine[conn[0]] | mine[conn[1]] | mine[conn[2]] | mine[conn[3]]
    conn = conn[:, keep]
    used = np.zeros(coord.shape[1], dtype=bool)
    used[conn.reshape(-1)] = True
    nd = np.nonzero(used)[0]
    remap = np.full(coord.shape[1], -1, dtype=np.int32)
    remap[nd] = np.arange(nd.size, dtype=np.int32)
    conn = remap[conn]
    coord = coord[:, nd]; bcflag = bcflag[nd]; node_key = node_key[nd]; owner = owner[nd]
    # renumbering_mesh: the same keys and stable sorts as regular_box()
    f = 1e-3
    wn = coord[0] + f * coord[1] + f * f * coord[2]
    we = wn[conn[0]] + wn[conn[1]] + wn[conn[2]] + wn[conn[3]]
    nd_idx = np.argsort(wn, kind="stable").astype(np.int32)
    el_idx = np.argsort(we, kind="stable")
    nd_inv = np.empty(nd_idx.size, dtype=np.int32)
    nd_inv[nd_idx] = np.arange(nd_idx.size, dtype=np.int32)
    coord = coord[:, nd_idx]
    conn = nd_inv[conn[:, el_idx]]
    return (np.ascontiguousarray(coord), np.ascontiguousarray(conn.astype(np.int32)), bcflag[nd_idx],
            node_key[nd_idx], owner[nd_idx])


def _part1by2(x):
    x = x.astype(np.int64) & 0x1fffff
    x = (x | (x << 32)) & 0x1f00000000ffff
    x = (x | (x << 16)) & 0x1f0000ff0000ff
    x = (x | (x << 8)) & 0x100f00f00f00f00f
    x = (x | (x << 4)) & 0x10c30c30c30c30c3
    x = (x | (x << 2)) & 0x1249249249249249
    return x


def _morton3(i, j, k):
    return _part1by2(i) | (_part1by2(j) << 1) | (_part1by2(k) << 2)


def build_support(nnode, conn):
    """create_support, mesh.cxx:3287-3329.  Returns idx [N+1], arr [4E], lidx [4E] (int32)."""
    E = conn.shape[1]
    keys = conn.T.reshape(-1)                      # (e, i) order == the reference's fill order
    order = np.argsort(keys, kind="stable")
    arr = (order // 4).astype(np.int32)
    lidx = (order % 4).astype(np.int32)
    counts = np.bincount(keys, minlength=nnode)
    idx = np.zeros(nnode + 1, dtype=np.int64)
    np.cumsum(counts, out=idx[1:])
    assert idx[-1] == 4 * E
    return idx.astype(np.int32), arr, lidx


def boundary_topology(coord, conn, bcflag, support=None, elem_key=None):
    """bnodes, bfacets, connectivity_surface, top_elems, surface info, normals and edges.
    Returns a dict with the DesMesh keys (dynearthsol_b200.core.DesGpu.upload_mesh).
    elem_key [E]: sort key of the elements for the facet lists (default: the element id).  A
    rank that stores its elements in another order than the global one (multigpu: interior
    elements first) passes the global order here, so that every per-node sum over facets keeps
    the single-GPU order."""
    N = coord.shape[1]
    E = conn.shape[1]
    m = {"nnode": N, "nelem": E, "connectivity": conn, "bcflag": bcflag}
    if support is None:
        support = build_support(N, conn)
    m["support_idx"], m["support_arr"], m["support_lidx"] = support
    for b in range(10):
        bit = np.uint32(1 << b)
        m["bnodes%d" % b] = np.nonzero(bcflag & bit)[0].astype(np.int32)
        onb = (bcflag & bit) != 0
        es, fs = [], []
        if m["bnodes%d" % b].size:
            for f in range(4):
                hit = onb[conn[NODE_OF_FACET[f, 0]]] & onb[conn[NODE_OF_FACET[f, 1]]] & onb[conn[NODE_OF_FACET[f, 2]]]
                e = np.nonzero(hit)[0].astype(np.int32)
                es.append(e)
                fs.append(np.full(e.size, f, dtype=np.int32))
        if es:
            e = np.concatenate(es); f = np.concatenate(fs)
            o = np.lexsort((f, e if elem_key is None else elem_key[e]))
            pairs = np.stack([e[o], f[o]], axis=1).reshape(-1)
        else:
            pairs = np.zeros(0, dtype=np.int32)
        m["bfacets%d" % b] = np.ascontiguousarray(pairs, dtype=np.int32)

    # connectivity_surface, mesh.cxx:3249-3262 (conn_t: 4 columns SoA, 3 used)
    top = m["bfacets5"].reshape(-1, 2)
    etop = top.shape[0]
    cs = np.zeros((4, etop), dtype=np.int32)
    for j in range(3):
        cs[j] = conn[NODE_OF_FACET[top[:, 1], j], top[:, 0]]
    m["connectivity_surface"] = cs
    # create_surface_info, mesh.cxx:3031-3101: top nodes sorted by x
    tn = m["bnodes5"]
    tn = tn[np.argsort(coord[0, tn], kind="stable")]
    m["top_nodes"] = tn.astype(np.int32)
    arctop = np.full(N, -1, dtype=np.int32)
    arctop[tn] = np.arange(tn.size, dtype=np.int32)
    ean = arctop[cs[:3]]
    m["elem_and_nodes"] = np.ascontiguousarray(ean, dtype=np.int32)
    # create_support_surf, mesh.cxx:2934-2964
    keys = ean.T.reshape(-1)
    order = np.argsort(keys, kind="stable")
    m["support_surf_arr"] = (order // 3).astype(np.int32)
    sidx = np.zeros(tn.size + 1, dtype=np.int64)
    np.cumsum(np.bincount(keys, minlength=tn.size), out=sidx[1:])
    m["support_surf_idx"] = sidx.astype(np.int32)
    # create_top_elems, mesh.cxx:2882-2931: ascending unique support elements of top nodes
    idx, arr = m["support_idx"], m["support_arr"]
    istop = np.zeros(E, dtype=bool)
    onz1 = (bcflag & 32) != 0
    for v in range(4):
        istop |= onz1[conn[v]]
    m["top_elems"] = np.nonzero(istop)[0].astype(np.int32)

    # create_boundary_normals, bc.cxx:94-224 (box boundaries: first facet only)
    bn = np.zeros((3, 10), dtype=np.float64)
    for b in range(10):
        pf = m["bfacets%d" % b]
        if pf.size == 0:
            continue
        e, f = int(pf[0]), int(pf[1])
        p = [coord[:, conn[NODE_OF_FACET[f, k], e]] for k in range(3)]
        v01 = p[1] - p[0]
        v02 = p[2] - p[0]
        nrm = np.array([(v01[1] * v02[2] - v01[2] * v02[1]) / 2,
                        (v01[2] * v02[0] - v01[0] * v02[2]) / 2,
                        (v01[0] * v02[1] - v01[1] * v02[0]) / 2])
        ln = np.sqrt(nrm[0] * nrm[0] + nrm[1] * nrm[1] + nrm[2] * nrm[2])
        bn[:, b] = nrm / ln
    m["bnormals"] = bn
    slot = np.full(100, -1, dtype=np.int32)
    ev = []
    eps = 1e-15
    for i in range(10):
        if m["bfacets%d" % i].size == 0:
            continue
        for j in range(i + 1, 10):
            if m["bfacets%d" % j].size == 0:
                continue
            if abs(bn[2, i]) < eps and abs(bn[2, j]) < eps:
                s = [0.0, 0.0, 1.0]
            else:
                s = [bn[1, i] * bn[2, j] - bn[2, i] * bn[1, j],
                     bn[2, i] * bn[0, j] - bn[0, i] * bn[2, j],
                     bn[0, i] * bn[1, j] - bn[1, i] * bn[0, j]]
            slot[i * 10 + j] = len(ev) // 3
            ev += s
    m["edge_slot"] = slot
    m["edge_vec"] = np.asarray(ev, dtype=np.float64)
    return m


def tet_volumes(coord, conn):
    """tetrahedron_volume, geometry.cxx:36-56 (same operation order)."""
    d0, d1, d2, d3 = (coord[:, conn[i]] for i in range(4))
    x01 = d0[0] - d1[0]; x12 = d1[0] - d2[0]; x23 = d2[0] - d3[0]
    y01 = d0[1] - d1[1]; y12 = d1[1] - d2[1]; y23 = d2[1] - d3[1]
    z01 = d0[2] - d1[2]; z12 = d1[2] - d2[2]; z23 = d2[2] - d3[2]
    return (x01 * (y23 * z12 - y12 * z23) + x12 * (y01 * z23 - y23 * z01) + x23 * (y12 * z01 - y01 * z12)) / 6


def initial_fields(coord, conn, *, rho0, bulk_modulus, gravity, xlength, ylength, zlength, resolution,
                   weakzone=None, temperature=273.0):
    """Lithostatic stress/strain (ic.cxx:322-362, ref_pressure_option 0), planar weak zone
    (ic.cxx:29-69, 497-528; weakzone = dict(azimuth, inclination, halfwidth, depth_min, depth_max,
    xcenter, ycenter, zcenter, plstrain, y_min=0, y_max=1)), uniform temperature.
    Returns a dict of DesState arrays (SoA) and the compensation pressure."""
    N = coord.shape[1]
    E = conn.shape[1]
    zc = np.zeros(E)
    for i in range(4):
        zc += coord[2, conn[i]]
    zc /= 4
    p = rho0 * gravity * (-zc)
    stress = np.zeros((6, E))
    strain = np.zeros((6, E))
    for i in range(3):
        stress[i] = -p
        strain[i] = -p / bulk_modulus / 3
    plstrain = np.zeros(E)
    if weakzone is not None:
        w = dict(y_min=0.0, y_max=1.0)
        w.update(weakzone)
        deg = np.pi / 180
        az = np.tan(w["azimuth"] * deg)
        incl = 1 / np.tan(w["inclination"] * deg)
        hw = w["halfwidth"] * resolution
        x0 = (w["xcenter"] * xlength, w["ycenter"] * ylength, -w["zcenter"] * zlength)
        ymin, ymax = w["y_min"] * ylength, w["y_max"] * ylength
        zmin, zmax = -w["depth_max"] * zlength, -w["depth_min"] * zlength
        c = np.zeros((3, E))
        for i in range(4):
            c += coord[:, conn[i]]
        c /= 4
        inside = ((c[2] > zmin) & (c[2] < zmax) & (c[1] > ymin) & (c[1] < ymax) &
                  (np.abs((c[0] - x0[0]) - az * (c[1] - x0[1]) + incl * (c[2] - x0[2])) < hw))
        plstrain[inside] = w["plstrain"] * 1.0
    vol = tet_volumes(coord, conn)
    st = dict(coord=coord, coord0=coord.copy(), vel=np.zeros((3, N)),
              temperature=np.full(N, float(temperature)),
              stress=stress, strain=strain, strain_rate=np.zeros((6, E)), plstrain=plstrain,
              delta_plstrain=np.zeros(E), volume=vol, volume_old=vol.copy(),
              radiogenic_source=np.zeros(E))
    return st, rho0 * gravity * zlength


# benchmarks-cores/test-3d.cfg physics ([control]/[bc]/[ic]/[mat] blocks), restated as keywords
TEST3D_PHYSICS = dict(
    gravity=10.0, inertial_scaling=1e4, damping_factor=0.8, damping_option=1, is_quasi_static=1,
    has_thermal_diffusion=1, is_using_mixed_stress=1, surface_process_option=1,
    surface_diffusivity=1e-6, dt_fraction=1.0, has_moving_mesh=1,
    has_winkler_foundation=1, has_water_loading=0, surface_temperature=273.0,
    vbc_types=[1, 1, 1, 1, 0, 0, 1, 1, 1, 1],
    vbc_values=[-1e-9, 1e-9, 0, 0, 0, 0, 0, 0, 0, 0],
    max_vbc_val=1e-9, rheol_type=5, nmat=1, visc_min=1e24, visc_max=1e24,
    tension_max=1e9, therm_diff_max=5e-6,
    rho0=[2700.0], alpha=[0.0], bulk_modulus=[50e9], shear_modulus=[30e9],
    visc_exponent=[3.05], visc_coefficient=[1.25e-1], visc_activation_energy=[3.76e5],
    visc_activation_volume=[0.0], heat_capacity=[1000.0], therm_cond=[3.0],
    pls0=[0.0], pls1=[0.5], cohesion0=[4.4e7], cohesion1=[4e6],
    friction_angle0=[30.0], friction_angle1=[30.0], dilation_angle0=[0.0], dilation_angle1=[0.0],
    porosity=[0.0])

TEST3D_WEAKZONE = dict(azimuth=15.0, inclination=-60.0, halfwidth=1.2, depth_min=0.5, depth_max=1.0,
                       xcenter=0.5, ycenter=0.5, zcenter=0.0, plstrain=0.5)
