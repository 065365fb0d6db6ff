"""One process per GPU: the explicit step on an element-partitioned mesh, ghost-node values
exchanged over torch.distributed (NCCL on the GPUs, gloo in the CPU tests) between the nodal
barriers of a step.  See partition.py for the scheme and DESIGN.md section 7.

torch is plumbing here: device buffers of the C-ABI library are wrapped as tensors (no copy)
so that torch.distributed can send them; the numerical path is the CUDA library alone.
"""
import json
import os
import sys
import time

import numpy as np

from . import boxmesh as bm, core, partition


class _DevPtr:
    """Expose a raw device pointer of the library as a CUDA array (no copy)."""

    def __init__(self, ptr, n):
        self.__cuda_array_interface__ = {"shape": (int(n),), "typestr": "<f8", "data": (int(ptr), False),
                                         "version": 2}


def device_tensor(ptr, n, device):
    import torch
    if n == 0 or not ptr:
        return torch.empty(0, dtype=torch.float64, device=device)
    return torch.as_tensor(_DevPtr(ptr, n), device=device)


class HaloExchanger:
    """Point-to-point exchange of the packed ghost buffers: one message per neighbour and
    direction, `width` doubles per node.  Works on any torch.distributed backend."""

    def __init__(self, plan, send_buf, recv_buf, group=None):
        self.plan, self.send_buf, self.recv_buf, self.group = plan, send_buf, recv_buf, group

    def exchange(self, width):
        import torch.distributed as dist
        p = self.plan
        ops = []
        for q, r in enumerate(p.neighbours):
            so, sc = int(p.send_offsets[q]) * width, p.send_counts[q] * width
            ro, rc = int(p.recv_offsets[q]) * width, p.recv_counts[q] * width
            if rc:
                ops.append(dist.P2POp(dist.irecv, self.recv_buf[ro:ro + rc], r, self.group))
            if sc:
                ops.append(dist.P2POp(dist.isend, self.send_buf[so:so + sc], r, self.group))
        if ops:
            for w in dist.batch_isend_irecv(ops):
                w.wait()


HALO_WIDTH = (2, 1, 6, 2)   # des_gpu_halo_width(phase)


class DistModel:
    """The reference's step loop on one rank of an element-partitioned mesh."""

    def __init__(self, g, plan, device, group=None, n_interior=None, overlap=None):
        import torch
        self.g, self.plan, self.group = g, plan, group
        self.device = torch.device("cuda", device)
        # kernels and collectives share one stream (made torch's current stream, which the
        # NCCL process group orders its sends/receives against), so they are ordered
        self.stream = torch.cuda.Stream(self.device)
        torch.cuda.set_stream(self.stream)
        g.set_stream(self.stream.cuda_stream)
        g.set_halo(plan.send_nodes, plan.recv_nodes)
        sp, rp = g.halo_buffers()
        self.send_buf = device_tensor(sp, plan.send_nodes.size * 7, self.device)
        self.recv_buf = device_tensor(rp, plan.recv_nodes.size * 7, self.device)
        self.x = HaloExchanger(plan, self.send_buf, self.recv_buf, group)
        self.exchanges = 0
        # overlap of the exchanges that precede E-B and E-C with those kernels' run over the
        # elements that touch no ghost node (the first n_interior, partition.interior_first):
        # they go to a side stream while the messages are in flight (DES_OVERLAP=0 turns it off)
        if overlap is None:
            overlap = os.environ.get("DES_OVERLAP", "1") != "0"
        self.overlap = bool(overlap) and n_interior is not None and 0 < n_interior
        if self.overlap:
            g.set_interior(n_interior)
            self.side = torch.cuda.Stream(self.device)
            self.ev_main = torch.cuda.Event()
            self.ev_side = torch.cuda.Event()

    def exchange(self, phase):
        self.g.halo_pack(phase)
        self.x.exchange(HALO_WIDTH[phase])
        self.g.halo_unpack(phase)
        self.exchanges += 1

    def exchange_then_segment(self, phase, seg, dt, time):
        """exchange(phase) followed by segment seg, with the segment's element kernel on the
        interior elements running while the messages are in flight."""
        g = self.g
        if not self.overlap:
            self.exchange(phase)
            g.step_segment(seg, dt, time)
            return
        g.halo_pack(phase)
        self.ev_main.record(self.stream)
        self.side.wait_event(self.ev_main)              # the previous node kernel has finished
        g.step_segment_part(seg, 0, dt, time, self.side.cuda_stream)
        self.ev_side.record(self.side)
        self.x.exchange(HALO_WIDTH[phase])
        g.halo_unpack(phase)
        g.step_segment_part(seg, 1, dt, time)           # the elements that read ghost nodes
        self.stream.wait_event(self.ev_side)
        g.step_segment_part(seg, 2, dt, time)           # the node gather needs both
        self.exchanges += 1

    def step(self, dt, time=0.0, steps=0):
        g = self.g
        mixed = bool(g.param.is_using_mixed_stress)
        g.step_segment(0, dt, time)
        self.exchange_then_segment(0, 1, dt, time)
        if mixed:
            self.exchange_then_segment(1, 2, dt, time)
        else:
            g.step_segment(2, dt, time)
        self.exchange(2)
        g.step_segment(3, dt, time)
        if g.param.surface_process_option == 1 and g.param.has_moving_mesh:
            self.exchange(3)
        g.step_segment(4, dt, time)
        g.step_segment(5, dt, time)

    def compute_dt(self):
        """Local reductions, min/max over ranks, then the reference's scalar algebra
        (geometry.cxx:1597-1630)."""
        import ctypes as C
        import torch
        import torch.distributed as dist
        r = self.g.compute_dt_reductions()
        mn = torch.tensor([r.minl, r.dt_maxwell, r.dt_diffusion, r.global_dt_min], dtype=torch.float64,
                          device=self.device)
        mx = torch.tensor([r.global_max_vem, r.max_surf_vel], dtype=torch.float64, device=self.device)
        dist.all_reduce(mn, op=dist.ReduceOp.MIN, group=self.group)
        dist.all_reduce(mx, op=dist.ReduceOp.MAX, group=self.group)
        mn, mx = mn.cpu().numpy(), mx.cpu().numpy()
        r.minl, r.dt_maxwell, r.dt_diffusion, r.global_dt_min = [float(v) for v in mn]
        r.global_max_vem, r.max_surf_vel = float(mx[0]), float(mx[1])
        vmag = C.c_double()
        dt = self.g.lib.des_dt_from_reductions(C.byref(self.g.param), C.byref(r), C.byref(vmag))
        if not dt > 0:
            raise core.DesError(11, "dt <= 0")
        return dt

    def run(self, nsteps, dt, time=0.0, steps=0, dt_interval=10):
        for _ in range(nsteps):
            steps += 1
            time += dt
            self.step(dt, time, steps)
            if steps % dt_interval == 0:
                dt = self.compute_dt()
        return dt, time, steps


def local_mesh(coord, conn, bcflag, perm=None):
    """DesMesh dict of a rank's sub-mesh (box boundaries: flags decide what is a boundary
    facet, exactly as on the global mesh, so partition cuts create none).  conn is in the global
    relative order; perm (partition.interior_first) is the order the device stores the elements
    in -- the support and facet lists keep the GLOBAL order, which is the summation order."""
    sup = core.host_build_support(coord.shape[1], conn)
    if perm is None:
        return bm.boundary_topology(coord, conn, bcflag, support=sup)
    inv = np.empty(perm.size, dtype=np.int32)
    inv[perm] = np.arange(perm.size, dtype=np.int32)
    sup = (sup[0], inv[sup[1]], sup[2])
    return bm.boundary_topology(coord, np.ascontiguousarray(conn[:, perm]), bcflag, support=sup,
                                elem_key=np.asarray(perm))


def init_model(mesh, phys, state, comp_p, device):
    """What init() does after the mesh exists (dynearthsol.cxx:159-228), on one rank."""
    g = core.DesGpu(core.make_param(**phys), device)
    g.upload_mesh(mesh)
    g.upload_elemmarkers(np.full(mesh["nelem"], 4, dtype=np.int32))
    g.set_compensation_pressure(comp_p)
    g.upload_state(**state)
    return g


def build_box_rank(dims, rank, world, phys_overrides=None):
    """Rank-local structured box (generated directly, never the global mesh), its physics
    and initial fields.  dims = (xlength, ylength, zlength, resolution)."""
    xl, yl, zl, res = dims
    nx, ny, nz = bm.grid_dims(xl, yl, zl, res)
    boxes = partition.grid_boxes(nx, ny, nz, (xl, yl, zl), world)
    coord, conn, bcflag, key, owner = bm.regular_box_part(xl, yl, zl, nx, ny, nz, rank, boxes)
    plan = partition.halo_plan(rank, conn, owner, key)
    perm, n_interior = partition.interior_first(conn, owner, rank)
    mesh = local_mesh(coord, conn, bcflag, perm)
    conn = mesh["connectivity"]                  # device order from here on
    phys = dict(bm.TEST3D_PHYSICS)
    phys["zlength"] = zl
    if phys_overrides:
        phys.update(phys_overrides)
    state, comp_p = bm.initial_fields(coord, conn, rho0=phys["rho0"][0], bulk_modulus=phys["bulk_modulus"][0],
                                      gravity=phys["gravity"], xlength=xl, ylength=yl, zlength=zl,
                                      resolution=res, weakzone=bm.TEST3D_WEAKZONE)
    own_elems = int((owner[conn].min(axis=0) == rank).sum())     # each element counted on one rank
    return dict(mesh=mesh, phys=phys, state=state, comp_p=comp_p, plan=plan, owner=owner, key=key,
                nelem_local=conn.shape[1], nnode_local=coord.shape[1], nelem_owned=own_elems,
                n_interior=n_interior,
                cells=(nx - 1, ny - 1, nz - 1), nelem_global=(nx - 1) * (ny - 1) * (nz - 1) * 5,
                nnode_global=nx * ny * nz)


def bench_main(a, rank, world, local_rank, W, K):
    """bench.py under torchrun.  Default: the north star's multi-GPU configuration, the 200M-tet
    box (450x300x300 cells, BASELINE.json configs[2]) split over the N GPUs (strong scaling;
    the 1-GPU line is the 50M box, the largest single-GPU configuration).  --weak: the
    150x100x100 km box refined so that every GPU holds about as many tets as --size (50M):
    resolution / world^(1/3)."""
    import torch
    import torch.distributed as dist
    torch.cuda.set_device(local_rank)
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    from bench import SIZES, ClockSampler, alg_bytes_step, read_peaks      # bench.py is the main module's file
    strong = not bool(getattr(a, "weak", False))
    size = a.size or ("200M" if strong else "50M")
    xl, yl, zl, res = SIZES[size]
    if not strong:
        res = res / world ** (1.0 / 3.0)
    t0 = time.time()
    prob = build_box_rank((xl, yl, zl, res), rank, world)
    g = init_model(prob["mesh"], prob["phys"], prob["state"], prob["comp_p"], local_rank)
    m = DistModel(g, prob["plan"], local_rank, n_interior=prob["n_interior"])
    g.call("apply_vbcs", 0.0)
    g.call("compute_mass")
    dt0 = m.compute_dt()
    if a.verbose:
        print("# rank %d: %d local tets (%d owned), %d nodes, %d neighbours, send %d recv %d, setup %.1f s"
              % (rank, prob["nelem_local"], prob["nelem_owned"], prob["nnode_local"], len(prob["plan"].neighbours),
                 prob["plan"].send_nodes.size, prob["plan"].recv_nodes.size, time.time() - t0), file=sys.stderr)
    E = prob["nelem_global"]
    dt, t, s = m.run(W, dt0)
    torch.cuda.synchronize()
    dist.barrier()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    launches0 = g.launch_count()
    torch.cuda.synchronize()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    dt, t, s = m.run(K, dt, t, s)
    ev1.record()
    torch.cuda.synchronize()
    dist.barrier()
    ms = torch.tensor([ev0.elapsed_time(ev1)], dtype=torch.float64, device="cuda")
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms = float(ms.item())
    launches = g.launch_count() - launches0
    sampler.stop_flag = True
    value = E * K / (ms * 1e-3)

    # end to end: evolving state from pinned host memory each run, outputs back to the host
    from bench import EVOLVING, OUTPUT, pin
    e2e = None
    if not a.no_e2e:
        got = g.download_state(*EVOLVING)
        up, keep, down = {}, [], {}
        for k in EVOLVING:
            up[k], tk = pin(got[k]); keep.append(tk)
        for k in OUTPUT:
            down[k], tk = pin(np.empty_like(g.download_state(k)[k])); keep.append(tk)
        h2d = sum(v.nbytes for v in up.values())
        d2h = sum(v.nbytes for v in down.values())
        torch.cuda.synchronize(); dist.barrier()
        t0 = time.perf_counter()
        g.upload_state(**up)
        dt, t, s = m.run(K, dt, t, s)
        g.download_into(**down)
        torch.cuda.synchronize(); dist.barrier()
        te = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device="cuda")
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
        tot = torch.tensor([h2d, d2h], dtype=torch.float64, device="cuda")
        dist.all_reduce(tot)
        e2e = {"value": E * K / float(te.item()), "unit": "element-steps/s",
               "h2d_bytes_per_step": float(tot[0].item()) / K, "d2h_bytes_per_step": float(tot[1].item()) / K,
               "seconds": float(te.item()),
               "note": "every rank: upload evolving state (pinned) + K steps + download output fields"}
    stats = torch.tensor([prob["nelem_local"], prob["plan"].send_nodes.size], dtype=torch.float64, device="cuda")
    dist.all_reduce(stats)
    if rank == 0:
        sampler.join(timeout=2)
        peak, peak_src = read_peaks()
        r = prob["nnode_global"] / E
        alg = alg_bytes_step(r)
        line = {"metric": "3D tet element-steps/sec", "value": value, "unit": "element-steps/s", "n_gpus": world,
                "steps": K, "warmup": W, "ms_per_step": ms / K, "higher_is_better": True, "scaling": "strong" if strong else "weak",
                "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                "config": {"workload": "box %s %s%d (%dx%dx%d cells, %d tets, %d nodes): test-3d.cfg Mohr-Coulomb "
                                       "elasto-plastic + thermal + mixed stress + surface diffusion, compute_dt "
                                       "every 10 steps" % ((size, "split over " if strong else "x", world) + prob["cells"]
                                                           + (E, prob["nnode_global"])),
                           "element_order": "reference", "l2": "working set per GPU >> 126 MB L2",
                           "parallelism": "%d GPUs: RCB node-box partition, one-element halo, ghost-node exchange "
                                          "(NCCL send/recv) at the 3 nodal barriers + surface diffusion%s, allreduce(min/max) for dt"
                                          % (world, ", the two exchanges before E-B / E-C overlapped with those kernels' interior elements"
                                             if m.overlap else ""),
                           "redundant_element_fraction": float(stats[0].item()) / E - 1.0,
                           "ghost_nodes_sent_per_step_total": float(stats[1].item())},
                "clocks": sampler.summary(), "e2e": e2e, "gpu_launches": launches,
                "roofline": {"bound": "hbm", "kernel": "whole step (per-kernel breakdown: 1-GPU run)",
                             "achieved": value * alg / 1e9 / world, "peak": peak, "unit": "GB/s",
                             "frac": value * alg / 1e9 / world / peak, "traffic": None, "peak_source": peak_src,
                             "alg_bytes_per_elem_step": alg, "per": "GPU"},
                "cpu_baseline": None}
        print(json.dumps(line))
    g.close()
    dist.destroy_process_group()
