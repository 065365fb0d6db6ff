#!/usr/bin/env python
"""bench.py -- 3D tet element-steps/s of the explicit time-step core on B200.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--size 50M|...]

Workload (BASELINE.json configs[2], the configuration the metric is quoted on): the
150 x 100 x 100 km box on the reference's regular hex->5-tet generator at resolution
526.3158 m (285x190x190 cells, 51 442 500 tets, 10 433 566 nodes), Mohr-Coulomb
elasto-plastic + thermal diffusion + mixed stress + surface diffusion, physics of
benchmarks-cores/test-3d.cfg; synthetic mesh and initial fields (dynearthsol_b200/boxmesh.py).
A "step" is one explicit time step (dynearthsol.cxx:768-894), compute_dt every 10 steps.

value      device-resident: K steps timed with CUDA events on the launching stream.
e2e        through the C-ABI with host buffers: upload of the evolving state from pinned host
           memory + K steps + download of the output fields, all inside the timed region.
roofline   dominant kernel: algorithmic bytes per launch / its average
           launch duration (CUDA events, separate instrumented pass) vs the measured HBM peak.
cpu_baseline / --impl reference: the UNMODIFIED reference (oracle/_ref/libdesref.so) on the
           host cores, same physics, on a bounded sample (a smaller box it can initialise).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

SIZES = {  # name: (xlength, ylength, zlength, resolution)
    "50M": (150e3, 100e3, 100e3, 526.3158),
    "200M": (150e3, 100e3, 100e3, 333.3333),
    "6M": (150e3, 100e3, 100e3, 1063.83),
    "1M": (150e3, 100e3, 100e3, 1923.1),
    "100k": (150e3, 100e3, 100e3, 4166.7),
}

# Algorithmic (compulsory) bytes per element-step, SURVEY.md section 8(d), apportioned to the
# six kernels of the schedule; r = nnode / nelem.
def alg_bytes(r):
    return {
        "elem_a": 16 + 4 + 8 + 8 + 64 + 56 * r,                 # conn, markers, radiogenic, volume W, record W; coord+vel+T
        "node_a": 64 + 32 + 56 * r,                             # record R, CSR; T rw, mass/tmass/volume_n/ntmp W, bcflag, idx
        "elem_b": 16 + 4 + 96 + 96 + 48 + 16 + 8 + 8 + 8 + 8 + 8 * r,  # conn, markers, stress rw, strain rw, strain_rate W, plstrain rw, dpl W, volume R, dpressure W, dp*vol W; ntmp
        # (SURVEY's count; the schedule also moves, compulsorily under its nodal barriers: the NMD
        #  scalar handed from E-C to the next E-B, 8 W + 8 R, and E-C's re-read of stress, conn,
        #  volume, markers -- DESIGN.md section 4: 923 B per element-step instead of 803)
        "node_b": 8 + 32 + 20 * r,                              # dp*vol R, CSR; volume_n, ntmp W, idx
        "elem_c": 8 + 96 + 8 * r,                               # dpressure R, force record W; ntmp
        "node_c": 96 + 32 + 136 * r,                            # force record R, CSR; coord rw, vel rw, force W, mass, bcflag, idx
    }


def alg_bytes_step(r):
    """B_alg (EP + thermal + NMD, nmat=1) = 756 + 232 r."""
    return 756.0 + 232.0 * r


KCLASS_KEY = {"elem_a": "elem_a(", "node_a": "node_a(", "elem_b": "elem_b(", "node_b": "node_b(",
              "elem_c": "elem_c(", "node_c": "node_c("}


def read_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler(threading.Thread):
    """nvidia-smi clocks and throttle reasons during the timed region."""

    def __init__(self, gpu_index):
        super().__init__(daemon=True)
        self.idx = gpu_index
        self.rows = []
        self.stop_flag = False

    def run(self):
        q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
             "clocks_event_reasons.sw_power_cap")
        try:
            p = subprocess.Popen(["nvidia-smi", "-i", str(self.idx), "--query-gpu=" + q,
                                  "--format=csv,noheader,nounits", "-lms", "100"],
                                 stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except OSError:
            return
        while not self.stop_flag:
            line = p.stdout.readline()
            if not line:
                break
            self.rows.append([x.strip() for x in line.split(",")])
        p.terminate()

    def summary(self):
        sm = [float(r[0]) for r in self.rows if r and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) > 1 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names)
                   if any(len(r) > 2 + i and r[2 + i].lower().startswith("active") for r in self.rows)]
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


def build_problem(size, order="reference", verbose=False):
    """Mesh, topology, parameters and initial state for one box (host, numpy)."""
    from dynearthsol_b200 import boxmesh as bm, core
    xl, yl, zl, res = SIZES[size]
    nx, ny, nz = bm.grid_dims(xl, yl, zl, res)
    t0 = time.time()
    coord, conn, bcflag = bm.regular_box(xl, yl, zl, nx, ny, nz, order=order)
    sup = core.host_build_support(coord.shape[1], conn)
    mesh = bm.boundary_topology(coord, conn, bcflag, support=sup)
    mesh["coord"] = coord
    phys = dict(bm.TEST3D_PHYSICS)
    phys["zlength"] = zl
    state, comp_p = bm.initial_fields(coord, conn, rho0=phys["rho0"][0], bulk_modulus=phys["bulk_modulus"][0],
                                      gravity=phys["gravity"], xlength=xl, ylength=yl, zlength=zl,
                                      resolution=res, weakzone=bm.TEST3D_WEAKZONE)
    if verbose:
        print("# mesh %dx%dx%d cells, %d tets, %d nodes, host setup %.1f s"
              % (nx - 1, ny - 1, nz - 1, conn.shape[1], coord.shape[1], time.time() - t0), file=sys.stderr)
    return dict(mesh=mesh, phys=phys, state=state, comp_p=comp_p, cells=(nx - 1, ny - 1, nz - 1),
                nelem=conn.shape[1], nnode=coord.shape[1])


def make_model(prob, device):
    """Device model initialised the way init() does (dynearthsol.cxx:159-228): upload, apply_vbcs,
    compute_dt, compute_mass."""
    from dynearthsol_b200 import core
    g = core.DesGpu(core.make_param(**prob["phys"]), device)
    g.upload_mesh(prob["mesh"])
    g.upload_elemmarkers(np.full(prob["nelem"], 4, dtype=np.int32))
    g.set_compensation_pressure(prob["comp_p"])
    g.upload_state(**prob["state"])
    g.call("apply_vbcs", 0.0)
    g.call("compute_mass")
    dt = g.compute_dt()[0]
    return g, dt


def pin(a):
    """A pinned-host copy of a numpy array (torch is only the pinned allocator here)."""
    import torch
    t = torch.from_numpy(np.ascontiguousarray(a)).pin_memory()
    return t.numpy(), t


EVOLVING = ["coord", "vel", "temperature", "stress", "strain", "plstrain",
            "volume", "volume_old"]      # SURVEY.md section 8b "evolving state" (strain_rate is derived: E-B writes it)
OUTPUT = ["coord", "vel", "temperature", "stress", "strain", "strain_rate", "plstrain",
          "delta_plstrain", "force", "volume"]                        # what Output::_write reads


def run_steps(g, n, dt, t, s):
    return g.run(n, dt, t, s)


def read_traffic():
    """DRAM bytes per element and launch of each kernel, from the committed ncu --set full
    capture at 51.4M tets (profiles/r02_traffic_50M.json); None if absent."""
    p = os.path.join(ROOT, "profiles", "r02_traffic_50M.json")
    if os.path.exists(p):
        return json.load(open(p))
    return None


def host_ram_gb():
    try:
        import psutil
        return psutil.virtual_memory().total / 1e9
    except Exception:
        return 0.0


def reference_cpu(steps, warmup, sample="1M", target_seconds=None):
    """The unmodified reference on the host cores, same physics, same box generator.
    Returns a dict (rate, cores, nelem, steps, seconds, init_seconds, desc).  Runs in a
    subprocess (the reference keeps process-wide state).  The hot functions that are timed are
    the reference's own object code; only init()'s O(nseg*nelem) boundary-facet search is
    replaced by the support-based one of oracle/ref_harness.cxx (same arrays, tested), without
    which the reference cannot initialise a box of more than ~1 M tets."""
    lib = os.path.join(ROOT, "oracle", "_ref", "libdesref.so")
    if not os.path.exists(lib):
        return {"rate": None, "cores": 0, "desc": "oracle/_ref/libdesref.so absent"}
    xl, yl, zl, res = SIZES[sample]
    cfg = CFG_TEMPLATE % dict(xl=xl, yl=yl, zl=zl, res=res)
    import tempfile
    wd = tempfile.mkdtemp(prefix="desref_")
    open(os.path.join(wd, "sample.cfg"), "w").write(cfg)
    code = (
        "import sys, json, time; sys.path.insert(0, %r)\n"
        "from refharness import Ref\n"
        "t0 = time.time()\n"
        "r = Ref(%r, workdir=%r, fast_init=True)\n"
        "ti = time.time() - t0\n"
        "r.step(%d)\n"
        "n = %d\n"
        "t = r.step(n)\n"
        "target = %r\n"
        "if target and t < target / 2:\n"
        "    n = max(n, min(20000, int(target / (t / n))))   # bounded sample: ~target seconds of CPU work\n"
        "    t = r.step(n)\n"
        "print('RESULT', json.dumps({'t': t, 'n': n, 'ti': ti, 'nelem': r.int('nelem'), 'nnode': r.int('nnode')}))\n"
    ) % (os.path.join(ROOT, "tests"), os.path.join(wd, "sample.cfg"), wd, warmup, steps, target_seconds)
    cores = os.cpu_count() or 1
    env = dict(os.environ, OMP_NUM_THREADS=str(cores))
    out = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, env=env)
    for line in out.stdout.splitlines():
        if line.startswith("RESULT"):
            r = json.loads(line[7:])
            rate = r["nelem"] * r["n"] / r["t"]
            return {"rate": rate, "cores": cores, "nelem": r["nelem"], "nnode": r["nnode"], "steps": r["n"],
                    "seconds": r["t"], "init_seconds": r["ti"],
                    "desc": ("unmodified reference (g++ -O2 -fopenmp), %d-tet box (%s), same physics, "
                             "%d steps in %.1f s after %d warm-up, %d OpenMP threads; init %.0f s"
                             % (r["nelem"], sample, r["n"], r["t"], warmup, cores, r["ti"]))}
    return {"rate": None, "cores": cores, "desc": "reference run failed: " + out.stderr[-300:]}


def workload_name(size, cells, E, N):
    return ("box %s (%dx%dx%d cells, %d tets, %d nodes): test-3d.cfg Mohr-Coulomb "
            "elasto-plastic + thermal + mixed stress + surface diffusion, "
            "compute_dt every 10 steps" % ((size,) + tuple(cells) + (E, N)))


CFG_TEMPLATE = """[sim]
modelname = sample
max_steps = 10
output_step_interval = 100000000
is_outputting_averaged_fields = no
has_initial_checkpoint = no
[mesh]
meshing_option = 1
meshing_elem_shape = 1
xlength = %(xl)r
ylength = %(yl)r
zlength = %(zl)r
resolution = %(res)r
quality_check_step_interval = 100000
remeshing_option = 0
[control]
surface_process_option = 1
surface_diffusivity = 1e-6
dt_fraction = 1.0
inertial_scaling = 1e4
[bc]
vbc_x0 = 1
vbc_x1 = 1
vbc_val_x0 = -1e-9
vbc_val_x1 = 1e-9
vbc_y0 = 1
vbc_y1 = 1
vbc_val_y0 = 0
vbc_val_y1 = 0
has_water_loading = no
surface_temperature = 273
mantle_temperature = 273
[ic]
weakzone_option = 1
weakzone_azimuth = 15
weakzone_inclination = -60
weakzone_halfwidth = 1.2
weakzone_depth_min = 0.5
weakzone_depth_max = 1.0
weakzone_xcenter = 0.5
weakzone_ycenter = 0.5
weakzone_zcenter = 0
weakzone_plstrain = 0.5
[mat]
rheology_type = elasto-plastic
rho0 = [2700]
alpha = [0]
bulk_modulus = [50e9]
shear_modulus = [30e9]
pls0 = [0]
pls1 = [0.5]
cohesion0 = [4.4e7]
cohesion1 = [4e6]
friction_angle0 = [30]
friction_angle1 = [30]
min_viscosity = 1e24
"""


def driver_bench(a, K, W, metric, unit):
    """dynearthsol3d.b200 (the reference's object code + the bridge) on a generated box .cfg."""
    import re
    import tempfile
    size = a.size or "6M"
    xl, yl, zl, res = SIZES[size]
    exe = os.path.join(ROOT, "bridge", "_build", "dynearthsol3d.b200")
    if not os.path.exists(exe):
        print(json.dumps({"impl": "driver", "unavailable": "bridge/_build/dynearthsol3d.b200 not built"}))
        return
    wd = tempfile.mkdtemp(prefix="desdrv_")
    # the reference's "Compute" bucket also holds the final download and end() (freeing the host
    # arrays): a few hundred steps make that a small share
    nsteps = max(K + W, 300)
    cfg = (CFG_TEMPLATE % dict(xl=xl, yl=yl, zl=zl, res=res)).replace("max_steps = 10", "max_steps = %d" % nsteps)
    open(os.path.join(wd, "drv.cfg"), "w").write(cfg)
    env = dict(os.environ, OMP_NUM_THREADS=str(os.cpu_count() or 1), DES_BRIDGE_NO_INITIAL_FRAME="1")
    t0 = time.time()
    r = subprocess.run([exe, "drv.cfg"], cwd=wd, env=env, capture_output=True, text=True)
    wall = time.time() - t0
    m = re.search(r"Compute\s*:\s*\S+.*?/\s*(\d+)\s*=\s*([0-9.eE+-]+)\s*s/step", r.stdout)
    mm = re.search(r"nodes=(\d+), elements=(\d+)", r.stdout)
    im = re.search(r"Initiate:\s*(\d+):(\d+):([0-9.]+)", r.stdout)
    if r.returncode != 0 or not m or not mm:
        print(json.dumps({"impl": "driver", "unavailable": "driver run failed: " + (r.stderr or r.stdout)[-300:]}))
        return
    N, E = int(mm.group(1)), int(mm.group(2))
    s_step = float(m.group(2))
    init_s = (int(im.group(1)) * 3600 + int(im.group(2)) * 60 + float(im.group(3))) if im else None
    print(json.dumps({"impl": "driver", "metric": metric, "value": E / s_step, "unit": unit, "n_gpus": 1,
                      "steps": int(m.group(1)), "warmup": 0, "ms_per_step": s_step * 1e3, "higher_is_better": True,
                      "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                      "config": {"workload": "box %s (%d tets, %d nodes) from a .cfg through dynearthsol3d.b200: the reference's "
                                             "init() (markers, initial fields) with create_boundary_facets / create_support "
                                             "swapped for the library's builders, step loop on the GPU, compute_dt every 10 steps; "
                                             "Compute = run time - init - output, over ALL steps (no warm-up excluded)" % (size, E, N),
                                 "init_seconds": init_s, "wall_seconds": wall}}))
    import shutil
    shutil.rmtree(wd, ignore_errors=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--size", default=None, help="box size key: " + ", ".join(SIZES))
    ap.add_argument("--order", default="reference", choices=["reference", "morton"])
    ap.add_argument("--cpu-sample", default=None,
                    help="box the CPU reference runs on (default: --impl reference: the named workload when "
                         "the host has the RAM for it, else 6M; cpu_baseline leg: 1M)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--driver", action="store_true",
                    help="measure the path through the reference's own driver instead of the ctypes mirror: "
                         "bridge/_build/dynearthsol3d.b200 reads a .cfg (meshing_option = 1 box of --size, default 6M), "
                         "init() runs on the host with the library's topology builders, the step loop on the GPU; "
                         "the value is the reference's own \"Compute ... s/step\" line")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--verbose", action="store_true")
    ap.add_argument("--strong", action="store_true",
                    help="N > 1: keep the --size mesh as it is and split it over the ranks (strong scaling). "
                         "This is the default for N > 1: the north star's 200M-tet box split over the N GPUs")
    ap.add_argument("--weak", action="store_true",
                    help="N > 1: weak scaling instead -- the box refined so that every GPU holds the --size mesh "
                         "(default 50M per GPU)")
    a = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    W = max(a.warmup, 3)
    K = a.steps
    metric = "3D tet element-steps/sec"
    unit = "element-steps/s"

    if a.impl == "reference":
        if rank != 0:
            return
        size = a.size or "50M"
        # the named workload itself when the host can hold it (the reference needs ~65 GB for the
        # 51.4M-tet box: 4 markers per element, vector<vector<int>> marker counts), else the
        # largest box that fits; --cpu-sample overrides
        sample = a.cpu_sample
        if sample is None:
            sample = size if host_ram_gb() >= 160 else ("6M" if host_ram_gb() >= 24 else "1M")
        res_ = reference_cpu(K, min(W, 3), sample)
        rate, cores, desc = res_["rate"], res_["cores"], res_["desc"]
        xl, yl, zl, res = SIZES[size]
        from dynearthsol_b200 import boxmesh as bm
        nx, ny, nz = bm.grid_dims(xl, yl, zl, res)
        E_named, N_named = (nx - 1) * (ny - 1) * (nz - 1) * 5, nx * ny * nz
        same = rate is not None and res_["nelem"] == E_named
        # ms_per_step is MEASURED: wall time of the K timed steps of the box that was run
        ms_step = (res_["seconds"] / res_["steps"] * 1e3) if rate else None
        line = {"impl": "reference", "metric": metric, "value": rate, "unit": unit, "n_gpus": a.gpus,
                "steps": res_.get("steps", K), "warmup": min(W, 3), "ms_per_step": ms_step, "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                "config": {"workload": workload_name(size, (nx - 1, ny - 1, nz - 1), E_named, N_named),
                           "sample": desc, "same_config": same,
                           "sample_nelem": res_.get("nelem"), "host_ram_gb": round(host_ram_gb(), 1),
                           "element_order": "reference", "parallelism": "%d OpenMP threads" % cores},
                "cpu_baseline": {"value": rate, "unit": unit, "cores": cores, "kind": "reference", "sample": desc},
                "e2e": {"value": rate, "unit": unit, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
        if rate is None:
            line["unavailable"] = desc
        print(json.dumps(line))
        return

    if world > 1:
        from dynearthsol_b200 import multigpu
        return multigpu.bench_main(a, rank, world, local_rank, W, K)

    if a.driver:
        return driver_bench(a, K, W, metric, unit)

    size = a.size or "50M"
    prob = build_problem(size, a.order, a.verbose)
    E, N = prob["nelem"], prob["nnode"]
    g, dt0 = make_model(prob, local_rank)

    # ---- device-resident timing -------------------------------------------------------
    dt, t, s = g.run(W, dt0)
    g.sync()
    sampler = ClockSampler(local_rank)
    sampler.start()
    launches0 = g.launch_count()
    g.event_record(0)
    dt, t, s = g.run(K, dt, t, s)
    g.event_record(1)
    ms = g.event_elapsed_ms(0, 1)
    launches = g.launch_count() - launches0
    sampler.stop_flag = True
    value = E * K / (ms * 1e-3)

    # ---- per-kernel pass (CUDA events around every launch) -----------------------------
    g.enable_kernel_timing(True)
    nk = min(K, 20)
    dt, t, s = g.run(nk, dt, t, s)
    kt = g.kernel_times()
    g.enable_kernel_timing(False)
    r = N / E
    ab = alg_bytes(r)
    peak, peak_src = read_peaks()
    per_kernel = {}
    tot_ms = sum(m for _, m, _ in kt)
    for key, pat in KCLASS_KEY.items():
        for name, m, n in kt:
            if name.startswith(pat):
                avg = m / n
                per_kernel[key] = {"avg_ms": avg, "launches_per_step": n / nk, "share": m / tot_ms,
                                   "alg_bytes_per_elem": ab[key],
                                   "achieved_gbs": ab[key] * E / (avg * 1e-3) / 1e9}
    dom = max(per_kernel, key=lambda k: per_kernel[k]["share"])
    step_alg = alg_bytes_step(r)
    tr = read_traffic()
    traffic = None
    if tr:
        for key, v in tr["dram_bytes_per_elem"].items():
            if key in per_kernel:
                per_kernel[key]["dram_bytes_per_elem_ncu"] = v
        if dom in tr["dram_bytes_per_elem"]:
            traffic = tr["dram_bytes_per_elem"][dom] * E      # bytes per launch (the capture is at 51.4M tets)
    roofline = {"bound": "hbm", "kernel": dom, "achieved": per_kernel[dom]["achieved_gbs"], "peak": peak,
                "unit": "GB/s", "frac": per_kernel[dom]["achieved_gbs"] / peak, "traffic": traffic,
                "traffic_source": tr["source"] if tr else None,
                "peak_source": peak_src,
                "step": {"alg_bytes_per_elem_step": step_alg, "achieved_gbs": value * step_alg / 1e9,
                         "frac": value * step_alg / 1e9 / peak,
                         # what a three-phase schedule has to move and what ncu measured
                         "compulsory_bytes_per_elem_step_under_the_nodal_barriers": step_alg + 120.0,
                         "frac_on_compulsory_bytes": value * (step_alg + 120.0) / 1e9 / peak,
                         "dram_bytes_per_elem_step_ncu": tr["dram_bytes_per_elem_step"] if tr and "dram_bytes_per_elem_step" in tr else None},
                "kernels": per_kernel}

    # ---- end to end through the C-ABI with pinned host buffers -------------------------
    e2e = None
    if not a.no_e2e:
        got = g.download_state(*EVOLVING)
        up, keep = {}, []
        for k in EVOLVING:
            up[k], tk = pin(got[k]); keep.append(tk)
        down = {}
        for k in OUTPUT:
            down[k], tk = pin(np.empty_like(g.download_state(k)[k])); keep.append(tk)
        h2d = sum(v.nbytes for v in up.values())
        d2h = sum(v.nbytes for v in down.values())
        g.sync()
        t0 = time.perf_counter()
        g.upload_state(**up)
        dt, t, s = g.run(K, dt, t, s)
        g.download_into(**down)
        g.sync()
        te = time.perf_counter() - t0
        e2e = {"value": E * K / te, "unit": unit, "h2d_bytes_per_step": h2d / K, "d2h_bytes_per_step": d2h / K,
               "seconds": te, "note": "upload evolving state (pinned) + K steps + download output fields"}

    # ---- CPU baseline: the unmodified reference on the host cores -----------------------
    cpu = None
    if not a.no_cpu:
        res_ = reference_cpu(10, 2, a.cpu_sample or "1M", target_seconds=12.0)
        cpu = {"value": res_["rate"], "unit": unit, "cores": res_["cores"], "kind": "reference", "sample": res_["desc"]}

    sampler.join(timeout=2)
    line = {"metric": metric, "value": value, "unit": unit, "n_gpus": 1, "steps": K, "warmup": W,
            "ms_per_step": ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic",
            "config": {"workload": workload_name(size, prob["cells"], E, N),
                       "element_order": a.order, "l2": "working set (%.1f GB) >> 126 MB L2" % (E * 1160 / 1e9),
                       "parallelism": "1 GPU"},
            "clocks": sampler.summary(), "e2e": e2e, "gpu_launches": launches,
            "roofline": roofline, "cpu_baseline": cpu}
    print(json.dumps(line))
    g.close()


if __name__ == "__main__":
    main()
