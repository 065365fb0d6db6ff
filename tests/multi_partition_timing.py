"""What the in-library partition costs on a large mesh: the 6.2M-tet box (bench.py's generator)
on a group handle of P parts (all on device 0 when the box has one GPU), partition and upload
seconds as des_gpu_partition_info reports them, 20 steps against the single-device handle.

    python tests/multi_partition_timing.py out.json [size] [parts]
"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
from dynearthsol_b200 import core


def initial_state(prob):
    """What init() leaves (dynearthsol.cxx:159-228): boundary velocities applied, nodal mass
    computed -- done once on a single-device handle (the twins), then uploaded like the
    reference's own host arrays."""
    g, dt = bench.make_model(prob, 0)
    st = dict(prob["state"])
    st.update(g.download_state("vel", "mass", "tmass", "volume_n", "volume"))
    g.close()
    return st


def run(prob, devices, nsteps):
    g = core.DesGpu(core.make_param(**prob["phys"]), 0, devices=devices)
    t0 = time.time()
    g.upload_mesh(prob["mesh"])
    t_mesh = time.time() - t0
    g.upload_elemmarkers(np.full(prob["nelem"], 4, dtype=np.int32))
    g.set_compensation_pressure(prob["comp_p"])
    t0 = time.time()
    g.upload_state(**prob["state"])
    t_state = time.time() - t0
    dt = g.compute_dt()[0]
    g.run(3, dt)
    g.sync()
    t0 = time.time()
    dt, t, s = g.run(nsteps, dt, 3 * dt, 3)
    g.sync()
    ms = (time.time() - t0) / nsteps * 1e3
    t0 = time.time()
    out = g.download_state("vel", "stress", "plstrain")
    t_down = time.time() - t0
    info = g.partition_info()
    g.close()
    return out, dict(upload_mesh_s=t_mesh, upload_state_s=t_state, download_3_fields_s=t_down, ms_per_step_wall=ms, **info)


def main():
    import torch
    out = sys.argv[1]
    size = sys.argv[2] if len(sys.argv) > 2 else "6M"
    parts = int(sys.argv[3]) if len(sys.argv) > 3 else 4
    prob = bench.build_problem(size)
    prob["state"] = initial_state(prob)
    ndev = max(1, torch.cuda.device_count())
    a, ia = run(prob, None, 20)
    b, ib = run(prob, [i % ndev for i in range(parts)], 20)
    same = all(np.array_equal(a[k], b[k]) for k in a)
    res = {"mesh": "box %s: %d tets, %d nodes" % (size, prob["nelem"], prob["nnode"]), "devices_in_box": ndev,
           "single": ia, "group": ib, "bit_identical_after_23_steps": bool(same)}
    json.dump(res, open(out, "w"), indent=1)
    print(json.dumps(res, indent=1))
    assert same


if __name__ == "__main__":
    main()
