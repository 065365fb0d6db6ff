"""torchrun worker of tests/test_multigpu.py::test_ranks_nccl_bit_identical: 12 steps of a box on
WORLD_SIZE GPUs (one process per GPU, NCCL), every rank saving the values of the nodes it owns
and of the elements it reports (lowest owner among the element's nodes)."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import torch
import torch.distributed as dist
from dynearthsol_b200 import multigpu as mg

BOX = tuple(float(x) for x in sys.argv[2:6]) if len(sys.argv) >= 6 else (24e3, 10e3, 8e3, 2e3)
rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr)
dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
prob = mg.build_box_rank(BOX, rank, world)
g = mg.init_model(prob["mesh"], prob["phys"], prob["state"], prob["comp_p"], lr)
m = mg.DistModel(g, prob["plan"], lr, n_interior=prob["n_interior"])
g.call("apply_vbcs", 0.0)
g.call("compute_mass")
dt = m.compute_dt()
dt, t, s = m.run(12, dt)
N, E = prob["nnode_local"], prob["nelem_local"]
out = g.download_state("coord", "vel", "temperature", "force", "mass", "stress", "plstrain", "volume")
own = prob["owner"] == rank
conn = prob["mesh"]["connectivity"].reshape(4, E)
eown = prob["owner"][conn].min(axis=0) == rank
ekey = np.sort(prob["key"][conn], axis=0)           # an element is named by its sorted global node keys
save = {"key": prob["key"][own], "ekey": ekey[:, eown], "dt": np.array([dt])}
for k in ("coord", "vel", "force"):
    save[k] = out[k].reshape(3, N)[:, own]
for k in ("temperature", "mass"):
    save[k] = out[k][own]
save["stress"] = out["stress"].reshape(6, E)[:, eown]
for k in ("plstrain", "volume"):
    save[k] = out[k][eown]
np.savez(os.path.join(sys.argv[1], "rank%d.npz" % rank), **save)
g.close()
dist.barrier()
dist.destroy_process_group()
