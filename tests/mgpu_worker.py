"""torchrun worker of tests/test_multigpu.py::test_two_ranks_nccl_bit_identical: 12 steps of the
small box on WORLD_SIZE GPUs, every rank saving its local result."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import torch
import torch.distributed as dist
from dynearthsol_b200 import multigpu as mg

BOX = (24e3, 10e3, 8e3, 2e3)
rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr)
dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
prob = mg.build_box_rank(BOX, rank, world)
g = mg.init_model(prob["mesh"], prob["phys"], prob["state"], prob["comp_p"], lr)
m = mg.DistModel(g, prob["plan"], lr, n_interior=prob["n_interior"])
g.call("apply_vbcs", 0.0)
g.call("compute_mass")
dt = m.compute_dt()
m.run(12, dt)
out = g.download_state("coord", "vel", "temperature", "force", "mass")
np.savez(os.path.join(sys.argv[1], "rank%d.npz" % rank), key=prob["key"], owner=prob["owner"], **out)
g.close()
dist.barrier()
dist.destroy_process_group()
