"""Cost of Output::average_fields on the device (the reference's default: averaged output on):
ms/step of the plain step loop, of step + des_gpu_average_fields riding on the next E-B
(default), and of the same with a flush after every step (what a per-step download forced)."""
import json
import os
import sys
import time

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import bench

size = sys.argv[1] if len(sys.argv) > 1 else "6M"
K = int(sys.argv[2]) if len(sys.argv) > 2 else 40
prob = bench.build_problem(size)
g, dt = bench.make_model(prob, 0)
out = {"size": size, "nelem": prob["nelem"], "steps": K}


def loop(mode):
    t, s = 0.0, 0
    for w in range(2):      # warm-up, then timed
        g.sync()
        g.event_record(0)
        for i in range(K):
            t += dt; s += 1
            g.step(dt, t, s)
            if mode != "plain":
                g.average_fields(first=(i == 0))
            if mode == "flush":
                g.call("compute_volume")        # any twin call completes the deferred end-of-step work
        g.event_record(1)
        ms = g.event_elapsed_ms(0, 1) / K
    return ms


for mode in ("plain", "avg", "flush"):
    out[mode + "_ms_per_step"] = loop(mode)
print(json.dumps(out))
