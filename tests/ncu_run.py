"""A bare stepping run for profilers: builds the box (bench.py's generator), runs N steps.
    python tests/ncu_run.py 50M 8"""
import os
import sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench

size, steps = sys.argv[1], int(sys.argv[2])
prob = bench.build_problem(size)
g, dt = bench.make_model(prob, 0)
g.run(steps, dt)
g.sync()
print("ran", steps, "steps of", prob["nelem"], "tets")
g.close()
