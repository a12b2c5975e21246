#!/usr/bin/env python3
"""Short driver for ncu: one cfg2-shaped run (mode/size/steps from argv) through the C ABI.
usage: prof_run.py [mode] [init_n] [poisson_steps] [poisson_timestep] [capacity] [warm_steps]
`warm_steps` Poisson steps are run first so that collisions are switched on in the profiled part."""
import importlib
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
eng = importlib.import_module("particle-simulation_b200")
a = [int(v) for v in sys.argv[1:]] + [None] * 6
mode, n, P, pt, cap, warm = a[0] or 0, a[1] or 1000000, a[2] or 20, a[3] or 10, a[4] or 100000000, a[5] or 0
with eng.Sim(mode=mode, init_n=n, poisson_steps=P + warm, capacity=cap, poisson_timestep=pt) as sim:
    sim.init_uniform(n)
    if warm:
        sim.run_steps(warm)
    before = sim.summary()["pushes"]
    sim.set_profiling(2)
    sim.run_steps(P)
    print("pushes_in_profiled_launch", sim.summary()["pushes"] - before)
    print(sim.summary())
    print(sim.profile())
