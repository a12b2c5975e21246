#!/usr/bin/env python3
"""Wall clock of picb200_create / destroy for the cfg2 configuration (what picb200_run pays besides the run)."""
import importlib, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
eng = importlib.import_module("particle-simulation_b200")
kw = dict(mode=0, init_n=1000000, poisson_steps=100, block_size=256, capacity=100000000, sleep_time_ns=100, poisson_timestep=10)
for it in range(4):
    t0 = time.perf_counter(); sim = eng.Sim(**kw); t1 = time.perf_counter()
    sim.init_uniform(1000000); sim.run_steps(100); s = sim.summary(); t2 = time.perf_counter()
    sim.close(); t3 = time.perf_counter()
    print("create %.2f ms  init+run+summary %.2f ms (gpu %.2f)  destroy %.2f ms" % (1e3 * (t1 - t0), 1e3 * (t2 - t1), s["gpu_ms"], 1e3 * (t3 - t2)))
